import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(
    os.path.dirname(os.path.abspath(__file__)))))))
if _ROOT not in sys.path:
    sys.path.append(_ROOT)
from commefficient_b200.models.fixup import FixupBottleneck, FixupResNet  # noqa: E402,F401
