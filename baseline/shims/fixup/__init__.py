"""Import stub for the external ``fixup`` package (hongyi-zhang/Fixup), which the reference's
models/fixup_resnet9.py:6 and models/fixup_resnet.py:4 import at package-import time.  The
baseline benchmark runs ResNet9, so these classes are never executed there."""
import _ref_compat  # noqa: F401
