import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box via gpurun)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(autouse=True)
def _fresh_runtime():
    """Each test gets its own single-rank runtime."""
    yield
    from commefficient_b200.parallel import runtime
    rt = runtime.get_runtime()
    if rt is not None:
        rt.shutdown()


@pytest.fixture(autouse=True)
def _deterministic_inputs(request):
    """Every test draws its random inputs from a seed derived from its own name, so a rare unlucky draw
    (e.g. a heavy hitter that lands next to zero, a near-tie at a top-k threshold) cannot fail a run that
    stops at the first failure, and a failure reproduces."""
    import zlib

    import numpy as np
    import torch
    seed = zlib.crc32(request.node.nodeid.encode()) & 0x7FFFFFFF
    torch.manual_seed(seed)             # seeds the CUDA generators too
    np.random.seed(seed % (2 ** 32))
    yield


@pytest.fixture(autouse=True)
def _default_nn_state():
    """The tensor-core switch and the precision are process-wide: start every test from the same state
    (native layers off, bf16 operands) so that no test's result depends on which test ran before it."""
    import torch
    from commefficient_b200.ops import nn as nnops
    nnops.set_native(False)
    nnops._precision = "bf16"
    # (set_precision("tf32") turns TF32 on for the library matmuls as well: a test's fp32 oracle must not inherit that)
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = True
    try:
        from commefficient_b200.ops import tc
        tc._PRECISION = "bf16"
    except Exception:
        pass
    yield
