"""Compatibility injections that let the UNMODIFIED reference import on this image
(torch 2.11 / torchvision 0.26 / numpy 2.3).  Nothing here is on the reference's compute
path except the two transport notes below.  Imported by every shim module so that the
reference's spawned worker processes get the same environment as its parameter server.

  * ``torchvision.models.utils.load_state_dict_from_url`` (removed upstream; imported by the
    reference's models/resnets.py:3) -> re-exported from torch.hub;
  * ``np.float`` (removed in numpy 1.24; used by the reference's TableLogger) -> float;
  * CEB_REF_ONE_GPU=1 (set by the harness only for the 1-GPU arm): NCCL refuses two ranks
    on one device ("Duplicate GPU detected"), and the reference needs a PS rank plus a
    worker rank; the process group is created with gloo instead and ``reduce`` on CUDA
    tensors (unsupported by gloo) is issued as ``all_reduce``.  The reference's code is
    unchanged; only the transport underneath its one collective differs for N=1.
"""
import os
import sys
import types

_done = False


def install():
    global _done
    if _done:
        return
    _done = True
    trace_after = int(os.environ.get("CEB_REF_TRACE_AFTER", "0") or 0)
    if trace_after > 0:
        import faulthandler                      # debugging aid: dump stacks if a process hangs
        faulthandler.dump_traceback_later(trace_after, exit=False)
    import numpy as np
    if not hasattr(np, "float"):
        np.float = float
    try:
        import torchvision.models as tvm
        if "torchvision.models.utils" not in sys.modules:
            mod = types.ModuleType("torchvision.models.utils")
            from torch.hub import load_state_dict_from_url
            mod.load_state_dict_from_url = load_state_dict_from_url
            sys.modules["torchvision.models.utils"] = mod
            tvm.utils = mod
    except Exception:
        pass
    if os.environ.get("CEB_REF_ONE_GPU") == "1":
        import torch.distributed as dist
        if not getattr(dist, "_ceb_patched", False):
            real_init = dist.init_process_group
            real_all_reduce = dist.all_reduce

            def init_process_group(backend=None, *a, **kw):
                if backend == "nccl":
                    backend = "gloo"
                return real_init(backend, *a, **kw)

            def reduce(tensor, dst, *a, **kw):
                return real_all_reduce(tensor)

            dist.init_process_group = init_process_group
            dist.reduce = reduce
            dist._ceb_patched = True


install()
