"""ctypes binding to the in-tree native library ``lib/libceb200.so``.

The library is plain C ABI (``include/commefficient_b200.h``) compiled by
``build_native.sh`` with ``nvcc -gencode arch=compute_100a,code=sm_100a``; it has no
dependency on libtorch, so it builds in seconds and loads anywhere a CUDA runtime
exists.  Tensors are passed as raw device pointers together with the current
PyTorch CUDA stream.

If a CUDA device is present but the library is missing we fail loudly instead of
silently falling back to PyTorch ops -- the fallback exists only for CPU tensors.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG = os.path.dirname(_HERE)
_ROOT = os.path.dirname(_PKG)
LIB_PATH = os.path.join(_PKG, "lib", "libceb200.so")

_lib: Optional[ctypes.CDLL] = None

c_i64 = ctypes.c_int64
c_i32 = ctypes.c_int
c_f32 = ctypes.c_float
c_u64 = ctypes.c_ulonglong
c_ptr = ctypes.c_void_p

_SIGNATURES = {
    "ceb_version": ([], c_i32),
    "ceb_launch_count": ([], c_u64),
    "ceb_reset_launch_count": ([], None),
    "ceb_add_launch_count": ([c_u64], None),
    "ceb_select_scratch_bytes": ([], c_i64),
    "ceb_topk_select": ([c_ptr, c_i64, c_i64, c_ptr, c_i64, c_ptr, c_ptr, c_ptr], c_i32),
    "ceb_sketch_accumulate": ([c_ptr, c_ptr, c_f32, c_f32, c_i64, c_ptr, c_i64, c_ptr, c_i32, c_ptr,
                               c_i32, c_i64, c_ptr], c_i32),
    "ceb_sketch_accumulate_range": ([c_ptr, c_ptr, c_f32, c_f32, c_i64, c_i64, c_i64, c_ptr, c_i64, c_ptr, c_i32, c_ptr,
                                     c_i32, c_i64, c_ptr], c_i32),
    "ceb_sketch_l2_scale": ([c_ptr, c_i32, c_i64, c_f32, c_ptr, c_ptr, c_ptr], c_i32),
    "ceb_axpy_dev": ([c_ptr, c_ptr, c_i64, c_f32, c_ptr, c_i32, c_ptr], c_i32),
    "ceb_sketch_server_update": ([c_ptr, c_ptr, c_ptr, c_f32, c_f32, c_i32, c_i64, c_ptr], c_i32),
    "ceb_sketch_unsketch_topk": ([c_ptr, c_ptr, c_i64, c_ptr, c_i32, c_i32, c_i64, c_i64, c_i64, c_ptr,
                                  c_ptr, c_i64, c_ptr, c_ptr, c_ptr], c_i32),
    "ceb_sketch_query": ([c_ptr, c_ptr, c_i64, c_ptr, c_i32, c_i32, c_i64, c_i64, c_ptr, c_i64, c_ptr,
                          c_ptr], c_i32),
    "ceb_sketch_apply_zero": ([c_ptr, c_ptr, c_i64, c_ptr, c_i64, c_ptr, c_i32, c_i64, c_ptr, c_ptr,
                               c_ptr, c_f32, c_ptr, c_ptr, c_ptr], c_i32),
    "ceb_dense_server_update": ([c_ptr, c_ptr, c_f32, c_f32, c_i64, c_ptr, c_f32, c_ptr, c_f32, c_u64,
                                 c_u64, c_ptr, c_ptr], c_i32),
    "ceb_true_topk_step": ([c_ptr, c_ptr, c_ptr, c_f32, c_f32, c_i64, c_i64, c_ptr, c_f32, c_ptr, c_ptr,
                            c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_i32, c_ptr], c_i32),
    "ceb_dense_apply_zero": ([c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_f32, c_ptr, c_ptr, c_ptr,
                              c_ptr, c_i32, c_i64, c_ptr], c_i32),
    "ceb_client_step": ([c_ptr, c_f32, c_ptr, c_f32, c_ptr, c_i64, c_i64, c_ptr, c_ptr, c_i64, c_ptr,
                         c_ptr, c_ptr], c_i32),
    "ceb_scatter_add_zero": ([c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr], c_i32),
    "ceb_clip_by_norm": ([c_ptr, c_i64, c_f32, c_f32, c_ptr, c_ptr], c_i32),
    "ceb_add_gaussian_noise": ([c_ptr, c_i64, c_f32, c_u64, c_u64, c_ptr], c_i32),
    "ceb_fedavg_delta": ([c_ptr, c_ptr, c_ptr, c_f32, c_i64, c_ptr], c_i32),
    "ceb_count_changed": ([c_ptr, c_ptr, c_ptr, c_i64, c_ptr, c_ptr], c_i32),
    "ceb_tracker_record": ([c_ptr, c_i64, c_ptr, c_ptr, c_i64, c_ptr], c_i32),
    "ceb_tracker_record_dev": ([c_ptr, c_i64, c_ptr, c_ptr, c_ptr, c_ptr], c_i32),
    "ceb_sketch_apply_zero_dev": ([c_ptr, c_ptr, c_i64, c_ptr, c_i64, c_ptr, c_i32, c_i64, c_ptr, c_ptr, c_ptr, c_ptr,
                                   c_ptr, c_ptr], c_i32),
    "ceb_tracker_counts": ([c_ptr, c_i32, c_ptr, c_i32, c_ptr, c_ptr], c_i32),
}

# Entry points added by optional translation units (peer memory, tcgen05 GEMM/conv);
# bound lazily by the modules that use them.
_OPTIONAL_SIGNATURES = {}


def register_signatures(table):
    _OPTIONAL_SIGNATURES.update(table)
    if _lib is not None:
        _bind(_lib, table, optional=True)


class NativeError(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile every ``csrc/*.cu`` for sm_100a into ``lib/libceb200.so``."""
    script = os.path.join(_ROOT, "build_native.sh")
    res = subprocess.run(["bash", script], cwd=_ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                         text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
    if res.returncode != 0:
        raise NativeError("native build failed (see output above)")
    return LIB_PATH


def _bind(lib, table, optional=False):
    for name, (argtypes, restype) in table.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            if optional:
                continue
            raise NativeError("libceb200.so is stale: missing symbol %s (run build_native.sh)" % name)
        fn.argtypes = argtypes
        fn.restype = restype


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise NativeError(
                "native library %s not found; run `bash build_native.sh` (or "
                "`python -c 'import __graft_entry__ as g; g.build()'`)" % LIB_PATH)
        l = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        l.ceb_last_error.restype = ctypes.c_char_p
        _bind(l, _SIGNATURES)
        _bind(l, _OPTIONAL_SIGNATURES, optional=True)
        _lib = l
    return _lib


def last_error() -> str:
    return lib().ceb_last_error().decode("utf-8", "replace")


def check(code: int, what: str = "") -> None:
    if code != 0:
        raise NativeError("%s failed (%d): %s" % (what or "native call", code, last_error()))


def ptr(t: Optional[torch.Tensor]):
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return None
    return ctypes.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def call(name: str, *args):
    fn = getattr(lib(), name)
    check(fn(*args), name)


def launch_count() -> int:
    return int(lib().ceb_launch_count())


def add_launch_count(n: int) -> None:
    lib().ceb_add_launch_count(int(n))


def reset_launch_count() -> None:
    lib().ceb_reset_launch_count()


def require_f32_contig(t: torch.Tensor, name: str) -> torch.Tensor:
    if t.dtype != torch.float32 or not t.is_contiguous():
        raise TypeError("%s must be a contiguous float32 tensor" % name)
    return t


_scratch_cache = {}


def select_scratch(device) -> torch.Tensor:
    """Per-device, per-stream scratch for top-k style calls (reused across calls on
    the same stream, which serialises them)."""
    key = (torch.device(device).index, torch.cuda.current_stream(device).cuda_stream)
    buf = _scratch_cache.get(key)
    if buf is None:
        nbytes = int(lib().ceb_select_scratch_bytes())
        buf = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        _scratch_cache[key] = buf
    return buf
