"""Single-GPU emulation of multi-rank collectives (the real kernels, N virtual ranks).

Kernels that wait on each other must never run as separate launches on one GPU, so the
unit tests, ``smoke()`` and the 1-GPU benchmark drive all ranks from ONE cooperative grid
(``gridDim.y`` = rank).  See src/device/device_capi.inc ``RabitEmuCollective``.
"""
import ctypes
import os

import torch

from .. import rabit as _rabit

ALLREDUCE, BROADCAST, ALLGATHER = 0, 1, 2
ONESHOT, TWOSHOT, TMA = 1, 2, 3
_TORCH_DTYPE = {torch.int8: 0, torch.uint8: 1, torch.int32: 2, torch.int64: 4,
                torch.float32: 6, torch.float64: 7}
_lib_cache = None


def load_library():
    """The CUDA-enabled librabit.so; raises loudly if it was built without nvcc."""
    global _lib_cache
    if _lib_cache is not None:
        return _lib_cache
    lib = _rabit.get_lib()
    if lib is None:
        path = os.path.join(_rabit.lib_dir(), "librabit.so")
        if not os.path.isfile(path):
            raise RuntimeError("%s missing: run `python build.py`" % path)
        lib = ctypes.CDLL(path, ctypes.RTLD_GLOBAL)
    if not hasattr(lib, "RabitEmuCollective"):
        raise RuntimeError("librabit.so was built without CUDA (python build.py --no-cuda); "
                           "the device kernels are required on a GPU box")
    c = ctypes
    lib.RabitEmuCollective.argtypes = [c.c_int, c.c_int, c.c_int, c.POINTER(c.c_void_p), c.POINTER(c.c_void_p),
                                       c.c_size_t, c.c_int, c.c_int, c.c_int, c.POINTER(c.c_size_t),
                                       c.POINTER(c.c_size_t), c.POINTER(c.c_uint), c.c_int, c.c_int,
                                       c.POINTER(c.c_int)]
    lib.RabitEmuCollective.restype = c.c_int
    lib.RabitDeviceLaunchCount.restype = c.c_ulonglong
    lib.RabitEmuStreamAllreduce.argtypes = [c.c_int, c.POINTER(c.c_void_p), c.POINTER(c.c_void_p), c.c_size_t, c.c_int,
                                            c.c_int, c.c_size_t, c.c_int, c.c_int, c.POINTER(c.c_int)]
    lib.RabitEmuStreamAllreduce.restype = c.c_int
    lib.RabitEmuSetEpoch.argtypes = [c.c_int, c.c_size_t, c.c_uint]
    lib.RabitEmuSetEpoch.restype = c.c_int
    _lib_cache = lib
    return lib


def _ptrs(tensors):
    arr = (ctypes.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = None if t is None else t.data_ptr()
    return arr


def collective(kind, bufs, algo=ONESHOT, op=2, root=0, keeps=None, slices=None, codes=None, blocks=0,
               timeout_ms=5000):
    """Run one emulated collective over ``bufs`` (one CUDA tensor per virtual rank, in place).

    Returns the list of per-rank kernel status codes (0 ok, 1 consensus mismatch, 2 timeout).
    """
    lib = load_library()
    n = len(bufs)
    t0 = bufs[0]
    for t in bufs:
        assert t.is_cuda and t.is_contiguous() and t.dtype == t0.dtype
    dtype = _TORCH_DTYPE[t0.dtype]
    size = t0.numel() if kind == ALLREDUCE else t0.numel() * t0.element_size()
    sb = se = None
    if slices is not None:
        sb = (ctypes.c_size_t * n)(*[s[0] for s in slices])
        se = (ctypes.c_size_t * n)(*[s[1] for s in slices])
    cc = (ctypes.c_uint * n)(*codes) if codes is not None else None
    status = (ctypes.c_int * n)()
    torch.cuda.synchronize()
    rc = lib.RabitEmuCollective(kind, algo, n, _ptrs(bufs), _ptrs(keeps) if keeps is not None else None,
                                size, dtype, op, root, sb, se, cc, blocks, timeout_ms, status)
    if rc != 0:
        raise RuntimeError("RabitEmuCollective failed with %d" % rc)
    return list(status)


def stream_allreduce(bufs, op=2, keeps=None, chunk_bytes=1 << 20, fail_chunk=-1, blocks=0):
    """The chunked (streamed) allreduce over ``bufs``: every chunk is copied into the emulated rank's stage
    and reduced by one launch of the StreamP2P kernel.  ``fail_chunk`` >= 0 injects a consensus mismatch in
    that chunk; the call then undoes the op from the stage copy like the multi-GPU path does.

    Returns the per-rank status of the op (0 ok, 1 mismatch, 2 timeout).
    """
    lib = load_library()
    n = len(bufs)
    t0 = bufs[0]
    for t in bufs:
        assert t.is_cuda and t.is_contiguous() and t.dtype == t0.dtype
    status = (ctypes.c_int * n)()
    torch.cuda.synchronize()
    rc = lib.RabitEmuStreamAllreduce(n, _ptrs(bufs), _ptrs(keeps) if keeps is not None else None, t0.numel(),
                                     _TORCH_DTYPE[t0.dtype], op, chunk_bytes, fail_chunk, blocks, status)
    if rc != 0:
        raise RuntimeError("RabitEmuStreamAllreduce failed with %d" % rc)
    return list(status)


def set_epoch(nranks, payload_bytes, epoch):
    """Test hook: move the op counter of the emulated ranks to ``epoch`` (flag rows made consistent with it)."""
    rc = load_library().RabitEmuSetEpoch(nranks, payload_bytes, epoch & 0xFFFFFFFF)
    if rc != 0:
        raise RuntimeError("RabitEmuSetEpoch failed with %d" % rc)


def launch_count():
    return int(load_library().RabitDeviceLaunchCount())
