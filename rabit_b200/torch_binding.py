"""PyTorch binding: rabit collectives on ``torch`` tensors.

CUDA tensors are handed to the engine by address; the engine recognises device pointers
and runs the fused sm_100a kernels on them (robust path: consensus fused into the kernel,
result cached on the GPU).  CPU tensors go through their numpy view to the host engine.
``*_async`` variants only enqueue on the current CUDA stream (no fault tolerance, no host
synchronisation) and are what training loops and CUDA-graph capture want.

No counterpart in the reference (it predates GPU tensors); the semantics of each call are
those of ``rabit::Allreduce/Broadcast/Allgather`` (reference include/rabit/rabit.h).
"""
import ctypes
import os

import torch

from . import rabit as _rabit

# The native binding (src/python/torch_ext.cc, built by build.py as rabit_b200/_torch_ext*.so): same calls without
# the per-call ctypes / Python work (~4-5 us each). RABIT_TORCH_CTYPES=1 forces the ctypes path below.
try:
    if os.environ.get("RABIT_TORCH_CTYPES"):
        raise ImportError("disabled")
    from . import _torch_ext as _ext
except ImportError:
    _ext = None


_bound = [None]     # the ctypes library whose entry points the native module currently calls


def native_binding():
    """True when the pybind11 / libtorch extension carries the tensor calls (False: ctypes fallback)."""
    return _ext is not None

_TORCH_DTYPE = {
    torch.int8: 0, torch.uint8: 1, torch.int32: 2, torch.int64: 4,
    torch.float32: 6, torch.float64: 7,
}
_ITEMSIZE = {0: 1, 1: 1, 2: 4, 3: 4, 4: 8, 5: 8, 6: 4, 7: 8}


def _lib():
    lib = _rabit.get_lib()
    if lib is None:
        raise RuntimeError("call rabit_b200.rabit.init() first")
    if not getattr(lib, "_rabit_device_declared", False):
        c = ctypes
        lib.RabitDeviceAllreduceAsync.argtypes = [c.c_void_p, c.c_size_t, c.c_int, c.c_int]
        lib.RabitDeviceBroadcastAsync.argtypes = [c.c_void_p, c.c_size_t, c.c_int]
        lib.RabitDeviceAllgatherAsync.argtypes = [c.c_void_p, c.c_size_t, c.c_size_t, c.c_size_t]
        lib.RabitDeviceSetStream.argtypes = [c.c_void_p]
        lib.RabitDeviceLaunchCount.restype = c.c_ulonglong
        lib.RabitDeviceDescribe.argtypes = [c.c_char_p, c.c_size_t]
        lib._rabit_device_declared = True
    if _ext is not None and _bound[0] is not lib:
        # hand the native module the entry points of THIS library (a process may hold several librabit*.so)
        names = ["RabitDeviceAllreduceAsync", "RabitDeviceBroadcastAsync", "RabitDeviceAllgatherAsync", "RabitDeviceSynchronize",
                 "RabitDeviceSetStream", "RabitAllreduce", "RabitBroadcast", "RabitDeviceLaunchCount"]
        _ext.bind_library(dict((n, ctypes.cast(getattr(lib, n), ctypes.c_void_p).value) for n in names))
        _bound[0] = lib
    return lib


def _dtype_enum(t):
    if t.dtype not in _TORCH_DTYPE:
        raise TypeError("dtype %s is not supported by rabit allreduce" % t.dtype)
    return _TORCH_DTYPE[t.dtype]


def _check(t):
    if not t.is_contiguous():
        raise ValueError("rabit collectives need contiguous tensors")


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_last_stream = [None]


def _bind_stream(lib, t):
    """Make the engine launch on torch's current stream of the tensor's device (cheap: raw handle, and
    the C call is skipped while the handle does not change)."""
    if not t.is_cuda:
        return
    if _raw_stream is not None:
        h = _raw_stream(t.device.index if t.device.index is not None else torch.cuda.current_device())
    else:
        h = torch.cuda.current_stream(t.device).cuda_stream
    if h != _last_stream[0]:
        lib.RabitDeviceSetStream(ctypes.c_void_p(h))
        _last_stream[0] = h


def describe():
    """One line about the device backend (arena kind, NVLS on/off ...)."""
    lib = _lib()
    buf = ctypes.create_string_buffer(1024)
    lib.RabitDeviceDescribe(buf, 1024)
    return buf.value.decode()


def _reset_stream_cache():
    _last_stream[0] = None
    if _ext is not None:
        _ext.reset_stream_cache()
        _ext.unbind_library()
        _bound[0] = None


def device_active():
    return bool(_lib().RabitDeviceActive())


def launch_count():
    """Number of this library's kernels launched so far in this process."""
    return int(_lib().RabitDeviceLaunchCount())


def allreduce_(tensor, op=_rabit.SUM, prepare_fun=None):
    """Fault-tolerant in-place allreduce; returns ``tensor``.  Blocks until the result is in
    ``tensor`` on this rank (rabit semantics)."""
    lib = _lib()
    _check(tensor)
    if _ext is not None and prepare_fun is None:
        return _ext.allreduce_(tensor, op)
    keep = None
    fptr = None
    if prepare_fun is not None:
        proto = ctypes.CFUNCTYPE(None, ctypes.c_void_p)

        def pfunc(_):
            prepare_fun(tensor)
            if tensor.is_cuda:
                torch.cuda.current_stream(tensor.device).synchronize()
        keep = proto(pfunc)
        fptr = ctypes.cast(keep, ctypes.c_void_p)
    if tensor.is_cuda:
        _bind_stream(lib, tensor)
    lib.RabitAllreduce(ctypes.c_void_p(tensor.data_ptr()), tensor.numel(), _dtype_enum(tensor), op, fptr, None)
    return tensor


def broadcast_(tensor, root=0):
    """Fault-tolerant in-place broadcast of the tensor's bytes from ``root``."""
    lib = _lib()
    _check(tensor)
    if _ext is not None:
        return _ext.broadcast_(tensor, root)
    if tensor.is_cuda:
        _bind_stream(lib, tensor)
    lib.RabitBroadcast(ctypes.c_void_p(tensor.data_ptr()), tensor.numel() * tensor.element_size(), root)
    return tensor


def allgather_(out, begin, count, prev_count=None):
    """Fault-tolerant ring-ordered allgather: this rank owns ``out.view(-1)[begin:begin+count]``."""
    lib = _lib()
    _check(out)
    if out.is_cuda:
        _bind_stream(lib, out)
    if prev_count is None:
        prev_count = count
    lib.RabitAllgather(ctypes.c_void_p(out.data_ptr()), out.numel(), begin, count, prev_count, _dtype_enum(out))
    return out


def allreduce_async_(tensor, op=_rabit.SUM):
    """Enqueue an in-place allreduce on the current CUDA stream and return immediately."""
    lib = _lib()
    if _ext is not None:
        return _ext.allreduce_async_(tensor, op)
    _check(tensor)
    if not tensor.is_cuda:
        raise ValueError("allreduce_async_ needs a CUDA tensor")
    _bind_stream(lib, tensor)
    rc = lib.RabitDeviceAllreduceAsync(ctypes.c_void_p(tensor.data_ptr()), tensor.numel(), _dtype_enum(tensor), op)
    if rc != 0:
        raise RuntimeError("RabitDeviceAllreduceAsync failed (%d)" % rc)
    return tensor


def broadcast_async_(tensor, root=0):
    lib = _lib()
    if _ext is not None:
        return _ext.broadcast_async_(tensor, root)
    _check(tensor)
    _bind_stream(lib, tensor)
    rc = lib.RabitDeviceBroadcastAsync(ctypes.c_void_p(tensor.data_ptr()), tensor.numel() * tensor.element_size(), root)
    if rc != 0:
        raise RuntimeError("RabitDeviceBroadcastAsync failed (%d)" % rc)
    return tensor


def allgather_async_(out, begin, count):
    """Enqueue a ring-ordered allgather on the current CUDA stream: this rank owns
    ``out.view(-1)[begin:begin+count]``, every other range is filled in from its owner."""
    lib = _lib()
    if _ext is not None:
        return _ext.allgather_async_(out, begin, count)
    _check(out)
    if not out.is_cuda:
        raise ValueError("allgather_async_ needs a CUDA tensor")
    _bind_stream(lib, out)
    es = out.element_size()
    rc = lib.RabitDeviceAllgatherAsync(ctypes.c_void_p(out.data_ptr()), out.numel() * es, begin * es, (begin + count) * es)
    if rc != 0:
        raise RuntimeError("RabitDeviceAllgatherAsync failed (%d)" % rc)
    return out


def synchronize():
    """Wait for the enqueued collectives; raises if a peer was lost or timed out."""
    lib = _lib()
    if _ext is not None:
        return _ext.synchronize()
    rc = lib.RabitDeviceSynchronize()
    if rc != 0:
        raise RuntimeError("device collective failed: %s" %
                           {1: "consensus mismatch", 2: "peer timeout", 3: "peer lost"}.get(rc, rc))


def set_tuning(algo=None, blocks=0, trace=False):
    """Runtime override of the allreduce algorithm ('auto', 'oneshot', 'twoshot', 'tma', 'nvls'),
    the block cap (0 = default) and the phase trace (see :func:`get_trace`)."""
    lib = _lib()
    lib.RabitDeviceSetTuning.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int]
    lib.RabitDeviceSetTuning(None if algo is None else algo.encode(), int(blocks), int(bool(trace)))


def get_trace():
    """Phase timestamps (ns, %globaltimer) of block 0 of the last traced kernel:
    [start, staged, entry barrier passed, reduce-scatter done, mid barrier passed, end]."""
    lib = _lib()
    out = (ctypes.c_ulonglong * 6)()
    lib.RabitDeviceGetTrace(out)
    return list(out)


class _SymBlock(object):
    """CUDA-array-interface view of a piece of the symmetric heap (keeps nothing alive: the heap
    belongs to the engine and lives until finalize())."""

    def __init__(self, ptr, numel, dtype):
        typestr = {torch.float32: "<f4", torch.float64: "<f8", torch.int32: "<i4", torch.int64: "<i8",
                   torch.uint8: "|u1", torch.int8: "|i1"}[dtype]
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (ptr, False),
                                         "version": 2, "strides": None}


def symmetric_empty(numel, dtype=torch.float32, device=None):
    """A 1-D CUDA tensor inside the engine's symmetric heap (``rabit_device_heap=<bytes>`` at init).

    Collective: every rank must allocate the same sizes in the same order.  Collectives on such a
    tensor are zero-copy (peers address it directly over NVLink / through the multicast mapping).
    """
    lib = _lib()
    lib.RabitDeviceSymAlloc.restype = ctypes.c_void_p
    lib.RabitDeviceSymAlloc.argtypes = [ctypes.c_size_t]
    nbytes = numel * torch.empty((), dtype=dtype).element_size()
    ptr = lib.RabitDeviceSymAlloc(nbytes)
    if not ptr:
        raise MemoryError("symmetric heap exhausted or not configured (pass rabit_device_heap=<bytes> to init)")
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device())
    return torch.as_tensor(_SymBlock(ptr, numel, dtype), device=device)
