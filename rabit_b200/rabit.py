"""ctypes binding with the API of the reference's ``python/rabit.py``.

Same functions and constants (``init/finalize/get_rank/get_world_size/tracker_print/
get_processor_name/broadcast/allreduce/load_checkpoint/checkpoint/version_number``,
``MAX/MIN/SUM/BITOR``, ``DTYPE_ENUM__``), so scripts written for the reference run after
``import rabit_b200.rabit as rabit``.  Differences:

* works on Python 3 (the reference calls ``.encode`` on a ``c_char_p`` and passes ``str``
  where ``bytes`` are needed);
* ``init(args=None)`` forwards ``sys.argv`` as its docstring always promised, so
  ``mock=...`` / ``rabit_*=...`` given on the command line reach the engine;
* ``allreduce`` also accepts ``torch`` CUDA tensors (see :mod:`rabit_b200.torch_binding`),
  which are reduced in place by the sm_100a kernels.
"""
import ctypes
import os
import pickle
import sys

import numpy as np

__version__ = "1.0"

_LIB = None
_LIB_PATH = None

MAX, MIN, SUM, BITOR = 0, 1, 2, 3

DTYPE_ENUM__ = {
    np.dtype("int8"): 0,
    np.dtype("uint8"): 1,
    np.dtype("int32"): 2,
    np.dtype("uint32"): 3,
    np.dtype("int64"): 4,
    np.dtype("uint64"): 5,
    np.dtype("float32"): 6,
    np.dtype("float64"): 7,
}

# 'mpi' names the library-backed, non-fault-tolerant engine: MPI in the reference (python/rabit.py:47-71,
# librabit_mpi), NCCL here (src/device/engine_nccl.cu); 'nccl' is the same library under its own name
_LIB_NAMES = {"standard": "librabit.so", "mock": "librabit_mock.so", "base": "librabit_base.so",
              "empty": "librabit_empty.so", "mpi": "librabit_nccl.so", "nccl": "librabit_nccl.so"}


def lib_dir():
    """Directory holding the in-tree shared objects (``<repo>/lib``)."""
    here = os.path.dirname(os.path.abspath(__file__))
    for cand in (os.path.join(here, "..", "lib"), os.path.join(here, "lib")):
        if os.path.isdir(cand):
            return os.path.abspath(cand)
    return os.path.abspath(os.path.join(here, "..", "lib"))


def _declare(lib):
    c = ctypes
    lib.RabitInit.argtypes = [c.c_int, c.POINTER(c.c_char_p)]
    lib.RabitInit.restype = c.c_bool
    lib.RabitFinalize.restype = c.c_bool
    lib.RabitGetRank.restype = c.c_int
    lib.RabitGetWorldSize.restype = c.c_int
    lib.RabitIsDistributed.restype = c.c_int
    lib.RabitGetRingPrevRank.restype = c.c_int
    lib.RabitVersionNumber.restype = c.c_int
    lib.RabitTrackerPrint.argtypes = [c.c_char_p]
    lib.RabitGetProcessorName.argtypes = [c.c_char_p, c.POINTER(c.c_ulong), c.c_ulong]
    lib.RabitBroadcast.argtypes = [c.c_void_p, c.c_ulong, c.c_int]
    lib.RabitAllreduce.argtypes = [c.c_void_p, c.c_size_t, c.c_int, c.c_int, c.c_void_p, c.c_void_p]
    lib.RabitAllgather.argtypes = [c.c_void_p, c.c_size_t, c.c_size_t, c.c_size_t, c.c_size_t, c.c_int]
    lib.RabitLoadCheckPoint.argtypes = [c.POINTER(c.POINTER(c.c_char)), c.POINTER(c.c_ulong),
                                        c.POINTER(c.POINTER(c.c_char)), c.POINTER(c.c_ulong)]
    lib.RabitLoadCheckPoint.restype = c.c_int
    lib.RabitCheckPoint.argtypes = [c.c_char_p, c.c_ulong, c.c_char_p, c.c_ulong]
    lib.RabitGetLastError.restype = c.c_char_p
    lib.RabitDeviceAvailable.restype = c.c_int


def _loadlib(lib="standard", lib_dll=None):
    global _LIB, _LIB_PATH
    if _LIB is not None:
        return _LIB
    if lib_dll is not None:
        _LIB = lib_dll
    else:
        if lib not in _LIB_NAMES:
            raise ValueError("unknown rabit lib %r (standard|mock|base|empty|mpi|nccl)" % (lib,))
        path = os.path.join(lib_dir(), _LIB_NAMES[lib])
        if not os.path.isfile(path):
            raise RuntimeError("cannot find %s; run `python build.py` in the repository root" % path)
        _LIB = ctypes.CDLL(path, ctypes.RTLD_GLOBAL)
        _LIB_PATH = path
    _declare(_LIB)
    return _LIB


def _unloadlib():
    global _LIB, _LIB_PATH
    _LIB = None
    _LIB_PATH = None


def get_lib():
    """The loaded ``ctypes.CDLL`` (None before :func:`init`)."""
    return _LIB


def init(args=None, lib="standard", lib_dll=None):
    """Start the engine.

    args: list of ``name=value`` strings (str or bytes); None means ``sys.argv``.
    lib: 'standard' (robust engine), 'mock' (fault injection), 'base', 'empty', 'mpi' / 'nccl' (library-backed
    comparator engine, not fault tolerant; needs the CUDA build).
    """
    if args is None:
        args = sys.argv
    L = _loadlib(lib, lib_dll)
    enc = [a if isinstance(a, bytes) else str(a).encode() for a in args]
    arr = (ctypes.c_char_p * len(enc))(*enc)
    if not L.RabitInit(len(enc), arr):
        raise RuntimeError("RabitInit failed: %s" % L.RabitGetLastError().decode())


def finalize():
    """Leave the job and unload the engine."""
    _LIB.RabitFinalize()
    _unloadlib()
    tb = sys.modules.get(__package__ + ".torch_binding")
    if tb is not None:
        tb._reset_stream_cache()


def get_rank():
    """Rank of this process, 0 .. world size - 1 (assigned by the tracker, stable across restarts)."""
    return _LIB.RabitGetRank()


def get_world_size():
    """Number of ranks in the job (1 without a tracker)."""
    return _LIB.RabitGetWorldSize()


def is_distributed():
    """True when a tracker is configured, False for a single stand-alone process."""
    return bool(_LIB.RabitIsDistributed())


def get_ring_prev_rank():
    """Rank of the ring predecessor (whose local model this rank keeps a replica of)."""
    return _LIB.RabitGetRingPrevRank()


def version_number():
    """Number of checkpoints committed so far."""
    return _LIB.RabitVersionNumber()


def tracker_print(msg):
    """Print a line in the tracker's log."""
    if not isinstance(msg, str):
        msg = str(msg)
    _LIB.RabitTrackerPrint(msg.encode())


def get_processor_name():
    """Host name of this worker."""
    buf = ctypes.create_string_buffer(256)
    n = ctypes.c_ulong()
    _LIB.RabitGetProcessorName(buf, ctypes.byref(n), 256)
    return buf.raw[:n.value].decode()


def broadcast(data, root):
    """Broadcast any picklable object from ``root``; returns it on every rank."""
    rank = get_rank()
    length = ctypes.c_ulong()
    payload = None
    if root == rank:
        assert data is not None, "need to pass in data when broadcasting"
        payload = pickle.dumps(data, protocol=pickle.HIGHEST_PROTOCOL)
        length.value = len(payload)
    _LIB.RabitBroadcast(ctypes.byref(length), ctypes.sizeof(ctypes.c_ulong), root)
    if root != rank:
        buf = ctypes.create_string_buffer(length.value)
        _LIB.RabitBroadcast(ctypes.cast(buf, ctypes.c_void_p), length.value, root)
        data = pickle.loads(buf.raw)
    else:
        buf = ctypes.create_string_buffer(payload, len(payload))
        _LIB.RabitBroadcast(ctypes.cast(buf, ctypes.c_void_p), length.value, root)
    return data


def _is_torch_tensor(x):
    mod = type(x).__module__
    return mod.startswith("torch") and hasattr(x, "data_ptr")


def allreduce(data, op, prepare_fun=None):
    """Allreduce a numpy array (returns a NEW flattened array, like the reference) or a
    torch tensor (reduced IN PLACE and returned; CUDA tensors run the fused kernels).

    prepare_fun(data) lazily fills ``data``; it is skipped when the result is recovered
    from a peer after a restart.
    """
    if _is_torch_tensor(data):
        from . import torch_binding
        return torch_binding.allreduce_(data, op, prepare_fun)
    if not isinstance(data, np.ndarray):
        raise TypeError("allreduce only takes in numpy.ndarray or torch.Tensor")
    buf = data.flatten()  # always a copy: the caller's array is only touched by prepare_fun
    if buf.dtype not in DTYPE_ENUM__:
        raise TypeError("data type %s not supported" % str(buf.dtype))
    keep = None
    fptr, farg = None, None
    if prepare_fun is not None:
        proto = ctypes.CFUNCTYPE(None, ctypes.c_void_p)

        def pfunc(_):
            prepare_fun(data)
            buf[:] = data.ravel()
        keep = proto(pfunc)
        fptr = ctypes.cast(keep, ctypes.c_void_p)
    _LIB.RabitAllreduce(buf.ctypes.data_as(ctypes.c_void_p), buf.size, DTYPE_ENUM__[buf.dtype], op, fptr, farg)
    return buf


def allgather(data):
    """Gather equally sized numpy arrays from all ranks; returns the concatenation (new API)."""
    if not isinstance(data, np.ndarray) or data.dtype not in DTYPE_ENUM__:
        raise TypeError("allgather takes a numpy.ndarray of a supported dtype")
    n, r = get_world_size(), get_rank()
    flat = data.ravel()
    out = np.empty(flat.size * n, dtype=flat.dtype)
    out[r * flat.size:(r + 1) * flat.size] = flat
    _LIB.RabitAllgather(out.ctypes.data_as(ctypes.c_void_p), out.size, r * flat.size, flat.size, flat.size,
                        DTYPE_ENUM__[flat.dtype])
    return out


def _load_blob(ptr, length):
    if length == 0:
        return None
    return pickle.loads(ctypes.string_at(ptr, length))


def load_checkpoint(with_local=False):
    """Returns (version, global_model[, local_model]); version 0 means nothing stored yet."""
    gptr = ctypes.POINTER(ctypes.c_char)()
    glen = ctypes.c_ulong()
    if with_local:
        lptr = ctypes.POINTER(ctypes.c_char)()
        llen = ctypes.c_ulong()
        version = _LIB.RabitLoadCheckPoint(ctypes.byref(gptr), ctypes.byref(glen),
                                           ctypes.byref(lptr), ctypes.byref(llen))
        if version == 0:
            return (version, None, None)
        return (version, _load_blob(gptr, glen.value), _load_blob(lptr, llen.value))
    version = _LIB.RabitLoadCheckPoint(ctypes.byref(gptr), ctypes.byref(glen), None, None)
    if version == 0:
        return (version, None)
    return (version, _load_blob(gptr, glen.value))


def checkpoint(global_model, local_model=None):
    """Commit a new version of the (picklable) models."""
    sg = pickle.dumps(global_model, protocol=pickle.HIGHEST_PROTOCOL)
    if local_model is None:
        _LIB.RabitCheckPoint(sg, len(sg), None, 0)
    else:
        sl = pickle.dumps(local_model, protocol=pickle.HIGHEST_PROTOCOL)
        _LIB.RabitCheckPoint(sg, len(sg), sl, len(sl))
