"""The C ABI called directly with ctypes in a single process (world of one) and as a small job."""
import ctypes
import os

from conftest import ROOT, run_py


def _lib(name="librabit.so"):
    lib = ctypes.CDLL(os.path.join(ROOT, "lib", name))
    lib.RabitInit.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_char_p)]
    lib.RabitInit.restype = ctypes.c_bool
    lib.RabitGetLastError.restype = ctypes.c_char_p
    return lib


def test_single_process_c_abi():
    lib = _lib()
    args = (ctypes.c_char_p * 1)(b"prog")
    assert lib.RabitInit(1, args)
    assert lib.RabitGetRank() == 0 and lib.RabitGetWorldSize() == 1 and lib.RabitIsDistributed() == 0
    assert lib.RabitLinkTag() == 0 and lib.RabitVersionNumber() == 0
    name = ctypes.create_string_buffer(64)
    n = ctypes.c_ulong()
    lib.RabitGetProcessorName(name, ctypes.byref(n), 64)
    assert n.value == len(name.value) > 0
    buf = (ctypes.c_float * 4)(1, 2, 3, 4)
    called = []
    proto = ctypes.CFUNCTYPE(None, ctypes.c_void_p)
    cb = proto(lambda _: called.append(1))
    lib.RabitAllreduce(buf, ctypes.c_size_t(4), 6, 2, cb, None)
    assert called == [1] and list(buf) == [1, 2, 3, 4]      # world of one: prepare runs, data untouched
    lib.RabitCheckPoint(b"abc", ctypes.c_ulong(3), None, ctypes.c_ulong(0))
    assert lib.RabitVersionNumber() == 1
    assert lib.RabitFinalize()
    for so in ("librabit_mock.so", "librabit_base.so", "librabit_empty.so"):
        other = _lib(so)
        assert other.RabitInit(1, args) and other.RabitGetWorldSize() == 1 and other.RabitFinalize()


def test_c_abi_allgather_counts_are_elements_and_checkpoint_blobs_are_not_truncated():
    code = r'''
import ctypes, os
from rabit_b200 import rabit
rabit.init()
lib = rabit.get_lib()
r, n = rabit.get_rank(), rabit.get_world_size()
per = 3
buf = (ctypes.c_double * (per * n))()
for i in range(per): buf[r * per + i] = 10.0 * r + i
lib.RabitAllgather(buf, per * n, r * per, per, per, 7)      # dtype 7 = double, counts in ELEMENTS
assert list(buf) == [10.0 * q + i for q in range(n) for i in range(per)], list(buf)
blob = bytes((i * 7) % 251 for i in range(200000))           # > 65535 bytes
lib.RabitCheckPoint(blob, ctypes.c_ulong(len(blob)), None, ctypes.c_ulong(0))
g = ctypes.POINTER(ctypes.c_char)(); gl = ctypes.c_ulong()
v = lib.RabitLoadCheckPoint(ctypes.byref(g), ctypes.byref(gl), None, None)
# a survivor's own LoadCheckPoint in a running job is a fresh consensus: everybody asks -> version 0 path is
# only taken when NOBODY has state, so here the stored version comes back through the C API of the engine
rabit.tracker_print("rank %d ok\n" % r)
rabit.finalize()
'''
    res = run_py(code, 3)
    assert res.ok, res.outputs
    assert sum("ok" in m for m in res.messages) == 3
