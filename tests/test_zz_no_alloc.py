"""The allocation-free allreduce entry point, called straight through the C ABI.

RabitEmuAllreduceNoAlloc is meant for hosts where a library may not allocate GPU memory on its own
(the device's memory already belongs to a framework pool or a fixed budget): it runs the two-shot
kernel out of a static __device__ array inside librabit.so.  The test feeds the same inputs to it
and to the ordinary entry point (RabitEmuCollective, which cudaMallocs its arenas) and requires
identical results — the inputs are small integers, so sums are exact and the comparison is
bit-for-bit, the tolerance all GPU tests of this repository use.  It does so twice in a row with
different sizes, because the second call finds the scratch as the first one left it.

The file name sorts last on purpose: if this test breaks, a `pytest -x` run has already been
through every other test.  It loads the library itself, so it also passes alone in a fresh process.
"""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_DTYPES = {torch.float32: 6, torch.float64: 7, torch.int32: 2, torch.int64: 4}


def _load():
    path = os.path.join(ROOT, "lib", "librabit.so")
    lib = ctypes.CDLL(path, ctypes.RTLD_GLOBAL)
    c = ctypes
    if not hasattr(lib, "RabitEmuAllreduceNoAlloc"):
        raise RuntimeError("%s has no device code: build with `python build.py`" % path)
    lib.RabitEmuAllreduceNoAlloc.argtypes = [c.c_int, c.POINTER(c.c_void_p), c.c_size_t, c.c_int, c.c_int,
                                             c.c_void_p, c.POINTER(c.c_int)]
    lib.RabitEmuAllreduceNoAlloc.restype = c.c_int
    lib.RabitEmuCollective.argtypes = [c.c_int, c.c_int, c.c_int, c.POINTER(c.c_void_p), c.POINTER(c.c_void_p),
                                       c.c_size_t, c.c_int, c.c_int, c.c_int, c.POINTER(c.c_size_t),
                                       c.POINTER(c.c_size_t), c.POINTER(c.c_uint), c.c_int, c.c_int,
                                       c.POINTER(c.c_int)]
    lib.RabitEmuCollective.restype = c.c_int
    return lib


def _ptrs(tensors):
    arr = (ctypes.c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr


def _both(lib, nranks, count, dtype, op, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    src = [torch.randint(-40, 40, (count,), device="cuda", generator=g).to(dtype) for _ in range(nranks)]
    a = [t.clone() for t in src]
    b = [t.clone() for t in src]
    torch.cuda.synchronize()
    st_a = (ctypes.c_int * nranks)()
    rc = lib.RabitEmuAllreduceNoAlloc(nranks, _ptrs(a), count, _DTYPES[dtype], op, None, st_a)
    assert rc == 0 and list(st_a) == [0] * nranks
    st_b = (ctypes.c_int * nranks)()
    rc = lib.RabitEmuCollective(0, 2, nranks, _ptrs(b), None, count, _DTYPES[dtype], op, 0, None, None, None,
                                0, 5000, st_b)
    assert rc == 0 and list(st_b) == [0] * nranks
    torch.cuda.synchronize()
    for r in range(nranks):
        assert torch.equal(a[r], b[r]), "rank %d: no-alloc result differs from the allocating entry point" % r
    # and both are right
    st = torch.stack(src)
    want = {0: st.max(0).values, 1: st.min(0).values, 2: st.sum(0).to(dtype)}[op]
    assert torch.equal(a[0], want)


def test_no_alloc_matches_allocating_entry_point_twice_in_a_row():
    lib = _load()
    _both(lib, 8, 65536, torch.float32, 2, seed=1)    # 256 KiB per rank: exactly the limit
    _both(lib, 3, 1001, torch.float32, 2, seed=2)     # second call: smaller, ragged, stale scratch behind it
    _both(lib, 4, 7777, torch.float64, 0, seed=3)
    _both(lib, 2, 5, torch.int64, 1, seed=4)


def test_no_alloc_rejects_oversized_requests():
    lib = _load()
    bufs = [torch.zeros(65537, device="cuda") for _ in range(2)]
    assert lib.RabitEmuAllreduceNoAlloc(2, _ptrs(bufs), 65537, 6, 2, None, None) == -3
    many = [torch.zeros(16, device="cuda") for _ in range(9)]
    assert lib.RabitEmuAllreduceNoAlloc(9, _ptrs(many), 16, 6, 2, None, None) == -3
    assert lib.RabitEmuAllreduceNoAlloc(2, _ptrs(bufs), 16, 6, 3, None, None) == -2   # BitOR on float
    torch.cuda.synchronize()
    assert float(bufs[0].sum()) == 0.0
