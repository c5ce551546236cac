"""GPU numerics: the real collective kernels, N virtual ranks on one GPU, vs plain torch.

Each test drives all ranks from one cooperative grid (rabit_b200/ops/emulate.py) and compares
with an fp32/fp64 torch reference of the same op.  Integer-valued inputs make float sums exact,
so the comparison is bit-for-bit and also proves every rank got the identical result
(rabit's recovery protocol depends on that, reference doc/guide.md:353-357).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

emu = None


def setup_module(module):
    from rabit_b200.ops import emulate
    global emu
    emu = emulate
    emu.load_library()  # raises if the CUDA build is missing: no silent fallback


def _inputs(n, count, dtype, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    if dtype.is_floating_point:
        return [torch.randint(-64, 64, (count,), device="cuda", generator=g).to(dtype) for _ in range(n)]
    info = torch.iinfo(dtype)
    lo, hi = max(info.min, -1000), min(info.max, 1000)
    if dtype in (torch.int8, torch.uint8):
        lo, hi = (0, 15) if dtype == torch.uint8 else (-7, 7)
    return [torch.randint(lo, hi, (count,), device="cuda", generator=g).to(dtype) for _ in range(n)]


def _reference(bufs, op):
    st = torch.stack(bufs)
    if op == 0:
        return st.max(0).values
    if op == 1:
        return st.min(0).values
    if op == 2:
        return st.sum(0).to(bufs[0].dtype)
    out = st[0].clone()
    for i in range(1, st.shape[0]):
        out |= st[i]
    return out


@pytest.mark.parametrize("algo", [1, 2, 3])
@pytest.mark.parametrize("n", [2, 8])
@pytest.mark.parametrize("count", [1, 5, 4096, 100003, 1 << 20])
def test_allreduce_sum_f32(algo, n, count):
    bufs = _inputs(n, count, torch.float32, seed=count)
    want = _reference(bufs, 2)
    st = emu.collective(emu.ALLREDUCE, bufs, algo=algo, op=2)
    assert st == [0] * n
    for r, b in enumerate(bufs):
        assert torch.equal(b, want), "rank %d differs" % r


@pytest.mark.parametrize("dtype", [torch.int8, torch.uint8, torch.int32, torch.int64, torch.float32, torch.float64])
@pytest.mark.parametrize("op", [0, 1, 2, 3])
@pytest.mark.parametrize("algo", [1, 2, 3])
def test_allreduce_all_types_ops(dtype, op, algo):
    if op == 3 and dtype.is_floating_point:
        pytest.skip("BitOR is undefined on floating point (reference c_api.cc:24-33)")
    n, count = 4, 70001
    bufs = _inputs(n, count, dtype, seed=op * 7 + algo)
    want = _reference(bufs, op)
    st = emu.collective(emu.ALLREDUCE, bufs, algo=algo, op=op)
    assert st == [0] * n
    for b in bufs:
        assert torch.equal(b, want)


@pytest.mark.parametrize("algo", [1, 2, 3])
def test_allreduce_keep_copy_and_unaligned(algo):
    """`keep` (the result-cache copy fused into the epilogue) and a user pointer that is only
    4-byte aligned."""
    n, count = 4, 33333
    raw = [torch.zeros(count + 1, device="cuda") for _ in range(n)]
    src = _inputs(n, count, torch.float32, seed=5)
    bufs = []
    for r in range(n):
        view = raw[r][1:]
        view.copy_(src[r])
        bufs.append(view)
    keeps = [torch.empty(count, device="cuda") for _ in range(n)]
    want = _reference(src, 2)
    st = emu.collective(emu.ALLREDUCE, bufs, algo=algo, op=2, keeps=keeps)
    assert st == [0] * n
    for b, k in zip(bufs, keeps):
        assert torch.equal(b, want) and torch.equal(k, want)


def test_pair_sum_histogram_f64():
    """XGBoost-style grad/hess histogram: 1000 x 256 bins of {double, double} = f64 sum over 2x."""
    n = 8
    bufs = _inputs(n, 1000 * 256 * 2, torch.float64, seed=11)
    want = _reference(bufs, 2)
    st = emu.collective(emu.ALLREDUCE, bufs, algo=2, op=2)
    assert st == [0] * n
    for b in bufs:
        assert torch.equal(b, want)


def test_repeated_ops_alternate_stages():
    """epoch parity double buffering: many back-to-back ops of different sizes and algorithms"""
    n = 4
    for i, (count, algo) in enumerate([(1000, 1), (1 << 18, 2), (77, 2), (1 << 18, 3), (4097, 2), (1 << 20, 1), (9, 1)] * 3):
        bufs = _inputs(n, count, torch.float32, seed=100 + i)
        want = _reference(bufs, 2)
        assert emu.collective(emu.ALLREDUCE, bufs, algo=algo, op=2) == [0] * n
        assert all(torch.equal(b, want) for b in bufs)


@pytest.mark.parametrize("nbytes", [1, 17, 4096, 1000003])
@pytest.mark.parametrize("root", [0, 2])
def test_broadcast(nbytes, root):
    n = 4
    bufs = [torch.randint(0, 255, (nbytes,), device="cuda", dtype=torch.uint8) for _ in range(n)]
    want = bufs[root].clone()
    keeps = [torch.empty(nbytes, device="cuda", dtype=torch.uint8) for _ in range(n)]
    assert emu.collective(emu.BROADCAST, bufs, root=root, keeps=keeps) == [0] * n
    for b, k in zip(bufs, keeps):
        assert torch.equal(b, want) and torch.equal(k, want)


@pytest.mark.parametrize("per", [1, 13, 4096, 250001])
def test_allgather(per):
    n = 4
    total = per * n
    truth = torch.randint(0, 255, (total,), device="cuda", dtype=torch.uint8)
    bufs = []
    for r in range(n):
        b = torch.zeros(total, device="cuda", dtype=torch.uint8)
        b[r * per:(r + 1) * per] = truth[r * per:(r + 1) * per]
        bufs.append(b)
    slices = [(r * per, (r + 1) * per) for r in range(n)]
    assert emu.collective(emu.ALLGATHER, bufs, slices=slices) == [0] * n
    for b in bufs:
        assert torch.equal(b, truth)


def test_allgather_uneven_slices():
    n = 3
    cuts = [0, 5, 1005, 1300]
    truth = torch.randint(0, 255, (cuts[-1],), device="cuda", dtype=torch.uint8)
    bufs = []
    for r in range(n):
        b = torch.zeros(cuts[-1], device="cuda", dtype=torch.uint8)
        b[cuts[r]:cuts[r + 1]] = truth[cuts[r]:cuts[r + 1]]
        bufs.append(b)
    slices = [(cuts[r], cuts[r + 1]) for r in range(n)]
    assert emu.collective(emu.ALLGATHER, bufs, slices=slices) == [0] * n
    for b in bufs:
        assert torch.equal(b, truth)


def test_fused_consensus_mismatch_leaves_data_untouched():
    """Ranks that disagree on the protocol code must all back out with status 1 and NOT reduce."""
    n = 4
    bufs = _inputs(n, 5000, torch.float32, seed=3)
    before = [b.clone() for b in bufs]
    st = emu.collective(emu.ALLREDUCE, bufs, algo=2, op=2, codes=[32, 32, 64, 32])
    assert st == [1] * n
    for b, a in zip(bufs, before):
        assert torch.equal(b, a)
    # and the arena is still usable afterwards
    want = _reference(bufs, 2)
    assert emu.collective(emu.ALLREDUCE, bufs, algo=2, op=2, codes=[96] * n) == [0] * n
    assert all(torch.equal(b, want) for b in bufs)


def test_library_kernels_were_launched():
    assert emu.launch_count() > 0


@pytest.mark.parametrize("dtype", [torch.int8, torch.uint8, torch.int32, torch.int64, torch.float32, torch.float64])
@pytest.mark.parametrize("op", [0, 1, 2, 3])
@pytest.mark.parametrize("n", [2, 8])
def test_low_latency_tagged_word_allreduce(dtype, op, n):
    """algo 9: data and flag travel in the same 8-byte word, no barrier (<= 32 KiB per rank)"""
    if op == 3 and dtype.is_floating_point:
        pytest.skip("BitOR is undefined on floating point")
    for nbytes in (8, 40, 4096, 12296, 32768):
        count = nbytes // torch.empty((), dtype=dtype).element_size()
        bufs = _inputs(n, count, dtype, seed=nbytes + op)
        keeps = [torch.zeros_like(b) for b in bufs]
        want = _reference(bufs, op)
        assert emu.collective(emu.ALLREDUCE, bufs, algo=9, op=op, keeps=keeps) == [0] * n
        for b, k in zip(bufs, keeps):
            assert torch.equal(b, want) and torch.equal(k, want)


def test_low_latency_consensus_mismatch_and_reuse():
    n = 4
    bufs = _inputs(n, 256, torch.float32, seed=9)
    before = [b.clone() for b in bufs]
    assert emu.collective(emu.ALLREDUCE, bufs, algo=9, op=2, codes=[5, 5, 5, 7]) == [1] * n
    assert all(torch.equal(b, a) for b, a in zip(bufs, before))
    want = _reference(bufs, 2)
    # many back-to-back ops alternate the two slot parities; mix with a barrier kernel in between
    assert emu.collective(emu.ALLREDUCE, bufs, algo=9, op=2) == [0] * n
    assert all(torch.equal(b, want) for b in bufs)
    for i in range(6):
        bufs = _inputs(n, 100 + i, torch.float32, seed=20 + i)
        want = _reference(bufs, 2)
        assert emu.collective(emu.ALLREDUCE, bufs, algo=9 if i % 3 else 1, op=2) == [0] * n
        assert all(torch.equal(b, want) for b in bufs)
