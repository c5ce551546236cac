"""GPU numerics of the streamed (chunked) allreduce and of the op counter's roll-over, N virtual ranks on one GPU.

The streamed allreduce cuts a large message into chunks, copies each into the stage and reduces it with one
kernel launch per chunk (src/device/coll_stream.cuh); the stage copy doubles as the backup that makes the op
all-or-nothing.  Reference of the op: torch sum / max over the rank axis (fp32, integer-valued inputs: exact).
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

emu = None


def setup_module(module):
    from rabit_b200.ops import emulate
    global emu
    emu = emulate
    emu.load_library()


def _inputs(n, count, dtype, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return [torch.randint(-64, 64, (count,), device="cuda", generator=g).to(dtype) for _ in range(n)]


@pytest.mark.parametrize("n", [1, 2, 8])
@pytest.mark.parametrize("count,chunk", [(1 << 20, 1 << 20), (1 << 20, 256 << 10), (3 * (1 << 18) + 5, 1 << 20),
                                         (100003, 64 << 10), (7, 4096)])
def test_stream_allreduce_sum_f32(n, count, chunk):
    bufs = _inputs(n, count, torch.float32, seed=count + n)
    keeps = [torch.zeros(count, device="cuda") for _ in range(n)]
    want = torch.stack(bufs).sum(0)
    assert emu.stream_allreduce(bufs, op=2, keeps=keeps, chunk_bytes=chunk) == [0] * n
    for r in range(n):
        assert torch.equal(bufs[r], want), "rank %d result differs" % r
        assert torch.equal(keeps[r], want), "rank %d cache copy differs" % r


@pytest.mark.parametrize("dtype,op", [(torch.float64, 2), (torch.int32, 0), (torch.int64, 1), (torch.uint8, 3),
                                      (torch.float32, 0)])
def test_stream_allreduce_types(dtype, op):
    n, count = 4, 300001
    if dtype == torch.uint8:
        g = torch.Generator(device="cuda").manual_seed(5)
        bufs = [torch.randint(0, 255, (count,), device="cuda", generator=g).to(dtype) for _ in range(n)]
    else:
        bufs = _inputs(n, count, dtype, seed=op + 11)
    st = torch.stack(bufs)
    if op == 0:
        want = st.max(0).values
    elif op == 1:
        want = st.min(0).values
    elif op == 2:
        want = st.sum(0).to(dtype)
    else:
        want = st[0].clone()
        for i in range(1, n):
            want |= st[i]
    assert emu.stream_allreduce(bufs, op=op, chunk_bytes=512 << 10) == [0] * n
    for r in range(n):
        assert torch.equal(bufs[r], want), "rank %d differs" % r


def test_stream_unaligned_user_buffer():
    n, count = 4, 200001
    base = [torch.zeros(count + 1, device="cuda") for _ in range(n)]
    bufs = [b[1:] for b in base]                       # 4-byte aligned only
    src = _inputs(n, count, torch.float32, seed=3)
    for b, s in zip(bufs, src):
        b.copy_(s)
    want = torch.stack(src).sum(0)
    assert emu.stream_allreduce(bufs, op=2, chunk_bytes=256 << 10) == [0] * n
    for r in range(n):
        assert torch.equal(bufs[r], want)
        assert float(base[r][0]) == 0.0


@pytest.mark.parametrize("fail_chunk", [0, 2])
def test_stream_failure_is_undone(fail_chunk):
    """a chunk that fails its consensus leaves: later chunks skip (poisoned op), the caller's buffers are
    restored from the stage copy although earlier chunks had already been written back"""
    n, count = 4, 1 << 20
    bufs = _inputs(n, count, torch.float32, seed=9)
    orig = [b.clone() for b in bufs]
    st = emu.stream_allreduce(bufs, op=2, chunk_bytes=1 << 20, fail_chunk=fail_chunk)
    assert st == [1] * n, st
    for r in range(n):
        assert torch.equal(bufs[r], orig[r]), "rank %d: input not restored" % r
    # and the next op works (the poisoned kernels still advanced the op counter consistently)
    want = torch.stack(orig).sum(0)
    assert emu.stream_allreduce(bufs, op=2, chunk_bytes=1 << 20) == [0] * n
    for r in range(n):
        assert torch.equal(bufs[r], want)


@pytest.mark.parametrize("algo", [1, 2, 9])
def test_op_counter_rollover(algo):
    """the 32-bit op counter in the flag barriers is compared wrap-safe: run across 2^32"""
    n, count = 4, 2048
    emu.set_epoch(n, count * 4, 0xFFFFFFFF - 3)
    for i in range(8):
        bufs = _inputs(n, count, torch.float32, seed=100 + i)
        want = torch.stack(bufs).sum(0)
        assert emu.collective(emu.ALLREDUCE, bufs, algo=algo, op=2) == [0] * n, "op %d" % i
        for b in bufs:
            assert torch.equal(b, want), "op %d" % i
    emu.set_epoch(n, count * 4, 0)
