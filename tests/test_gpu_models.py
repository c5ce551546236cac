"""GPU tests of the python-level helpers on one GPU (world of one, kernels forced on)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_histogram_allreducer_single_gpu():
    from rabit_b200 import rabit, torch_binding as rt
    from rabit_b200.models.histogram import HistogramAllreducer
    torch.cuda.set_device(0)
    rabit.init(["rabit_device=on", "rabit_device_force=1"])
    try:
        h = HistogramAllreducer(num_features=20, num_bins=16, device="cuda")
        rows = 500
        g = torch.Generator(device="cuda").manual_seed(0)
        bins = torch.randint(0, 16, (rows, 20), device="cuda", generator=g)
        grad = torch.randint(-5, 5, (rows,), device="cuda", generator=g).double()
        hess = torch.ones(rows, device="cuda", dtype=torch.float64)
        before = rt.launch_count()
        out = h.allreduce(lazy_build=lambda hist: h.build(bins, grad, hess))
        assert rt.launch_count() > before          # went through the device kernels
        want = torch.zeros(20, 16, 2, dtype=torch.float64, device="cuda")
        for f in range(20):
            want[f, :, 0].index_add_(0, bins[:, f], grad)
            want[f, :, 1].index_add_(0, bins[:, f], hess)
        assert torch.equal(out, want)
        # checkpoint helpers round-trip a module through rabit's versioned checkpoint API
        from rabit_b200.models import checkpoint as ck
        lin = torch.nn.Linear(4, 3).cuda()
        version, extra = ck.load(lin)
        assert version == 0
        ck.save(lin, extra={"step": 7})
        assert rabit.version_number() == 1
    finally:
        rabit.finalize()


def test_collectives_replay_from_a_cuda_graph():
    """The op epoch lives in device memory, so a captured graph of collectives can be replayed and
    freely mixed with eager launches (world of one, kernels forced on)."""
    from rabit_b200 import rabit, torch_binding as rt
    torch.cuda.set_device(0)
    rabit.init(["rabit_device=on", "rabit_device_force=1"])
    try:
        x = torch.arange(10000, device="cuda", dtype=torch.float32)
        y = torch.ones(300, device="cuda", dtype=torch.float64)
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            rt.allreduce_async_(x)          # warm up outside capture (arena sizing, lazy init)
            rt.allreduce_async_(y)
        rt.synchronize()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        before = rt.launch_count()
        with torch.cuda.graph(g, stream=side):
            rt.allreduce_async_(x)
            rt.allreduce_async_(y, rabit.MAX)
            rt.broadcast_async_(x, 0)
        captured = rt.launch_count() - before
        assert captured == 3
        for _ in range(5):
            g.replay()
        torch.cuda.synchronize()
        rt.allreduce_async_(x)              # eager launch after replays: the epochs still line up
        rt.synchronize()
        assert torch.equal(x, torch.arange(10000, device="cuda", dtype=torch.float32))
        assert torch.equal(y, torch.ones(300, device="cuda", dtype=torch.float64))
    finally:
        rabit.finalize()
