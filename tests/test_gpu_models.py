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
