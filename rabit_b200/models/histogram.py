"""XGBoost-style gradient-histogram aggregation (BASELINE config 3).

Each worker builds, per tree level, a (features x bins) histogram of {grad, hess} float64 pairs from
its data shard; the level's split search needs the element-wise sum over workers.  With the reference
that is ``rabit::Reducer<GradPair, PairSum>`` over 16-byte structs on the CPU
(reference include/rabit/internal/rabit-inl.h:255-288); here the pair sum is a float64 Sum over twice
the elements, which the fused kernels (and ``multimem.ld_reduce ... add.f64`` on NVLS) execute directly.
"""
import torch

from .. import rabit


class HistogramAllreducer(object):
    """Owns a (features, bins, 2) float64 device buffer and allreduces it per level."""

    def __init__(self, num_features=1000, num_bins=256, device=None, symmetric=False):
        self.shape = (num_features, num_bins, 2)
        numel = num_features * num_bins * 2
        if symmetric:
            from .. import torch_binding as rt
            self.flat = rt.symmetric_empty(numel, torch.float64)
        else:
            self.flat = torch.zeros(numel, dtype=torch.float64, device=device or "cuda")
        self.hist = self.flat.view(*self.shape)

    def build(self, bin_index, grad, hess):
        """bin_index: (rows, features) int64 bin of every row per feature; grad/hess: (rows,) float64."""
        self.flat.zero_()
        f = torch.arange(self.shape[0], device=bin_index.device).unsqueeze(0).expand_as(bin_index)
        idx = ((f * self.shape[1] + bin_index) * 2).reshape(-1)
        self.flat.index_add_(0, idx, grad.unsqueeze(1).expand_as(bin_index).reshape(-1))
        self.flat.index_add_(0, idx + 1, hess.unsqueeze(1).expand_as(bin_index).reshape(-1))
        return self.hist

    def allreduce(self, lazy_build=None):
        """Sum the histogram over all workers (fault tolerant; `lazy_build(hist)` is skipped when the
        result is recovered from a peer's cache)."""
        prep = None if lazy_build is None else (lambda _t: lazy_build(self.hist))
        rabit.allreduce(self.flat, rabit.SUM, prep)
        return self.hist
