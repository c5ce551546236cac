"""Timing helpers that follow the measurement rules of this project: device time via CUDA events,
multi-GPU numbers as the maximum over ranks."""
import numpy as np


class CudaTimer(object):
    """``with CudaTimer() as t: ...`` -> t.ms (device time on the current stream)."""

    def __enter__(self):
        import torch
        self._torch = torch
        self.e0, self.e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        self.e0.record()
        return self

    def __exit__(self, *exc):
        self.e1.record()
        self._torch.cuda.synchronize()
        self.ms = self.e0.elapsed_time(self.e1)
        return False


def max_over_ranks(value):
    """Maximum of a python float over all rabit ranks (host allreduce)."""
    from .. import rabit
    return float(rabit.allreduce(np.array([value], dtype=np.float64), rabit.MAX)[0])
