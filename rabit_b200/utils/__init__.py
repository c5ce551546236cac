"""Small utilities."""
from .timing import CudaTimer, max_over_ranks  # noqa: F401
