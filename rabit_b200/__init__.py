"""rabit-b200: reliable allreduce / broadcast for B200 (sm_100a) with rabit's API and semantics.

    from rabit_b200 import rabit            # reference-compatible python API (numpy + torch tensors)
    from rabit_b200 import torch_binding    # stream-ordered collectives, symmetric buffers
    from rabit_b200.parallel import tracker, launcher
    from rabit_b200.ops import emulate      # N virtual ranks on one GPU (tests, smoke)
    from rabit_b200.models import histogram, checkpoint
"""
__version__ = "0.1.0"
