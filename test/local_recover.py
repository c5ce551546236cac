#!/usr/bin/env python
"""Python fault-injection test with global + local models (scenario of reference test/local_recover.py,
which never sees its mock= arguments because of SURVEY 2.7 #14; here init() forwards sys.argv, so the
kills are real).  usage: local_recover.py <ndata> [mock=...]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from rabit_b200 import rabit  # noqa: E402

n = int(sys.argv[1])
rabit.init(lib="mock")
rank, world = rabit.get_rank(), rabit.get_world_size()
version, model, local = rabit.load_checkpoint(with_local=True)
if version == 0:
    model = np.ones(n)
    local = np.ones(n) * (1 + rank)
assert np.array_equal(local, model + rank), "local model corrupt"
print("[%d] start at iter=%d" % (rank, version))
for it in range(version, 3):
    z = 97 + it
    a = np.zeros(n)

    def prepare(a, it=it, z=z):
        a[:] = (np.arange(n) * (rank + 1) + it) % z + local
    out = rabit.allreduce(a, rabit.MAX, prepare_fun=prepare)
    want = np.max([(np.arange(n) * (r + 1) + it) % z + model + r for r in range(world)], axis=0)
    assert np.array_equal(out, want), "[%d] max mismatch" % rank
    msg = rabit.broadcast({"iter": it, "root": it % world} if rank == it % world else None, it % world)
    assert msg == {"iter": it, "root": it % world}
    s = rabit.allreduce((np.arange(n) * (rank + 1)) % (113 + it) + local, rabit.SUM)
    want = np.sum([(np.arange(n) * (r + 1)) % (113 + it) + model + r for r in range(world)], axis=0)
    assert np.allclose(s, want), "[%d] sum mismatch" % rank
    model = s % 1000
    local = model + rank
    rabit.checkpoint(model, local)
    print("[%d] iter %d checkpointed (version %d)" % (rank, it, rabit.version_number()))
rabit.finalize()
