#!/usr/bin/env python
"""rabit-b200 guide: broadcast any picklable object (cf. reference guide/broadcast.py)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from rabit_b200 import rabit  # noqa: E402

rabit.init()
rank = rabit.get_rank()
s = None
if rank == 0:
    s = {"hello world": 100, 2: 3}
print("@node[%d] before-broadcast: s=\"%s\"" % (rank, str(s)))
s = rabit.broadcast(s, 0)
print("@node[%d] after-broadcast: s=\"%s\"" % (rank, str(s)))
rabit.finalize()
