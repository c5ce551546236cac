#!/usr/bin/env python
"""rabit-b200 guide: lazy preparation (cf. reference guide/lazy_allreduce.py).
   python -m rabit_b200.parallel.launcher -n 2 python guide/lazy_allreduce.py mock=0,0,1,0"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from rabit_b200 import rabit  # noqa: E402

rabit.init(lib="mock")
n = 3
rank = rabit.get_rank()
a = np.zeros(n)


def prepare(a):
    print("@node[%d] run prepare function" % rank)
    for i in range(n):
        a[i] = rank + i


print("@node[%d] before-allreduce: a=%s" % (rank, str(a)))
a = rabit.allreduce(a, rabit.MAX, prepare_fun=prepare)
print("@node[%d] after-allreduce-max: a=%s" % (rank, str(a)))
a = rabit.allreduce(a, rabit.SUM)
print("@node[%d] after-allreduce-sum: a=%s" % (rank, str(a)))
rabit.finalize()
