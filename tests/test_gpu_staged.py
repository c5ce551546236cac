"""Device pointers WITHOUT the NVLink arena (rabit_device=off, or ranks on several hosts): the engines
stage them through the host TCP collectives, with the same recovery semantics. One GPU is enough:
the ranks share it."""
import os
import sys

import pytest

from conftest import ROOT, run_job

pytestmark = pytest.mark.gpu


def _run(args, n, timeout=300):
    env = {"PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")}
    return run_job([sys.executable, os.path.join(ROOT, "tests", "gpu_worker.py")] + args, n, timeout=timeout,
                   env=env, per_rank_gpu=True)


def test_device_pointers_staged_through_host():
    res = _run(["20000", "staged=1", "rabit_device=off"], 3)
    assert res.ok, {i: o[-1500:] for i, o in res.outputs.items()}
    assert res.restarts == 0


def test_device_pointers_staged_recover_after_kill():
    res = _run(["20000", "staged=1", "rabit_device=off", "mock=1,1,1,0", "rabit_timeout=1", "rabit_timeout_sec=120"],
               3, timeout=600)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
    assert res.restarts == 1


def test_base_engine_staged():
    res = _run(["20000", "staged=1", "rabit_device=off", "rabit_engine=base", "iters=1"], 2)
    assert res.ok, {i: o[-1500:] for i, o in res.outputs.items()}
