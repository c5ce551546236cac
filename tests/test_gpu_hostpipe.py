"""Host buffers through the device (rabit_device_host_min): one rank, one GPU, kernels forced on
(rabit_device_force) so that the chunked H2D | kernel | D2H pipeline itself is exercised."""
import os
import sys

import pytest

from conftest import ROOT, run_job

pytestmark = pytest.mark.gpu


def test_host_buffer_pipeline_single_rank():
    env = {"PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")}
    res = run_job([sys.executable, os.path.join(ROOT, "tests", "gpu_worker.py"), "20000", "hostpipe=1", "iters=2",
                   "rabit_device_force=1", "rabit_device_host_min=1M", "rabit_device_host_chunk=1M"], 1,
                  timeout=300, env=env, per_rank_gpu=True)
    assert res.ok, {i: o[-2500:] for i, o in res.outputs.items()}
    # 3 MiB + 3 MiB + 4.8 MB in 1 MiB chunks, two iterations: the chunk ops are real kernel launches
    launches = [int(m.split("done,")[1].split()[0]) for m in res.messages if "kernel launches" in m]
    assert launches and launches[0] >= 2 * (4 + 4 + 5), res.messages
