"""Launcher: host-file parsing, task placement and the ssh cluster mode (with a stand-in `ssh` that runs the
remote command on this machine, so the whole path — command line, environment, restarts — is exercised)."""
import os
import stat
import sys

import pytest

from conftest import ROOT, bin_path
from rabit_b200.parallel.launcher import launch, place_tasks, read_host_file, ssh_command


def test_host_file_and_placement(tmp_path):
    hf = tmp_path / "hosts.txt"
    hf.write_text("# two boxes\nnodeA:2\nnodeB   # one slot\n\nnodeC:3\n")
    hosts = read_host_file(str(hf))
    assert hosts == [("nodeA", 2), ("nodeB", 1), ("nodeC", 3)]
    assert place_tasks(hosts, 5) == [("nodeA", 0), ("nodeA", 1), ("nodeB", 0), ("nodeC", 0), ("nodeC", 1)]
    with pytest.raises(ValueError):
        place_tasks(hosts, 7)
    argv = ssh_command(("ssh", "-p", "22"), "nodeA", "/work dir", {"DMLC_TASK_ID": "3", "X": "a b"}, ["./prog", "k=v w"])
    assert argv[:4] == ["ssh", "-p", "22", "nodeA"]
    assert argv[4] == "cd '/work dir' && env DMLC_TASK_ID=3 X='a b' ./prog 'k=v w'"


def _fake_ssh(tmp_path):
    script = tmp_path / "fake_ssh"
    script.write_text("#!/bin/sh\n# usage: fake_ssh <host> <remote command>\necho \"ssh-to:$1\"\nexec sh -c \"$2\"\n")
    script.chmod(script.stat().st_mode | stat.S_IEXEC)
    return str(script)


def test_ssh_cluster_runs_and_restarts_workers(built, tmp_path):
    ssh = _fake_ssh(tmp_path)
    hosts = [("nodeA", 2), ("nodeB", 2)]
    env = {"PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")}
    # a kill in version 1: the launcher restarts that worker through ssh again, the job recovers
    res = launch([bin_path("model_recover"), "2000", "mock=1,1,1,0", "rabit_bootstrap_cache=1"], 4, capture=True,
                 timeout=120, hosts=hosts, ssh=(ssh,), env=env)
    assert res.ok, res.outputs
    assert res.restarts == 1
    seen = sorted(line for o in res.outputs.values() for line in o.splitlines() if line.startswith("ssh-to:"))
    assert seen == ["ssh-to:nodeA", "ssh-to:nodeA", "ssh-to:nodeB", "ssh-to:nodeB", "ssh-to:nodeB"] or \
        seen == ["ssh-to:nodeA", "ssh-to:nodeA", "ssh-to:nodeA", "ssh-to:nodeB", "ssh-to:nodeB"]


def test_ssh_cluster_local_rank_is_per_host(tmp_path):
    ssh = _fake_ssh(tmp_path)
    res = launch([sys.executable, "-c", "import os; print('LR', os.environ['DMLC_TASK_ID'], os.environ['RABIT_LOCAL_RANK'], "
                  "os.environ.get('RABIT_DEVICE_ORDINAL'))"], 3, capture=True, timeout=60,
                 hosts=[("nodeA", 2), ("nodeB", 2)], ssh=(ssh,), per_rank_gpu=True)
    got = sorted(line for o in res.outputs.values() for line in o.splitlines() if line.startswith("LR"))
    assert got == ["LR 0 0 0", "LR 1 1 1", "LR 2 0 0"]


def test_standby_workers_take_over(built):
    """--standby: a pre-started process waits inside rabit::Init and is handed the life of a worker that died, also
    twice in a row for the same task; standbys that were never needed end with the job"""
    res = launch([bin_path("model_recover"), "3000", "rabit_bootstrap_cache=1", "mock=0,0,1,0", "mock=1,1,1,0", "mock=1,1,1,1",
                  "mock=3,2,1,0"], 4, capture=True, timeout=120, standby=True)
    assert res.ok, res.outputs
    assert res.restarts == 4
    assert sorted(res.attempts.values()) == [0, 1, 1, 2]


def test_standby_python_worker_sees_attempt_number():
    code = ("import sys\nfrom rabit_b200 import rabit\nimport numpy as np\n"
            "rabit.init(sys.argv, lib='mock')\n"
            "v, m = rabit.load_checkpoint()\n"
            "for it in range(v, 2):\n"
            "    a = rabit.allreduce(np.ones(4, dtype=np.float32), rabit.SUM)\n"
            "    assert a[0] == rabit.get_world_size()\n"
            "    rabit.checkpoint(it)\n"
            "rabit.finalize()\n")
    env = {"PYTHONPATH": ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")}
    res = launch([sys.executable, "-c", code, "mock=1,1,0,0"], 3, capture=True, timeout=120, standby=True, env=env)
    assert res.ok, res.outputs
    assert res.restarts == 1
