"""Local launcher: tracker + N worker processes on this machine, with restarts.

Stands in for ``dmlc-submit --cluster local --num-workers N --local-num-attempt K``
that the reference's tests use (reference test/test.mk:16).  Each worker gets
``DMLC_TRACKER_URI/PORT``, ``DMLC_TASK_ID`` (stable across restarts, this is what lets
the tracker give a restarted worker its old rank), ``DMLC_ROLE=worker`` and
``DMLC_NUM_ATTEMPT`` (how many times it has been restarted; the mock engine matches
it against the 4th field of ``mock=rank,version,seq,ntrial``).  A worker that exits
non-zero is started again until ``max_attempts`` is reached.

    python -m rabit_b200.parallel.launcher -n 4 ./guide/basic.rabit
    python -m rabit_b200.parallel.launcher -n 8 --gpus python my_script.py
"""
import argparse
import os
import subprocess
import sys
import threading
import time

from .tracker import Tracker


class JobResult(object):
    """Outcome of :func:`launch`."""

    def __init__(self):
        self.returncodes = {}     # task index -> final return code
        self.attempts = {}        # task index -> number of restarts used
        self.outputs = {}         # task index -> captured stdout+stderr (all lives)
        self.messages = []        # TrackerPrint lines
        self.wall = 0.0

    @property
    def ok(self):
        return all(rc == 0 for rc in self.returncodes.values())

    @property
    def restarts(self):
        return sum(self.attempts.values())


def launch(cmd, nworker, max_attempts=20, env=None, capture=False, timeout=None,
           per_rank_gpu=False, verbose=False, host="127.0.0.1"):
    """Run ``cmd`` (argv list) as ``nworker`` local processes under a fresh tracker.

    per_rank_gpu: export CUDA_VISIBLE_DEVICES/RABIT_LOCAL_RANK so task i uses GPU i.
    """
    tracker = Tracker(nworker, host=host, verbose=verbose).start()
    res = JobResult()
    lock = threading.Lock()
    deadline = None if timeout is None else time.time() + timeout
    procs = {}

    def run_task(i):
        attempt = 0
        chunks = []
        while True:
            wenv = dict(os.environ)
            if env:
                wenv.update(env)
            wenv.update(tracker.worker_env())
            wenv["DMLC_TASK_ID"] = str(i)
            wenv["DMLC_NUM_ATTEMPT"] = str(attempt)
            wenv["RABIT_LOCAL_RANK"] = str(i)
            if per_rank_gpu:
                wenv["RABIT_DEVICE_ORDINAL"] = str(i)
            kw = {}
            if capture:
                kw = dict(stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
            p = subprocess.Popen(cmd, env=wenv, **kw)
            with lock:
                procs[i] = p
            try:
                remaining = None if deadline is None else max(deadline - time.time(), 0.1)
                out, _ = p.communicate(timeout=remaining)
            except subprocess.TimeoutExpired:
                p.kill()
                out, _ = p.communicate()
                if capture and out:
                    chunks.append(out.decode(errors="replace"))
                rc = -9
                break
            if capture and out:
                chunks.append(out.decode(errors="replace"))
            rc = p.returncode
            if rc == 0 or attempt >= max_attempts:
                break
            attempt += 1
        with lock:
            res.returncodes[i] = rc
            res.attempts[i] = attempt
            res.outputs[i] = "".join(chunks)

    t0 = time.time()
    threads = [threading.Thread(target=run_task, args=(i,), daemon=True) for i in range(nworker)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    res.wall = time.time() - t0
    if res.ok:
        tracker.join(5.0)
    tracker.close()
    res.messages = list(tracker.messages)
    return res


def main():
    ap = argparse.ArgumentParser(description="run a rabit program as N local workers")
    ap.add_argument("-n", "--num-workers", type=int, required=True)
    ap.add_argument("--max-attempts", "--local-num-attempt", type=int, default=20, dest="max_attempts")
    ap.add_argument("--gpus", action="store_true", help="bind task i to GPU i")
    ap.add_argument("--timeout", type=float, default=None)
    ap.add_argument("command", nargs=argparse.REMAINDER)
    args = ap.parse_args()
    cmd = [c for c in args.command if c != "--"]
    if not cmd:
        ap.error("no command given")
    res = launch(cmd, args.num_workers, max_attempts=args.max_attempts, timeout=args.timeout,
                 per_rank_gpu=args.gpus, verbose=True)
    if not res.ok:
        print("launcher: some workers failed: %r" % res.returncodes, file=sys.stderr)
    sys.exit(0 if res.ok else 1)


if __name__ == "__main__":
    main()
