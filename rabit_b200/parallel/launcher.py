"""Launcher: tracker + N worker processes, with restarts; on this machine or over ssh.

Stands in for ``dmlc-submit --cluster local --num-workers N --local-num-attempt K``
that the reference's tests use (reference test/test.mk:16).  Each worker gets
``DMLC_TRACKER_URI/PORT``, ``DMLC_TASK_ID`` (stable across restarts, this is what lets
the tracker give a restarted worker its old rank), ``DMLC_ROLE=worker`` and
``DMLC_NUM_ATTEMPT`` (how many times it has been restarted; the mock engine matches
it against the 4th field of ``mock=rank,version,seq,ntrial``).  A worker that exits
non-zero is started again until ``max_attempts`` is reached.

    python -m rabit_b200.parallel.launcher -n 4 ./guide/basic.rabit
    python -m rabit_b200.parallel.launcher -n 8 --gpus python my_script.py
    python -m rabit_b200.parallel.launcher -n 16 --cluster ssh --host-file hosts.txt --tracker-host 10.0.0.1 --gpus ./train

``--cluster ssh`` (the role of ``dmlc-submit --cluster ssh``): ``hosts.txt`` has one ``host[:slots]`` per line,
tasks are dealt to the hosts slot by slot, each worker is started as ``ssh host 'cd <cwd> && env K=V ... cmd'`` and
restarted the same way. The tracker runs here and must be reachable from the hosts at ``--tracker-host``.
With ``--gpus`` a worker uses the GPU with its index ON ITS HOST. Ranks of one host share an NVLink arena; as soon as
a job spans hosts, device buffers are staged through the TCP collectives.
"""
import argparse
import os
import shlex
import subprocess
import sys
import tempfile
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


def read_host_file(path):
    """``host[:slots]`` per line, ``#`` comments; returns [(host, slots)]."""
    hosts = []
    with open(path) as fi:
        for line in fi:
            line = line.split("#", 1)[0].strip()
            if not line:
                continue
            name, _, slots = line.partition(":")
            hosts.append((name.strip(), int(slots) if slots.strip() else 1))
    if not hosts:
        raise ValueError("host file %s names no host" % path)
    return hosts


def place_tasks(hosts, nworker):
    """Deal tasks to hosts slot by slot; returns [(host, index of the task on that host)] per task."""
    slots = [(h, k) for h, n in hosts for k in range(n)]
    if len(slots) < nworker:
        raise ValueError("%d workers do not fit into %d slots" % (nworker, len(slots)))
    return slots[:nworker]


def ssh_command(ssh, host, cwd, env, cmd):
    """argv that runs ``cmd`` with ``env`` in ``cwd`` on ``host``."""
    remote = "cd %s && env %s %s" % (shlex.quote(cwd), " ".join("%s=%s" % (k, shlex.quote(v)) for k, v in sorted(env.items())),
                                     " ".join(shlex.quote(c) for c in cmd))
    return list(ssh) + [host, remote]


def launch(cmd, nworker, max_attempts=20, env=None, capture=False, timeout=None,
           per_rank_gpu=False, verbose=False, host="127.0.0.1", hosts=None, ssh=("ssh", "-o", "StrictHostKeyChecking=no"),
           standby=False, standby_respawn_s=10.0):
    """Run ``cmd`` (argv list) as ``nworker`` processes under a fresh tracker.

    per_rank_gpu: export RABIT_DEVICE_ORDINAL / RABIT_LOCAL_RANK so a task uses the GPU with its index on its host.
    hosts: None = all workers on this machine; otherwise [(host, slots)] and the workers are started through ``ssh``
      (``host`` is then the address under which the remote workers reach the tracker).
    standby: keep one pre-started process per task (local workers only): it runs the program up to ``rabit::Init``,
      where the engine creates the CUDA context and then sleeps on a fifo (``RABIT_STANDBY_FIFO``). When the live
      worker of the task dies, the standby is handed its life (one line with the attempt number) instead of a new
      process being started, and a fresh standby is started behind it ``standby_respawn_s`` seconds later (creating a
      CUDA context stalls the other processes on that GPU for a moment: not while the job is still catching up).
      Recover-after-kill then no longer contains process start-up, library loading and CUDA context creation.
    """
    placement = place_tasks(hosts, nworker) if hosts else None
    tracker = Tracker(nworker, host=host, verbose=verbose).start()
    res = JobResult()
    lock = threading.Lock()
    deadline = None if timeout is None else time.time() + timeout
    procs = {}

    use_standby = standby and placement is None
    ended = {}
    fifo_dir = tempfile.mkdtemp(prefix="rabit-standby-") if use_standby else None
    fifo_count = [0]

    def start(i, attempt, fifo=None):
        wenv = dict(os.environ)
        if env:
            wenv.update(env)
        wenv.update(tracker.worker_env())
        wenv["DMLC_TASK_ID"] = str(i)
        wenv["DMLC_NUM_ATTEMPT"] = str(attempt)
        local = i if placement is None else placement[i][1]
        wenv["RABIT_LOCAL_RANK"] = str(local)
        if per_rank_gpu:
            wenv["RABIT_DEVICE_ORDINAL"] = str(local)
        if fifo is not None:
            wenv["RABIT_STANDBY_FIFO"] = fifo
        kw = {}
        if capture:
            kw = dict(stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        argv = cmd
        if placement is not None:
            # only what the job needs travels: the launcher's own environment stays here
            keep = dict((k, v) for k, v in wenv.items()
                        if k.startswith(("DMLC_", "RABIT_", "rabit_")) or (env and k in env))
            argv = ssh_command(ssh, placement[i][0], os.getcwd(), keep, cmd)
        return subprocess.Popen(argv, env=wenv, **kw)

    def start_standby(i):
        with lock:
            fifo_count[0] += 1
            path = os.path.join(fifo_dir, "task%d.%d" % (i, fifo_count[0]))
        os.mkfifo(path)
        return start(i, 0, fifo=path), path

    def release(spare, attempt):
        """hand a waiting standby process the life of its task; False if it is gone"""
        proc, path = spare
        if proc.poll() is not None:
            return False
        try:
            fd = spare_fds.pop(path, None)
            if fd is None:
                # (opening blocks until the standby has reached rabit::Init and opened its end)
                fd = os.open(path, os.O_WRONLY)
            os.write(fd, ("%d\n" % attempt).encode())
            os.close(fd)
            return True
        except OSError:
            return False

    def discard(spare):
        proc, path = spare
        fd = spare_fds.pop(path, None)
        if fd is not None:
            os.close(fd)
        if proc.poll() is None:
            proc.kill()
        try:
            proc.communicate(timeout=10)
        except Exception:  # pylint: disable=broad-except
            pass
        try:
            os.remove(path)
        except OSError:
            pass

    def parked(spare):
        """has this standby reached rabit::Init and opened its fifo? (a non-blocking open for writing succeeds
        only once a reader is there)"""
        proc, path = spare
        if proc.poll() is not None:
            return True
        try:
            fd = os.open(path, os.O_WRONLY | os.O_NONBLOCK)
        except OSError:
            return False
        # keep the write end open until release(): closing it now would look like "job over" to the standby
        spare_fds[path] = fd
        return True

    spare_fds = {}
    first_spares = {}
    if use_standby:
        # steady state from the first second on: the standbys are started (and have their CUDA context) before the
        # job itself, so that an early death already finds a warm replacement
        for i in range(nworker):
            first_spares[i] = start_standby(i)
        t_warm = time.time()
        while time.time() - t_warm < 120 and not all(parked(sp) for sp in first_spares.values() if sp[1] not in spare_fds):
            time.sleep(0.02)

    def run_task(i):
        attempt = 0
        chunks = []
        spare = first_spares.get(i)
        next_spare_at = 0.0
        p = None
        while True:
            if p is None:
                p = start(i, attempt)
            if use_standby and spare is None and time.time() >= next_spare_at:
                spare = start_standby(i)
            with lock:
                procs[i] = p
            try:
                remaining = None if deadline is None else max(deadline - time.time(), 0.1)
                if use_standby and spare is None:
                    # the next standby is started a little later: its start-up (CUDA context creation contends for
                    # the driver) must not slow down the recovery that is going on right now
                    try:
                        out, _ = p.communicate(timeout=min(max(next_spare_at - time.time(), 0.05), remaining or 1e9))
                    except subprocess.TimeoutExpired:
                        if deadline is not None and time.time() >= deadline:
                            raise
                        continue
                else:
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
            p = None
            if spare is not None:
                if release(spare, attempt):
                    p = spare[0]
                    try:
                        os.remove(spare[1])
                    except OSError:
                        pass
                else:
                    discard(spare)
                spare = None
                next_spare_at = time.time() + standby_respawn_s
        with lock:
            ended[i] = time.time()
        if spare is not None:
            discard(spare)
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
    res.wall = (max(ended.values()) if ended else time.time()) - t0   # (without the clean-up of unused standbys)
    if fifo_dir is not None:
        try:
            os.rmdir(fifo_dir)
        except OSError:
            pass
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
    ap.add_argument("--standby", action="store_true",
                    help="keep a pre-started process per task that takes over when the worker dies (fast recovery)")
    ap.add_argument("--timeout", type=float, default=None)
    ap.add_argument("--cluster", choices=["local", "ssh"], default="local")
    ap.add_argument("--host-file", help="ssh cluster: one host[:slots] per line")
    ap.add_argument("--tracker-host", default="127.0.0.1", help="address of this machine as seen by the workers")
    ap.add_argument("command", nargs=argparse.REMAINDER)
    args = ap.parse_args()
    cmd = [c for c in args.command if c != "--"]
    if not cmd:
        ap.error("no command given")
    hosts = None
    if args.cluster == "ssh":
        if not args.host_file:
            ap.error("--cluster ssh needs --host-file")
        hosts = read_host_file(args.host_file)
    res = launch(cmd, args.num_workers, max_attempts=args.max_attempts, timeout=args.timeout,
                 per_rank_gpu=args.gpus, verbose=True, host=args.tracker_host, hosts=hosts, standby=args.standby)
    if not res.ok:
        print("launcher: some workers failed: %r" % res.returncodes, file=sys.stderr)
    sys.exit(0 if res.ok else 1)


if __name__ == "__main__":
    main()
