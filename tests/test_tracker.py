"""Tracker topology invariants the workers rely on."""
import pytest

from rabit_b200.parallel.tracker import build_topology


@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 7, 8, 10, 16, 31, 64])
def test_topology_invariants(n):
    tree, parent, ring = build_topology(n)
    assert sorted(tree) == list(range(n))
    if n == 1:
        assert tree[0] == [] and parent[0] == -1
        return
    # ring is r -> r+1 (mod n): required by the ring collectives
    for r in range(n):
        assert ring[r] == ((r - 1) % n, (r + 1) % n)
    # tree: symmetric adjacency, n-1 edges, connected, exactly one root, parent is a neighbour
    edges = set()
    for r, nb in tree.items():
        assert len(nb) <= 3
        for q in nb:
            assert r in tree[q]
            edges.add((min(r, q), max(r, q)))
    assert len(edges) == n - 1
    roots = [r for r in range(n) if parent[r] == -1]
    assert len(roots) == 1
    seen, todo = {roots[0]}, [roots[0]]
    while todo:
        for q in tree[todo.pop()]:
            if q not in seen:
                seen.add(q)
                todo.append(q)
    assert len(seen) == n
    for r in range(n):
        if parent[r] != -1:
            assert parent[r] in tree[r]


def _raw(tracker):
    import socket
    s = socket.create_connection((tracker.host, tracker.port))
    s.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
    return s


def test_tracker_survives_broken_clients(built):
    """Half-open handshakes, a wrong magic and a worker that dies while parked for the batch rank assignment
    must not take the (single-threaded) tracker down; the job that follows still rendezvouses and its
    replacement worker gets the rank of the one that vanished."""
    import os
    import struct
    import subprocess
    from conftest import bin_path
    from rabit_b200.parallel.tracker import Tracker, MAGIC

    t = Tracker(2).start()
    try:
        s = _raw(t); s.close()                                         # connects, says nothing
        s = _raw(t); s.sendall(struct.pack("@i", 0x1234)); s.close()   # wrong magic
        s = _raw(t); s.sendall(struct.pack("@i", MAGIC)); s.recv(4); s.sendall(struct.pack("@i", -1)); s.close()
        # a first-time worker that parks for the batch assignment and then dies
        ghost = _raw(t)
        task = b"0"
        ghost.sendall(struct.pack("@i", MAGIC))
        ghost.recv(4)
        ghost.sendall(struct.pack("@iii", -1, 2, len(task)) + task + struct.pack("@i", 5) + b"start")
        procs = []
        env = dict(os.environ)
        env.update(t.worker_env())
        env["DMLC_TASK_ID"] = "1"
        env["DMLC_NUM_ATTEMPT"] = "0"
        procs.append(subprocess.Popen([bin_path("basic")], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
        import time
        time.sleep(0.5)
        ghost.close()      # the tracker notices when it tries to hand it rank 0
        env0 = dict(env)
        env0["DMLC_TASK_ID"] = "0"
        env0["rabit_timeout"] = "1"
        env["rabit_timeout"] = "1"
        procs.append(subprocess.Popen([bin_path("basic")], env=env0, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
        outs = [p.communicate(timeout=120)[0].decode(errors="replace") for p in procs]
        assert [p.returncode for p in procs] == [0, 0], outs
        assert t.join(30), "tracker did not see both shutdowns"
        assert t.error is None
    finally:
        t.close()


def test_tracker_tells_late_recoverer_that_the_job_is_over():
    """once a rank has shut down a full re-rendezvous can never complete: "recover" is answered with rank -2"""
    import struct
    from rabit_b200.parallel.tracker import Tracker, MAGIC

    def hello(t, rank, cmd):
        s = _raw(t)
        s.sendall(struct.pack("@i", MAGIC))
        assert struct.unpack("@i", s.recv(4))[0] == MAGIC
        task = str(rank).encode()
        s.sendall(struct.pack("@iii", rank, 2, len(task)) + task + struct.pack("@i", len(cmd)) + cmd.encode())
        return s

    t = Tracker(2).start()
    try:
        hello(t, 0, "shutdown").close()
        s = hello(t, 1, "recover")
        got = b""
        while len(got) < 12:
            chunk = s.recv(12 - len(got))
            assert chunk, "tracker closed without answering"
            got += chunk
        assert struct.unpack("@3i", got) == (-2, -1, 2)
        s.close()
        hello(t, 1, "shutdown").close()
        assert t.join(20) and t.error is None
    finally:
        t.close()


def test_replacement_worker_after_the_job_ended_exits_cleanly(built):
    """a worker killed after the final barrier is restarted by the launcher; its peers are gone: the tracker says so,
    the replacement exits 0 and the tracker counts the rank as finished"""
    import os
    import struct
    import subprocess
    from conftest import bin_path
    from rabit_b200.parallel.tracker import Tracker, MAGIC

    t = Tracker(2).start()
    try:
        env = dict(os.environ)
        env.update(t.worker_env())
        procs = []
        for i in range(2):
            e = dict(env, DMLC_TASK_ID=str(i), DMLC_NUM_ATTEMPT="0")
            procs.append(subprocess.Popen([bin_path("basic")], env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
        for p in procs:
            p.communicate(timeout=60)
        assert [p.returncode for p in procs] == [0, 0]
        assert t.join(20)
    finally:
        t.close()
    # same, but rank 1 "dies" right after the last barrier: here, its shutdown notice never reaches the tracker
    t = Tracker(2).start()
    try:
        s = _raw(t)   # rank 0 says goodbye
        s.sendall(struct.pack("@i", MAGIC)); s.recv(4)
        s.sendall(struct.pack("@iii", 0, 2, 1) + b"0" + struct.pack("@i", 8) + b"shutdown")
        s.close()
        e = dict(os.environ, DMLC_TASK_ID="1", DMLC_NUM_ATTEMPT="1")
        e.update(t.worker_env())
        p = subprocess.Popen([bin_path("basic")], env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        out = p.communicate(timeout=60)[0].decode(errors="replace")
        assert p.returncode == 0 and "nothing to re-join" in out, out
    finally:
        t.close()


def test_worker_parked_in_accept_learns_that_the_job_is_over(built):
    """The job's last barrier can complete on some ranks and break on others; a rank that passed it leaves for good.
    A worker that is then waiting for that rank to dial it asks the tracker ("status") once a second and leaves as
    well, instead of sitting out its whole accept timeout (60 s here, 300 s by default)."""
    import os
    import struct
    import subprocess
    import time
    from conftest import bin_path
    from rabit_b200.parallel.tracker import Tracker, MAGIC

    def hello(t, rank, task, cmd):
        s = _raw(t)
        s.sendall(struct.pack("@i", MAGIC))
        assert struct.unpack("@i", s.recv(4))[0] == MAGIC
        s.sendall(struct.pack("@iii", rank, 2, len(task)) + task + struct.pack("@i", len(cmd)) + cmd.encode())
        return s

    t = Tracker(2).start()
    try:
        ghost = hello(t, -1, b"0", "start")          # will be rank 0: takes its assignment, then never shows up as a peer
        env = dict(os.environ)
        env.update(t.worker_env())
        env.update({"DMLC_TASK_ID": "1", "DMLC_NUM_ATTEMPT": "0", "rabit_timeout": "1", "rabit_timeout_sec": "60"})
        t0 = time.time()
        worker = subprocess.Popen([bin_path("basic")], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        ghost.settimeout(30)
        ghost.recv(4096)                              # rank, parent, world, neighbours, ring
        ghost.sendall(struct.pack("@i", 0))           # no surviving links
        ghost.recv(8)                                 # nothing to dial, one peer will dial us
        ghost.sendall(struct.pack("@ii", 0, 1))       # no errors; a port where nobody listens
        ghost.close()
        time.sleep(1.5)                               # the worker failed to reach "rank 0" and is parked in accept by now
        hello(t, 0, b"0", "shutdown").close()         # rank 0 has left for good
        out = worker.communicate(timeout=30)[0].decode(errors="replace")
        assert worker.returncode == 0, out
        assert time.time() - t0 < 20, out
        assert "nothing to re-join" in out, out
    finally:
        t.close()


def test_worker_the_tracker_forgot_starts_over(built):
    """The tracker drops a parked worker from its list when somebody could not reach it (it takes it for dead). If the
    worker is alive it would wait in accept() for diallers that are never sent: `status` tells it, and it starts the
    rendezvous over within seconds instead of sitting out its accept timeout (60 s here)."""
    import os
    import struct
    import subprocess
    import time
    from conftest import bin_path
    from rabit_b200.parallel.tracker import Tracker, MAGIC

    def hello(t, rank, task, cmd):
        s = _raw(t)
        s.settimeout(30)
        s.sendall(struct.pack("@i", MAGIC))
        assert struct.unpack("@i", s.recv(4))[0] == MAGIC
        s.sendall(struct.pack("@iii", rank, 2, len(task)) + task + struct.pack("@i", len(cmd)) + cmd.encode())
        return s

    t = Tracker(2).start()
    try:
        ghost = hello(t, -1, b"0", "start")           # rank 0; gives a port where nobody listens
        env = dict(os.environ)
        env.update(t.worker_env())
        env.update({"DMLC_TASK_ID": "1", "DMLC_NUM_ATTEMPT": "0", "rabit_timeout": "1", "rabit_timeout_sec": "60"})
        t0 = time.time()
        worker = subprocess.Popen([bin_path("basic")], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        ghost.recv(4096)
        ghost.sendall(struct.pack("@i", 0))
        ghost.recv(8)
        ghost.sendall(struct.pack("@ii", 0, 1))
        ghost.close()
        time.sleep(1.0)                               # the worker could not reach "rank 0" and is parked in accept
        again = hello(t, 0, b"0", "recover")          # "rank 0" comes back, is pointed at the worker ...
        again.recv(4096)
        again.sendall(struct.pack("@i", 0))
        again.recv(4096)                              # ... (its address) ...
        again.sendall(struct.pack("@i", 1))           # ... and claims it could not reach it: the tracker forgets the worker
        again.sendall(struct.pack("@i", 0))           # next round of the link plan: still no links
        again.recv(8)                                 # nobody to dial any more, one peer is to dial us
        again.sendall(struct.pack("@ii", 0, 2))
        again.close()
        out = worker.communicate(timeout=30)[0].decode(errors="replace")
        assert time.time() - t0 < 20, out
        assert "no longer sends anybody" in out, out
        assert worker.returncode != 0, out            # Init could not join: the launcher starts the worker again
    finally:
        t.close()
