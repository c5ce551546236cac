"""Rendezvous tracker for rabit-b200 workers.

The reference repository does not contain a tracker: it relies on dmlc-core's
``dmlc_tracker``.  This is an independent implementation of the
*worker-visible* protocol that ``src/host/comm.cc`` (and the reference's
``src/allreduce_base.cc:221-438``) speaks, so that both our workers and an
unmodified reference worker binary can rendezvous here:

* every connection: magic 0xff99 both ways, then rank, world_size, task_id, command
* ``start`` / ``recover``: the tracker answers with rank, tree parent, world size, the
  tree neighbours and the ring neighbours, then brokers the peer connections
* ``print``: one log line; ``shutdown``: the rank is done
* two answers that are not part of the reference protocol (the reference's workers never see them): rank ``-2``
  to a ``start`` / ``recover`` that arrives after some rank has shut down ("the job is over"), and the command
  ``status`` (answer: how many ranks have shut down, and whether the asking worker is still on the list of parked
  workers), which a worker waiting in ``accept()`` uses to find out that the neighbours it waits for have left for
  good or will never be sent its way

Topology: binary heap tree; the ring is a depth-first walk of that tree and ranks are
relabelled so that the ring is simply ``r -> r + 1 (mod n)`` (the ring collectives
require that, reference allreduce_base.cc:787-789).  A worker that comes back with a
``task_id`` the tracker has seen before gets its old rank: that is what makes recovery
after a kill work.
"""
import logging
import socket
import struct
import threading
import time

MAGIC = 0xFF99
log = logging.getLogger("rabit.tracker")


class _Conn(object):
    """One accepted worker connection with the wire helpers of the protocol."""

    def __init__(self, sock, addr):
        self.sock = sock
        self.host = addr[0]
        self.rank = -1
        self.world_size = -1
        self.task_id = "NULL"
        self.cmd = ""
        self.port = -1
        self.wait_accept = 0
        self.pending = set()      # ranks this worker still expects to be dialled by

    def recv_all(self, n):
        chunks = []
        while n > 0:
            b = self.sock.recv(n)
            if not b:
                raise ConnectionError("worker closed the tracker connection")
            chunks.append(b)
            n -= len(b)
        return b"".join(chunks)

    def recv_int(self):
        return struct.unpack("@i", self.recv_all(4))[0]

    def send_int(self, v):
        self.sock.sendall(struct.pack("@i", v))

    def recv_str(self):
        n = self.recv_int()
        return self.recv_all(n).decode() if n > 0 else ""

    def send_str(self, s):
        b = s.encode()
        self.send_int(len(b))
        if b:
            self.sock.sendall(b)

    def handshake(self):
        magic = self.recv_int()
        if magic != MAGIC:
            raise ConnectionError("bad magic %x from %s" % (magic, self.host))
        self.send_int(MAGIC)
        self.rank = self.recv_int()
        self.world_size = self.recv_int()
        self.task_id = self.recv_str()
        self.cmd = self.recv_str()


def build_topology(n):
    """Return (tree_neighbours, parent, ring) for ranks relabelled along the ring.

    ``tree_neighbours[r]`` is the list of tree-adjacent ranks of r, ``parent[r]`` its
    parent (-1 for the root) and ``ring[r] = (prev, next)``.
    """
    if n == 1:
        return {0: []}, {0: -1}, {0: (-1, -1)}
    children = {i: [c for c in (2 * i + 1, 2 * i + 2) if c < n] for i in range(n)}

    def walk(i):
        # depth-first; the last child's walk is reversed so that the walk returns close to
        # where it started and consecutive ring members tend to be tree neighbours
        order = [i]
        kids = children[i]
        for k, c in enumerate(kids):
            sub = walk(c)
            if k == len(kids) - 1:
                sub = sub[::-1]
            order.extend(sub)
        return order

    order = walk(0)
    relabel = {old: new for new, old in enumerate(order)}
    tree, parent, ring = {}, {}, {}
    for old in range(n):
        new = relabel[old]
        nb = [relabel[c] for c in children[old]]
        par = -1
        if old != 0:
            par = relabel[(old + 1) // 2 - 1]
            nb.append(par)
        tree[new] = sorted(nb)
        parent[new] = par
        ring[new] = ((new - 1) % n, (new + 1) % n)
    return tree, parent, ring


class Tracker(object):
    """Accept loop + link broker.  ``start()`` runs it on a daemon thread."""

    def __init__(self, nworker, host="127.0.0.1", port=0, port_end=None, verbose=False):
        self.nworker = nworker
        self.sock = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self.sock.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        if port_end is None:
            self.sock.bind((host, port))
        else:
            for p in range(port, port_end):
                try:
                    self.sock.bind((host, p))
                    break
                except OSError:
                    continue
            else:
                raise OSError("no free tracker port in [%d, %d)" % (port, port_end))
        self.sock.listen(256)
        self.host = host
        self.port = self.sock.getsockname()[1]
        self.verbose = verbose
        self.io_timeout = 120.0     # seconds a worker may stay silent in the middle of an exchange
        self.messages = []          # lines received through "print"
        self.thread = None
        self.finished = threading.Event()   # every rank has said "shutdown"
        self.error = None
        self.start_time = None
        self.end_time = None

    # environment a worker needs to find this tracker (the launcher adds task id / attempt)
    def worker_env(self):
        return {"DMLC_TRACKER_URI": self.host, "DMLC_TRACKER_PORT": str(self.port),
                "DMLC_ROLE": "worker", "DMLC_NUM_WORKER": str(self.nworker)}

    def _assign(self, c, rank, tree, parent, ring, wait_conn):
        """Tell worker c who it is and broker its links (see module docstring)."""
        n = self.nworker
        c.rank = rank
        want = set(tree[rank])
        prv, nxt = ring[rank]
        if prv not in (-1, rank):
            want.add(prv)
        if nxt not in (-1, rank):
            want.add(nxt)
        # one segment for the whole assignment (rank, parent, world, tree neighbours, ring neighbours)
        msg = [rank, parent[rank], n, len(tree[rank])] + list(tree[rank])
        msg += [prv if prv != rank else -1, nxt if nxt != rank else -1]
        c.sock.sendall(struct.pack("@%di" % len(msg), *msg))
        dialled = []
        while True:
            ngood = c.recv_int()
            good = set(c.recv_int() for _ in range(ngood))
            # peers we pointed this worker at that it could not reach are stale (they died while
            # parked): forget them, their replacement will dial this worker instead
            for r in dialled:
                if r not in good:
                    wait_conn.pop(r, None)
            bad = want - good
            dial = [r for r in sorted(bad) if r in wait_conn]
            out = struct.pack("@ii", len(dial), len(bad) - len(dial))
            for r in dial:
                h = wait_conn[r].host.encode()
                out += struct.pack("@i", len(h)) + h + struct.pack("@ii", wait_conn[r].port, r)
            c.sock.sendall(out)
            dialled = dial
            if c.recv_int() != 0:
                continue
            c.port = c.recv_int()
            # book-keeping by SET, not by count: a worker that restarts its rendezvous dials the same parked
            # peer a second time, which must not use up the slot of somebody the peer is still waiting for
            for r in dial:
                wait_conn[r].pending.discard(rank)
                if not wait_conn[r].pending:
                    del wait_conn[r]
            c.pending = set(bad) - set(dial)
            c.wait_accept = len(c.pending)
            if c.pending:
                wait_conn[rank] = c
            return

    def _serve(self):
        n = self.nworker
        tree, parent, ring = build_topology(n)
        job_rank = {}          # task_id -> rank, survives worker restarts
        shutdown = set()
        wait_conn = {}         # rank -> conn still expecting accept()s from later arrivals
        pending = []           # first-time workers waiting for the batch rank assignment
        todo = list(range(n))  # ranks not handed out yet
        while True:
            if len(shutdown) >= n and not self.finished.is_set():
                self.end_time = time.time()
                self.finished.set()
            # After the last shutdown notice the tracker lingers until close(): a worker that was killed between
            # its notice and its exit is restarted by the launcher, and the replacement must get an answer
            # ("the job is over") rather than a listener nobody serves.
            try:
                fd, addr = self.sock.accept()
            except OSError:
                break           # close() was called
            # the protocol is a chain of tiny writes followed by a read: without TCP_NODELAY every such
            # exchange stalls ~40 ms in Nagle + delayed ACK, which dominated recover-after-kill time
            fd.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)
            fd.settimeout(self.io_timeout)   # a frozen worker must not freeze the (single-threaded) tracker
            c = _Conn(fd, addr)
            try:
                c.handshake()
                if c.cmd == "print":
                    msg = c.recv_str()
                    self.messages.append(msg)
                    log.info(msg.rstrip("\n"))
                    if self.verbose:
                        print(msg, end="" if msg.endswith("\n") else "\n", flush=True)
                    continue
                if c.cmd == "status":
                    # (not part of the reference protocol) a worker parked in its accept loop asks whether the peers
                    # it is waiting for can still turn up: once a rank has left for good, they cannot
                    c.send_int(len(shutdown))
                    # ... and whether the tracker still has it on the list of workers that somebody will be sent to
                    # dial (0 while it waits in accept(): both sides lost track of each other, it had better re-join)
                    c.send_int(1 if c.rank in wait_conn else 0)
                    continue
                if c.cmd == "shutdown":
                    if c.rank >= 0:
                        shutdown.add(c.rank)
                        wait_conn.pop(c.rank, None)
                    continue
                if c.cmd not in ("start", "recover"):
                    raise ConnectionError("unknown tracker command %r" % c.cmd)
                if self.start_time is None:
                    self.start_time = time.time()
                if shutdown:
                    # Some ranks have left for good, so a full re-rendezvous can never complete. This only happens
                    # at the very end of a job: the final barrier round reached part of the ranks and failed on the
                    # caller ("recover"), or a worker was killed after it had taken part in that round and its
                    # replacement turns up ("start"). Tell it the job is over instead of parking it for ever
                    # (rank -2; not part of the reference protocol, whose workers would block here).
                    c.sock.sendall(struct.pack("@3i", -2, -1, n))
                    if c.cmd == "start":
                        # the replacement exits on this answer: count the rank it stands for as gone
                        gone = c.rank if c.rank >= 0 else job_rank.get(c.task_id, -1)
                        if gone >= 0:
                            shutdown.add(gone)
                    continue
                rank = c.rank
                if rank < 0 and c.task_id != "NULL":
                    rank = job_rank.get(c.task_id, -1)
                if rank < 0:
                    pending.append(c)
                    if len(pending) == len(todo):
                        # hand out ranks in one go; sorting by host keeps co-located workers adjacent
                        pending.sort(key=lambda x: x.host)
                        batch, pending = pending, []
                        for p in batch:
                            r = todo.pop(0)
                            if p.task_id != "NULL":
                                job_rank[p.task_id] = r
                            try:
                                self._assign(p, r, tree, parent, ring, wait_conn)
                            except (ConnectionError, OSError, struct.error) as e:
                                # died while parked: its replacement finds the rank through its task id,
                                # or (no task id) takes it from the free list again
                                log.warning("tracker: worker for rank %d vanished during assignment: %s", r, e)
                                if p.task_id == "NULL":
                                    todo.append(r)
                            finally:
                                try:
                                    p.sock.close()
                                except OSError:
                                    pass
                    continue
                # a restarted worker ("start") or a survivor re-joining ("recover")
                log.debug("%s rank %d (task %s); parked: %s", c.cmd, rank, c.task_id,
                          {r: sorted(w.pending) for r, w in wait_conn.items()})
                wait_conn.pop(rank, None)
                self._assign(c, rank, tree, parent, ring, wait_conn)
            except (ConnectionError, OSError, struct.error) as e:
                log.warning("tracker: dropped connection from %s: %s", addr[0], e)
            finally:
                if c not in pending:
                    try:
                        c.sock.close()
                    except OSError:
                        pass
        if self.end_time is None:
            self.end_time = time.time()

    def start(self):
        def run():
            try:
                self._serve()
            except Exception as e:  # pylint: disable=broad-except
                self.error = e
                log.error("tracker died: %r", e)
        self.thread = threading.Thread(target=run, name="rabit-tracker", daemon=True)
        self.thread.start()
        return self

    def join(self, timeout=None):
        """Wait until every rank has shut down (or the serving thread died); True if the job is complete."""
        deadline = None if timeout is None else time.time() + timeout
        while not self.finished.is_set() and self.thread.is_alive():
            left = None if deadline is None else deadline - time.time()
            if left is not None and left <= 0:
                break
            self.finished.wait(0.2 if left is None else min(0.2, left))
        return self.finished.is_set()

    def close(self):
        try:
            self.sock.shutdown(socket.SHUT_RDWR)   # wakes the accept() of the lingering serve loop
        except OSError:
            pass
        try:
            self.sock.close()
        except OSError:
            pass


def main():
    """``python -m rabit_b200.parallel.tracker -n 4`` prints the env and serves one job."""
    import argparse
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("-n", "--num-workers", type=int, required=True)
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--port", type=int, default=9091)
    args = ap.parse_args()
    t = Tracker(args.num_workers, args.host, args.port, args.port + 1000, verbose=True)
    for k, v in sorted(t.worker_env().items()):
        print("%s=%s" % (k, v), flush=True)
    t.start()
    t.join()


if __name__ == "__main__":
    main()
