/*!
 * rabit-b200: host transport implementation. See comm.h for the parity map.
 */
#include "comm.h"

#include "rabit/internal/timer.h"

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <cstddef>
#include <map>
#include <thread>

namespace rabit {
namespace engine {

namespace {
/*! \brief idle time after which a blocked primitive probes every link for a silent death */
const int kProbeMs = 100;
/*! \brief longest silence tolerated inside the dial / accept handshake of one connection (both ends are active in it) */
const int kHandshakeMs = 15000;
/*! \brief handshake cookie of the tracker protocol (reference allreduce_base.h kMagic) */
const int kTrackerMagic = 0xff99;

thread_local int g_last_sock_errno = 0;
thread_local char g_last_sock_what[96] = "";

Status ErrnoStatus() {
  g_last_sock_errno = errno;
  switch (errno) {
    case ECONNRESET: case EPIPE: case ECONNABORTED: case ENOTCONN: case ETIMEDOUT:
      return Status::kPeerClosed;
    default:
      return Status::kSockError;
  }
}
inline bool WouldBlock() { return errno == EAGAIN || errno == EWOULDBLOCK || errno == EINTR; }

/*! \brief largest multiple of unit that is <= v (at least one unit) */
inline size_t FloorTo(size_t v, size_t unit) { return std::max(unit, v / unit * unit); }
}  // namespace

int LastSocketErrno() { return g_last_sock_errno; }
const char *LastSocketWhat() { return g_last_sock_what; }

const char *StatusName(Status s) {
  switch (s) {
    case Status::kOk: return "ok";
    case Status::kPeerClosed: return "peer-closed";
    case Status::kSockError: return "socket-error";
    case Status::kTimeout: return "device-timeout";
    case Status::kMismatch: return "consensus-mismatch";
  }
  return "?";
}

HostComm::HostComm() {}
HostComm::~HostComm() {
  this->CloseLinks();
  if (bell_fd_ >= 0) ::close(bell_fd_);
}

void HostComm::CloseLinks(bool graceful) {
  if (graceful) {
    // orderly exit: leave one "bye" byte behind, so that a neighbour that is still inside its last collective
    // can tell this close from a crash (see LinksBroken)
    const char bye = 'B';
    for (Link &l : links_) if (!l.sock.IsClosed()) l.sock.Send(&bye, 1);
  }
  for (Link &l : links_) { l.sock.Close(); DropShm(&l); }
  links_.clear();
  tree_.clear();
  parent_slot_ = -1;
  ring_prev_ = ring_next_ = nullptr;
}

bool HostComm::LinksBroken() {
  for (Link &l : links_) {
    if (l.sock.IsClosed()) return true;
    // peek: 0 = the peer closed without saying bye (it died, or it tore its links down to recover);
    // pending bytes (payload of a running op, or the bye of a peer that finished) = fine
    char c;
    ssize_t n = recv(l.sock.fd, &c, 1, MSG_PEEK | MSG_DONTWAIT);
    if (n == 0) return true;
    if (n < 0 && !WouldBlock()) return true;
  }
  return false;
}

bool HostComm::WaitLinks(WaitSet *ws) {
  if (ws->left >= 0) {
    snprintf(g_last_sock_what, sizeof(g_last_sock_what), "rank %d has closed its end", ws->left);
    return false;
  }
  auto ring_ready = [](const WaitSet::Item &it) {
    return it.recv ? it.l->rx.Avail() >= it.need : it.l->tx.Space() >= it.l->tx.SendThreshold();
  };
  if (!ws->ready_now && !ws->shm.empty()) {
    // a short look before going to sleep: the other end usually is a few microseconds away
    if (spin_us_ > 0) {
      const double until = utils::GetTime() + spin_us_ * 1e-6;
      do {
        for (const WaitSet::Item &it : ws->shm) if (ring_ready(it)) { ws->ready_now = true; break; }
      } while (!ws->ready_now && utils::GetTime() < until);
    }
    if (!ws->ready_now) {
      // leave the note "wake me", then look once more (see ShmRing::SetConsumerWaiting)
      for (const WaitSet::Item &it : ws->shm) {
        if (it.recv) it.l->rx.SetConsumerWaiting(1); else it.l->tx.SetProducerWaiting(1);
      }
      for (const WaitSet::Item &it : ws->shm) if (ring_ready(it)) { ws->ready_now = true; break; }
    }
  }
  // bytes are waiting in shared memory and no socket needs asking: nothing to ask the kernel (but look at the
  // sockets every so often all the same: a peer's death must not go unnoticed however busy the rings are)
  if (ws->ready_now && !ws->tcp && ++fast_turns_ < 4096) return true;
  fast_turns_ = 0;
  utils::PollHelper *poll = &ws->poll;
  if (!ws->shm.empty() && bell_fd_ >= 0) poll->WatchRead(bell_fd_);
  for (Link &l : links_) if (!l.hup_seen && !l.sock.IsClosed()) poll->WatchHangup(l.sock.fd);
  // nothing moved for a while: a peer can also vanish without a FIN reaching us (host down), so look at every link
  const int rc = poll->Poll(ws->ready_now ? 0 : kProbeMs);
  if (!ws->shm.empty() && bell_fd_ >= 0 && poll->CheckRead(bell_fd_)) {
    char drain[64];
    while (recv(bell_fd_, drain, sizeof(drain), MSG_DONTWAIT) > 0) {}
  }
  if (rc == 0) return ws->ready_now || !LinksBroken();
  for (Link &l : links_) {
    if (l.hup_seen || l.sock.IsClosed() || !poll->CheckHangup(l.sock.fd)) continue;
    char c;
    const ssize_t n = recv(l.sock.fd, &c, 1, MSG_PEEK | MSG_DONTWAIT);
    if (n == 0 || (n < 0 && !WouldBlock())) {
      snprintf(g_last_sock_what, sizeof(g_last_sock_what), "hang-up of %d", l.rank);
      return false;
    }
    // closed behind unread bytes (payload of this op, or the bye of a peer that is done): whoever reads the
    // link finds out; do not wake up for it again
    l.hup_seen = true;
  }
  return true;
}

// ---------------------------------------------------------------- link I/O
/*! \brief one non-blocking receive; *got == 0 with kOk means "nothing yet" */
Status HostComm::RecvSome(Link *l, char *p, size_t max, size_t *got) {
  *got = 0;
  if (max == 0) return Status::kOk;
  if (l->shm) {
    *got = l->rx.Read(p, max);
    // a producer that went to sleep on a full ring is woken once a worthwhile piece is free again
    if (*got != 0 && l->rx.Space() >= l->rx.SendThreshold() && l->rx.TakeProducerWaiting()) RingBell(l);
    return Status::kOk;
  }
  ssize_t n = l->sock.Recv(p, max);
  if (n > 0) { *got = static_cast<size_t>(n); return Status::kOk; }
  if (n == 0) { snprintf(g_last_sock_what, sizeof(g_last_sock_what), "recv=0 from %d", l->rank); return Status::kPeerClosed; }
  if (WouldBlock()) return Status::kOk;
  snprintf(g_last_sock_what, sizeof(g_last_sock_what), "recv(%p,%lu) from %d", static_cast<void *>(p),
           static_cast<unsigned long>(max), l->rank);  // NOLINT(*)
  return ErrnoStatus();
}

/*! \brief one non-blocking send */
Status HostComm::SendSome(Link *l, const char *p, size_t max, size_t *put) {
  *put = 0;
  if (max == 0) return Status::kOk;
  if (l->shm) {
    *put = l->tx.Write(p, max);
    if (*put != 0 && l->tx.TakeConsumerWaiting()) RingBell(l);
    return Status::kOk;
  }
  ssize_t n = l->sock.Send(p, max);
  if (n >= 0) { *put = static_cast<size_t>(n); return Status::kOk; }
  if (WouldBlock()) return Status::kOk;
  snprintf(g_last_sock_what, sizeof(g_last_sock_what), "send(%p,%lu) to %d", static_cast<const void *>(p),
           static_cast<unsigned long>(max), l->rank);  // NOLINT(*)
  return ErrnoStatus();
}

void HostComm::RingBell(Link *l) {
  // (a full queue means wake-ups are already pending; a vanished peer is the hang-up watch's business)
  if (bell_fd_ >= 0 && l->bell_len != 0) {
    const char c = 1;
    sendto(bell_fd_, &c, 1, MSG_DONTWAIT | MSG_NOSIGNAL, reinterpret_cast<const sockaddr *>(&l->bell), l->bell_len);
  }
}

void HostComm::ReplaceLink(Link *l, const Link &fresh) {
  l->sock.Close();
  DropShm(l);
  *l = fresh;      // socket, rings and doorbell of the new connection (plain handles: the copy owns them now)
}

void HostComm::DropShm(Link *l) {
  l->tx.Close();
  l->rx.Close();
  l->shm = false;
  l->bell_len = 0;
}

void HostComm::WantRecv(WaitSet *ws, Link *l, size_t need) {
  if (!l->shm) { ws->poll.WatchRead(l->sock.fd); ws->tcp = true; return; }
  ws->shm.push_back(WaitSet::Item{l, true, need});
  if (l->rx.Avail() >= need) {
    ws->ready_now = true;
  } else if (l->hup_seen) {
    // The peer closed its connection (it finished and said bye, or it is re-joining) after writing all it was going
    // to write. With payload on the socket the reader would hit the end of the stream; a ring just stays empty.
    ws->left = l->rank;
  }
}

void HostComm::WantSend(WaitSet *ws, Link *l) {
  if (!l->shm) { ws->poll.WatchWrite(l->sock.fd); ws->tcp = true; return; }
  ws->shm.push_back(WaitSet::Item{l, false, 0});
  if (l->hup_seen) { ws->left = l->rank; return; }   // nobody will read it (a socket would answer with EPIPE)
  if (l->tx.Space() >= l->tx.SendThreshold()) ws->ready_now = true;
}

bool HostComm::CanRecv(const WaitSet &ws, Link *l) const {
  return l->shm ? l->rx.Avail() != 0 : ws.poll.CheckRead(l->sock.fd);
}

bool HostComm::CanSend(const WaitSet &ws, Link *l) const {
  return l->shm ? l->tx.Space() >= l->tx.SendThreshold() : ws.poll.CheckWrite(l->sock.fd);
}

bool HostComm::AllLinksSharedMemory() const {
  for (const Link &l : links_) if (!l.shm) return false;
  return !links_.empty();
}

Link *HostComm::FindLink(int r) {
  for (Link &l : links_) if (l.rank == r) return &l;
  return nullptr;
}

// ------------------------------------------------------------------ tracker
utils::TCPSocket HostComm::DialTracker() {
  utils::TCPSocket t;
  utils::SockAddr addr(cfg.tracker_uri.c_str(), cfg.tracker_port);
  for (int attempt = 1;; ++attempt) {
    t.Create();
    t.SetNoDelay(true);  // the handshake is many tiny writes followed by reads: avoid Nagle stalls
    if (t.Connect(addr)) break;
    t.Close();
    utils::Check(attempt < std::max(cfg.connect_retry, 1), "cannot reach tracker %s:%d after %d attempts",
                 cfg.tracker_uri.c_str(), cfg.tracker_port, attempt);
    fprintf(stderr, "[rabit] tracker %s:%d not reachable, retry %d\n", cfg.tracker_uri.c_str(),
            cfg.tracker_port, attempt);
    usleep(200000 * attempt);
  }
  int magic = kTrackerMagic;
  utils::Check(t.SendInt(magic) && t.RecvInt(&magic) && magic == kTrackerMagic,
               "tracker handshake failed (bad magic)");
  const int my_rank = rank_assigned_ ? rank : -1;
  utils::Check(t.SendInt(my_rank) && t.SendInt(world_known_ ? world : cfg.world_size_hint) &&
               t.SendStr(cfg.task_id), "tracker handshake failed (identity)");
  return t;
}

bool HostComm::Connect(const char *cmd) {
  if (!distributed()) { rank = 0; world = 1; return true; }
  try {
    utils::TCPSocket t = this->DialTracker();
    utils::Check(t.SendStr(cmd), "tracker: cannot send command");
    int new_rank = -1, nnbr = 0;
    utils::Check(t.RecvInt(&new_rank) && t.RecvInt(&parent_rank) && t.RecvInt(&world),
                 "tracker: cannot read assignment");
    if (new_rank == -2) {
      // our tracker's way of saying: peers have shut down for good, there is nothing to re-join
      job_over = true;
      utils::Error("the job is over: other ranks have already shut down");
    }
    utils::Check(new_rank >= 0, "tracker refused to assign a rank");
    utils::Assert(!rank_assigned_ || new_rank == rank, "tracker changed our rank from %d to %d", rank, new_rank);
    rank = new_rank; rank_assigned_ = true; world_known_ = true;
    // looking at the rings before sleeping only pays while every worker has a core to itself: with more workers than
    // cores (measured: 31 workers on 8 cores, 0.14-0.19 s without vs 0.19-0.40 s with) the one we wait for may be the
    // one we keep off the CPU
    const int cores = static_cast<int>(std::thread::hardware_concurrency());
    spin_us_ = (cfg.shm_spin_set || cores <= 0 || world <= cores) ? cfg.shm_spin_us : 0;
    std::vector<int> tree_nbrs;
    utils::Check(t.RecvInt(&nnbr), "tracker: cannot read neighbour count");
    for (int i = 0; i < nnbr; ++i) {
      int r; utils::Check(t.RecvInt(&r), "tracker: cannot read neighbour");
      tree_nbrs.push_back(r);
    }
    utils::Check(t.RecvInt(&ring_prev_rank) && t.RecvInt(&ring_next_rank), "tracker: cannot read ring");

    this->OpenDoorbell();
    utils::TCPSocket listener;
    struct CloseOnExit { utils::TCPSocket *s; ~CloseOnExit() { s->Close(); } } close_listener{&listener};
    listener.Create();
    int port;
    if (cfg.slave_port > 0) {
      port = listener.TryBindHost(cfg.slave_port, cfg.slave_port + cfg.nport_trial);
      utils::Check(port != -1, "no free port in [%d, %d)", cfg.slave_port, cfg.slave_port + cfg.nport_trial);
    } else {
      port = listener.BindAny();
    }
    listener.Listen();

    int nconn = 0, naccept = 0, nerr = 0;
    std::vector<int> dialled;
    do {
      // report which links survived, drop the rest
      std::vector<int> good;
      for (Link &l : links_) {
        if (l.sock.BadSocket()) l.sock.Close(); else good.push_back(l.rank);
      }
      utils::Check(t.SendInt(static_cast<int>(good.size())), "tracker: send failed");
      for (int r : good) utils::Check(t.SendInt(r), "tracker: send failed");
      utils::Check(t.RecvInt(&nconn) && t.RecvInt(&naccept), "tracker: cannot read link plan");
      nerr = 0;
      for (int i = 0; i < nconn; ++i) {
        std::string peer_host; int peer_port, peer_rank;
        utils::Check(t.RecvStr(&peer_host) && t.RecvInt(&peer_port) && t.RecvInt(&peer_rank),
                     "tracker: cannot read peer address");
        Link nl;
        nl.sock.Create();
        if (!nl.sock.Connect(utils::SockAddr(peer_host.c_str(), peer_port))) {
          nl.sock.Close(); ++nerr; continue;
        }
        // handshake: who I am, whom I mean to reach (a stale address in the tracker's table can belong to a
        // different worker by now: ephemeral ports are recycled quickly), and who answered
        int echoed = -1;
        nl.rank = peer_rank;
        if (!nl.sock.SendInt(rank) || !nl.sock.SendInt(peer_rank) || !nl.sock.RecvIntTimed(&echoed, kHandshakeMs) ||
            echoed != peer_rank ||
            !NegotiateSharedMemory(&nl, true)) {
          nl.sock.Close(); ++nerr; continue;
        }
        dialled.push_back(peer_rank);
        Link *old = FindLink(peer_rank);
        if (old != nullptr) ReplaceLink(old, nl); else links_.push_back(nl);
      }
      utils::Check(t.SendInt(nerr), "tracker: send failed");
    } while (nerr != 0);
    utils::Check(t.SendInt(port), "tracker: cannot send port");
    t.Close();

    // The tracker only says HOW MANY peers will dial us. A peer whose own rendezvous attempt failed half-way
    // dials again on its retry, so count distinct ranks: a repeated rank replaces its older socket.
    // Only neighbours we did not dial ourselves may connect, and only if they really meant us.
    std::vector<int> expected(tree_nbrs);
    if (ring_prev_rank >= 0) expected.push_back(ring_prev_rank);
    if (ring_next_rank >= 0) expected.push_back(ring_next_rank);
    std::vector<int> accepted;
    // A peer that died between the tracker's plan and its dial would leave us in accept() for good (the
    // reference blocks there, allreduce_base.cc:420-436); give up after a while so that the caller goes back
    // to the tracker, where the replacement process will turn up.
    const double give_up_at = utils::GetTime() + (cfg.timeout_enabled ? std::max(cfg.timeout_sec, 1) : 300);
    int not_listed = 0;
    while (static_cast<int>(accepted.size()) < naccept) {
      utils::PollHelper incoming;
      incoming.WatchRead(listener.fd);
      if (incoming.Poll(1000) == 0) {
        utils::Check(utils::GetTime() < give_up_at, "gave up waiting for %d more neighbour(s) to connect",
                     naccept - static_cast<int>(accepted.size()));
        // The peers we wait for may have finished meanwhile: the job's last barrier can complete on some ranks
        // and break on others, and a rank that passed it leaves without looking back. They will never dial.
        bool parked = true;
        if (this->JobIsOver(&parked)) {
          job_over = true;
          utils::Error("the job is over: other ranks have already shut down");
        }
        // The tracker sends diallers to the workers on its list. If this worker is not on it (twice in a row: the
        // last dialler may just be in its handshake with us) nobody will ever come: start over rather than sit
        // out the whole timeout. (It forgets a parked worker when somebody could not reach it, taking it for dead.)
        not_listed = parked ? 0 : not_listed + 1;
        utils::Check(not_listed < 2, "the tracker no longer sends anybody to dial this worker; starting the rendezvous over");
        continue;
      }
      Link nl;
      nl.sock = listener.Accept();
      int intended = -1;
      if (!nl.sock.SendInt(rank) || !nl.sock.RecvIntTimed(&nl.rank, kHandshakeMs) ||
          !nl.sock.RecvIntTimed(&intended, kHandshakeMs) || intended != rank ||
          std::find(expected.begin(), expected.end(), nl.rank) == expected.end() || !NegotiateSharedMemory(&nl, false)) {
        nl.sock.Close();  // dialler died, or was sent here by mistake
        continue;
      }
      // A neighbour that restarted its own rendezvous dials again: its new socket replaces the old one below,
      // but it does not count towards the peers we are still waiting for.
      const bool known = std::find(dialled.begin(), dialled.end(), nl.rank) != dialled.end() ||
                         std::find(accepted.begin(), accepted.end(), nl.rank) != accepted.end();
      if (!known) accepted.push_back(nl.rank);
      Link *old = FindLink(nl.rank);
      if (old != nullptr) ReplaceLink(old, nl); else links_.push_back(nl);
    }
    listener.Close();

    // classify
    tree_.clear(); parent_slot_ = -1; ring_prev_ = ring_next_ = nullptr;
    for (Link &l : links_) {
      utils::Assert(!l.sock.BadSocket(), "link to rank %d is not usable after rendezvous", l.rank);
      l.sock.SetNonBlock(true);
      l.sock.SetKeepAlive(true);
      if (cfg.tcp_no_delay) l.sock.SetNoDelay(true);
      if (std::find(tree_nbrs.begin(), tree_nbrs.end(), l.rank) != tree_nbrs.end()) {
        if (l.rank == parent_rank) parent_slot_ = static_cast<int>(tree_.size());
        tree_.push_back(&l);
      }
      if (l.rank == ring_prev_rank) ring_prev_ = &l;
      if (l.rank == ring_next_rank) ring_next_ = &l;
    }
    utils::Assert(parent_rank == -1 || parent_slot_ != -1, "no link to tree parent %d", parent_rank);
    utils::Assert(ring_prev_rank == -1 || ring_prev_ != nullptr, "no link to ring prev %d", ring_prev_rank);
    utils::Assert(ring_next_rank == -1 || ring_next_ != nullptr, "no link to ring next %d", ring_next_rank);
    ++link_generation;
    return true;
  } catch (const std::exception &e) {
    fprintf(stderr, "[rabit] rendezvous (%s) failed: %s\n", cmd, e.what());
    // leave nothing half-built behind: the caller retries from scratch
    for (Link &l : links_) { l.sock.Close(); DropShm(&l); }
    links_.clear();
    tree_.clear(); parent_slot_ = -1; ring_prev_ = ring_next_ = nullptr;
    return false;
  }
}

// ------------------------------------------------------ same-host data plane
namespace {
/*! \brief what the two ends of a fresh connection tell each other right after the rendezvous */
struct ShmHello {
  uint32_t magic;
  uint32_t want;        // this end is able and willing to use shared memory
  uint64_t token[2];    // HostToken(): equal on both ends = same host, same namespaces, same user
  int32_t pid;
  int32_t ring_fd;      // the sender's ring (this end -> peer) is file descriptor ring_fd of process pid
  uint64_t capacity;
  char bell[48];        // abstract unix datagram address of this end's doorbell (without the leading NUL)
};
const uint32_t kHelloMagic = 0x52424c31;  // "RBL1"
}  // namespace

void HostComm::OpenDoorbell() {
  if (bell_fd_ >= 0 || cfg.shm_ring < (64 << 10)) return;
  static std::atomic<int> serial{0};
  bell_fd_ = socket(AF_UNIX, SOCK_DGRAM | SOCK_NONBLOCK | SOCK_CLOEXEC, 0);
  bell_name_ = "rabit-bell-" + std::to_string(getpid()) + "-" + std::to_string(serial.fetch_add(1));
  sockaddr_un a; std::memset(&a, 0, sizeof(a));
  a.sun_family = AF_UNIX;
  std::memcpy(a.sun_path + 1, bell_name_.data(), bell_name_.size());
  if (bell_fd_ < 0 || bind(bell_fd_, reinterpret_cast<sockaddr *>(&a),
                           static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + bell_name_.size())) != 0) {
    if (bell_fd_ >= 0) ::close(bell_fd_);
    bell_fd_ = -1;
  }
}

/*!
 * \brief second half of the dial / accept handshake of one connection: both ends say whether they can share memory
 *  (same host, namespaces and user), show each other their ring, and confirm. It runs while BOTH ends are busy with
 *  exactly this connection (the dialler right after connect(), the acceptor right after accept()), so it never waits
 *  for a third worker. false = the connection broke.
 */
bool HostComm::NegotiateSharedMemory(Link *l, bool dialler) {
  const size_t capacity = cfg.shm_ring / 4096 * 4096;
  ShmHello mine; std::memset(&mine, 0, sizeof(mine));
  mine.magic = kHelloMagic;
  HostToken(mine.token);
  if (cfg.shm_test_split > 1) mine.token[1] ^= static_cast<uint64_t>(rank % cfg.shm_test_split) + 1;  // tests: a mix of media
  mine.pid = static_cast<int32_t>(getpid());
  mine.capacity = capacity;
  snprintf(mine.bell, sizeof(mine.bell), "%s", bell_name_.c_str());
  mine.want = bell_fd_ >= 0 && capacity >= (64 << 10) && l->tx.Create(capacity) ? 1u : 0u;
  mine.ring_fd = l->tx.fd();
  if (cfg.tcp_no_delay) l->sock.SetNoDelay(true);
  ShmHello peer;
  auto send_hello = [&]() { return l->sock.SendAll(&mine, sizeof(mine)) == sizeof(mine); };
  auto recv_hello = [&]() { return l->sock.RecvAllTimed(&peer, sizeof(peer), kHandshakeMs) == sizeof(peer) && peer.magic == kHelloMagic; };
  if (dialler ? !(send_hello() && recv_hello()) : !(recv_hello() && send_hello())) { DropShm(l); return false; }
  peer.bell[sizeof(peer.bell) - 1] = '\0';
  const size_t blen = strlen(peer.bell);
  char ok = 0, peer_ok = 0;
  if (mine.want != 0 && peer.want != 0 && peer.token[0] == mine.token[0] && peer.token[1] == mine.token[1] &&
      peer.capacity == capacity && blen != 0 && l->rx.Open(peer.pid, peer.ring_fd, capacity)) {
    std::memset(&l->bell, 0, sizeof(l->bell));
    l->bell.sun_family = AF_UNIX;
    std::memcpy(l->bell.sun_path + 1, peer.bell, blen);
    l->bell_len = static_cast<socklen_t>(offsetof(sockaddr_un, sun_path) + 1 + blen);
    ok = 1;
  }
  auto send_ok = [&]() { return l->sock.SendAll(&ok, 1) == 1; };
  auto recv_ok = [&]() { return l->sock.RecvAllTimed(&peer_ok, 1, kHandshakeMs) == 1; };
  if (dialler ? !(send_ok() && recv_ok()) : !(recv_ok() && send_ok())) { DropShm(l); return false; }
  if (ok != 0 && peer_ok != 0) {
    l->shm = true;
  } else {
    DropShm(l);
  }
  return true;
}

bool HostComm::JobIsOver(bool *parked) {
  if (parked != nullptr) *parked = true;
  // one quiet attempt: the tracker may be busy (it serves one rendezvous at a time, and that worker may be dialling
  // US right now: do not keep it waiting) or gone, and then the answer simply is "not known to be over"
  utils::TCPSocket t;
  t.Create();
  t.SetNoDelay(true);
  bool over = false;
  int magic = kTrackerMagic, n = 0;
  if (t.Connect(utils::SockAddr(cfg.tracker_uri.c_str(), cfg.tracker_port)) && t.SendInt(magic) &&
      t.RecvIntTimed(&magic, 500) && magic == kTrackerMagic && t.SendInt(rank_assigned_ ? rank : -1) &&
      t.SendInt(world_known_ ? world : cfg.world_size_hint) && t.SendStr(cfg.task_id) && t.SendStr("status") &&
      t.RecvIntTimed(&n, 500)) {
    over = n > 0;
    int listed = 1;
    if (parked != nullptr && t.RecvIntTimed(&listed, 500)) *parked = listed != 0;
  }
  t.Close();
  return over;
}

void HostComm::SendShutdown() {
  if (!distributed() || !rank_assigned_) return;
  try {
    utils::TCPSocket t = this->DialTracker();
    t.SendStr("shutdown");
    t.Close();
  } catch (const std::exception &) {
    // the tracker may already be gone when the job is torn down from outside
  }
}

void HostComm::TrackerPrint(const std::string &msg) {
  if (!distributed()) { utils::Printf("%s", msg.c_str()); fflush(stdout); return; }
  utils::TCPSocket t = this->DialTracker();
  t.SendStr("print");
  t.SendStr(msg);
  t.Close();
}

// -------------------------------------------------------------- collectives
Status HostComm::Allreduce(void *buf, size_t elem_bytes, size_t count, ReduceFunction *reducer) {
  if (world <= 1 || count == 0) return Status::kOk;
  if (UseRing(elem_bytes, count)) return AllreduceRing(buf, elem_bytes, count, reducer);
  return AllreduceTree(buf, elem_bytes, count, reducer);
}

Status HostComm::AllreduceTree(void *buf_, size_t eb, size_t count, ReduceFunction *reducer) {
  if (tree_.empty() || count == 0) return Status::kOk;
  char *buf = static_cast<char *>(buf_);
  const size_t total = eb * count;
  const bool is_root = parent_slot_ < 0;
  Link *parent = is_root ? nullptr : tree_[parent_slot_];
  struct Child { Link *l; size_t recvd; size_t sent_down; char *stage; };
  std::vector<Child> kids;
  for (int i = 0; i < static_cast<int>(tree_.size()); ++i) {
    if (i != parent_slot_) kids.push_back(Child{tree_[i], 0, 0, nullptr});
  }
  // per-child staging ring; always a whole number of items so an item never straddles the wrap
  const size_t cap = FloorTo(std::min(total, cfg.reduce_buffer), eb);
  if (scratch_.size() < cap * kids.size()) scratch_.resize(cap * kids.size());
  for (size_t i = 0; i < kids.size(); ++i) kids[i].stage = scratch_.data() + i * cap;

  const size_t fold_min = std::min(std::max(cfg.tree_reduce_minsize / eb * eb, eb), std::max(cap / 2 / eb * eb, eb));
  size_t reduced = kids.empty() ? total : 0;  // prefix of buf that holds the subtree's reduction
  size_t sent_up = 0, recv_down = 0;
  const MPI::Datatype dtype(eb);
  WaitSet ws;
  while (true) {
    if (is_root) { sent_up = reduced; recv_down = reduced; }
    bool done = recv_down == total;
    for (const Child &c : kids) done = done && c.sent_down == total;
    if (done) break;
    ws.Clear();
    for (const Child &c : kids) {
      if (c.recvd < total && c.recvd - reduced < cap) WantRecv(&ws, c.l);
      if (c.sent_down < recv_down) WantSend(&ws, c.l);
    }
    if (parent != nullptr) {
      if (sent_up < reduced) WantSend(&ws, parent);
      if (recv_down < sent_up) WantRecv(&ws, parent);
    }
    // a neighbour this primitive does not talk to may have died: nothing would ever wake us up, so probe
    // all links whenever nothing moved for a while
    if (!WaitLinks(&ws)) return Status::kPeerClosed;
    // up: children -> staging
    for (Child &c : kids) {
      if (c.recvd < total && c.recvd - reduced < cap && CanRecv(ws, c.l)) {
        const size_t off = c.recvd % cap;
        size_t want = std::min(total - c.recvd, cap - (c.recvd - reduced));
        want = std::min(want, cap - off);
        size_t got; Status s = RecvSome(c.l, c.stage + off, want, &got);
        if (!Ok(s)) return s;
        c.recvd += got;
      }
    }
    // fold whatever every child has delivered
    if (!kids.empty()) {
      size_t avail = total;
      for (const Child &c : kids) avail = std::min(avail, c.recvd);
      avail = avail / eb * eb;
      // rabit_tree_reduce_minsize: do not call the reducer (and wake the parent) for every few bytes that
      // trickle in; wait for a worthwhile piece unless the message ends first (reference allreduce_base.cc:496)
      if (avail != total && avail - reduced < fold_min) avail = reduced;
      while (reduced < avail) {
        const size_t off = reduced % cap;
        const size_t n = std::min(avail - reduced, cap - off);
        for (Child &c : kids) reducer(c.stage + off, buf + reduced, static_cast<int>(n / eb), dtype);
        reduced += n;
      }
    }
    if (parent != nullptr) {
      if (sent_up < reduced && CanSend(ws, parent)) {
        size_t put; Status s = SendSome(parent, buf + sent_up, reduced - sent_up, &put);
        if (!Ok(s)) return s;
        sent_up += put;
      }
      if (recv_down < sent_up && CanRecv(ws, parent)) {
        size_t got; Status s = RecvSome(parent, buf + recv_down, sent_up - recv_down, &got);
        if (!Ok(s)) return s;
        recv_down += got;
      }
    } else {
      recv_down = reduced;
    }
    // down: final values -> children
    for (Child &c : kids) {
      if (c.sent_down < recv_down && CanSend(ws, c.l)) {
        size_t put; Status s = SendSome(c.l, buf + c.sent_down, recv_down - c.sent_down, &put);
        if (!Ok(s)) return s;
        c.sent_down += put;
      }
    }
  }
  return Status::kOk;
}

Status HostComm::Broadcast(void *buf_, size_t size, int root) {
  if (tree_.empty() || size == 0) return Status::kOk;
  char *buf = static_cast<char *>(buf_);
  const int nl = static_cast<int>(tree_.size());
  // flood over the undirected tree: whoever hands us the first byte is our upstream
  int upstream = (rank == root) ? -2 : -1;
  size_t recvd = (rank == root) ? size : 0;
  std::vector<size_t> sent(nl, 0);
  WaitSet ws;
  while (true) {
    bool done = recvd == size;
    for (int i = 0; i < nl; ++i) if (i != upstream) done = done && sent[i] == size;
    if (done) break;
    ws.Clear();
    for (int i = 0; i < nl; ++i) {
      if (upstream == -1 || (i == upstream && recvd < size)) WantRecv(&ws, tree_[i]);
      if (upstream != -1 && i != upstream && sent[i] < recvd) WantSend(&ws, tree_[i]);
    }
    // a neighbour this primitive does not talk to may have died: nothing would ever wake us up, so probe
    // all links whenever nothing moved for a while
    if (!WaitLinks(&ws)) return Status::kPeerClosed;
    if (upstream == -1) {
      for (int i = 0; i < nl && upstream == -1; ++i) {
        if (!CanRecv(ws, tree_[i])) continue;
        size_t got; Status s = RecvSome(tree_[i], buf, size, &got);
        if (!Ok(s)) return s;
        if (got != 0) { upstream = i; recvd = got; }
      }
      continue;  // re-arm the poll set now that directions are known
    }
    if (upstream >= 0 && recvd < size && CanRecv(ws, tree_[upstream])) {
      size_t got; Status s = RecvSome(tree_[upstream], buf + recvd, size - recvd, &got);
      if (!Ok(s)) return s;
      recvd += got;
    }
    for (int i = 0; i < nl; ++i) {
      if (i == upstream || sent[i] >= recvd || !CanSend(ws, tree_[i])) continue;
      size_t put; Status s = SendSome(tree_[i], buf + sent[i], recvd - sent[i], &put);
      if (!Ok(s)) return s;
      sent[i] += put;
    }
  }
  return Status::kOk;
}

Status HostComm::AllgatherRing(void *buf_, size_t total, size_t slice_begin, size_t slice_end,
                               size_t size_prev_slice) {
  if (ring_prev_ == nullptr || ring_next_ == nullptr || total == 0) return Status::kOk;
  utils::Assert(slice_begin <= slice_end && slice_end <= total, "Allgather: slice [%lu,%lu) outside %lu",
                (unsigned long)slice_begin, (unsigned long)slice_end, (unsigned long)total);  // NOLINT(*)
  char *buf = static_cast<char *>(buf_);
  // Bytes stream "downwards" in rank: in from the successor, out to the predecessor. Seen
  // as a ring of `total` bytes both streams are contiguous and ascending: what arrives
  // continues right after our slice, what leaves starts at our slice.
  const size_t mine = slice_end - slice_begin;
  const size_t total_recv = total - mine;
  const size_t total_send = total - size_prev_slice;
  size_t recvd = 0, sent = 0;
  WaitSet ws;
  while (recvd < total_recv || sent < total_send) {
    const size_t sendable = std::min(total_send, mine + recvd);
    ws.Clear();
    if (recvd < total_recv) WantRecv(&ws, ring_next_);
    if (sent < sendable) WantSend(&ws, ring_prev_);
    // a neighbour this primitive does not talk to may have died: nothing would ever wake us up, so probe
    // all links whenever nothing moved for a while
    if (!WaitLinks(&ws)) return Status::kPeerClosed;
    if (recvd < total_recv && CanRecv(ws, ring_next_)) {
      const size_t pos = (slice_end + recvd) % total;
      size_t got; Status s = RecvSome(ring_next_, buf + pos, std::min(total_recv - recvd, total - pos), &got);
      if (!Ok(s)) return s;
      recvd += got;
    }
    if (sent < sendable && CanSend(ws, ring_prev_)) {
      const size_t pos = (slice_begin + sent) % total;
      size_t put; Status s = SendSome(ring_prev_, buf + pos, std::min(sendable - sent, total - pos), &put);
      if (!Ok(s)) return s;
      sent += put;
    }
  }
  return Status::kOk;
}

Status HostComm::ReduceScatterRing(void *buf_, size_t eb, size_t count, ReduceFunction *reducer, const void *src_) {
  if (ring_prev_ == nullptr || ring_next_ == nullptr || count == 0) return Status::kOk;
  char *buf = static_cast<char *>(buf_);
  const char *src = src_ != nullptr && src_ != buf_ ? static_cast<const char *>(src_) : nullptr;
  const size_t n = static_cast<size_t>(world);
  const size_t step = (count + n - 1) / n;
  const size_t total = count * eb;
  auto seg_begin = [&](size_t j) { return std::min((j % n) * step, count) * eb; };
  auto seg_len = [&](size_t j) { return std::min((j % n + 1) * step, count) * eb - seg_begin(j); };
  // Partial sums travel towards lower ranks. We emit segments r+1, r+2, ..., r+n-1 and absorb
  // r+2, ..., r+n (= r, the one we end up owning); as byte ranges of the ring both are contiguous.
  const size_t r = static_cast<size_t>(rank);
  const size_t send_start = seg_begin(r + 1);
  const size_t recv_start = seg_begin(r + 2);
  const size_t total_send = total - seg_len(r);
  const size_t total_recv = total - seg_len(r + 1);
  const size_t head = seg_len(r + 1);  // our own data: can leave before anything is folded
  const size_t cap = FloorTo(std::min(total, cfg.reduce_buffer), eb);
  if (scratch_.size() < cap) scratch_.resize(cap);
  char *stage = scratch_.data();
  size_t recvd = 0, reduced = 0, sent = 0;
  const MPI::Datatype dtype(eb);
  // between workers of one host the partial sums are folded where they lie, in the shared-memory ring
  const bool fold_in_ring = ring_next_->shm && eb <= ring_next_->rx.capacity() / 4;
  WaitSet ws;
  while (sent < total_send || reduced < total_recv) {
    size_t sendable = std::min(total_send, head + reduced);
    ws.Clear();
    const bool can_read = recvd < total_recv && recvd - reduced < cap;
    // (the fold out of a shared-memory ring works on whole items: do not wake up for a part of one)
    if (can_read) WantRecv(&ws, ring_next_, fold_in_ring ? eb : 1);
    if (sent < sendable) WantSend(&ws, ring_prev_);
    // a neighbour this primitive does not talk to may have died: nothing would ever wake us up, so probe
    // all links whenever nothing moved for a while
    if (!WaitLinks(&ws)) return Status::kPeerClosed;
    if (can_read && fold_in_ring) {
      // fold straight out of the shared-memory ring: no staging copy. One contiguous span per turn, so that the
      // partial sums it produces leave for the predecessor before the next span is folded.
      ShmRing &rx = ring_next_->rx;
      const char *span;
      size_t len = std::min(rx.Peek(&span), total_recv - reduced);
      const size_t pos = (recv_start + reduced) % total;
      len = std::min(len, total - pos) / eb * eb;
      bool copied = false;
      if (len == 0 && std::min(rx.Avail(), total_recv - reduced) >= eb) {
        // an item that straddles the wrap of the ring goes through a copy
        if (item_.size() < eb) item_.resize(eb);
        rx.Read(item_.data(), eb);
        span = item_.data();
        len = eb;
        copied = true;
      }
      if (len != 0) {
        if (src != nullptr) std::memcpy(buf + pos, src + pos, len);
        reducer(span, buf + pos, static_cast<int>(len / eb), dtype);
        if (!copied) rx.Consume(len);
        if (rx.Space() >= rx.SendThreshold() && rx.TakeProducerWaiting()) RingBell(ring_next_);
        reduced += len;
      }
      recvd = reduced;
    } else if (can_read && CanRecv(ws, ring_next_)) {
      const size_t off = recvd % cap;
      size_t want = std::min(total_recv - recvd, cap - (recvd - reduced));
      want = std::min(want, cap - off);
      size_t got; Status s = RecvSome(ring_next_, stage + off, want, &got);
      if (!Ok(s)) return s;
      recvd += got;
    }
    const size_t avail = recvd / eb * eb;
    while (reduced < avail) {
      const size_t off = reduced % cap;
      const size_t pos = (recv_start + reduced) % total;
      size_t len = std::min(avail - reduced, cap - off);
      len = std::min(len, total - pos);
      if (src != nullptr) std::memcpy(buf + pos, src + pos, len);  // our share of this piece, cache-hot for the fold
      reducer(stage + off, buf + pos, static_cast<int>(len / eb), dtype);
      reduced += len;
    }
    sendable = std::min(total_send, head + reduced);
    if (sent < sendable && CanSend(ws, ring_prev_)) {
      const size_t pos = (send_start + sent) % total;
      // the first `head` bytes of the stream are our own untouched data, the rest are folded partial sums
      const bool own = src != nullptr && sent < head;
      const size_t upto = own ? std::min(sendable, head) : sendable;
      size_t put; Status s = SendSome(ring_prev_, (own ? src : buf) + pos, std::min(upto - sent, total - pos), &put);
      if (!Ok(s)) return s;
      sent += put;
    }
  }
  return Status::kOk;
}

Status HostComm::AllreduceFrom(const void *src, void *dst, size_t eb, size_t count, ReduceFunction *reducer) {
  if (count == 0) return Status::kOk;
  if (world <= 1 || !UseRing(eb, count) || src == dst) {
    if (src != dst) std::memcpy(dst, src, eb * count);
    return Allreduce(dst, eb, count, reducer);
  }
  return AllreduceRing(dst, eb, count, reducer, src);
}

Status HostComm::AllreduceRing(void *buf, size_t eb, size_t count, ReduceFunction *reducer, const void *src) {
  Status s = ReduceScatterRing(buf, eb, count, reducer, src);
  if (!Ok(s)) return s;
  const size_t n = static_cast<size_t>(world);
  const size_t step = (count + n - 1) / n;
  auto seg_begin = [&](size_t j) { return std::min(j * step, count); };
  const size_t r = static_cast<size_t>(rank), p = (r + n - 1) % n;
  return AllgatherRing(buf, count * eb, seg_begin(r) * eb, seg_begin(r + 1) * eb,
                       (seg_begin(p + 1) - seg_begin(p)) * eb);
}

// --------------------------------------------------------------- ring movers
Status HostComm::RelayImpl(Link *from, Link *to, void *buf_, size_t size) {
  char *buf = static_cast<char *>(buf_);
  size_t recvd = from != nullptr ? 0 : size;
  size_t sent = to != nullptr ? 0 : size;
  WaitSet ws;
  while (recvd < size || sent < size) {
    ws.Clear();
    if (recvd < size) WantRecv(&ws, from);
    if (sent < recvd) WantSend(&ws, to);
    // a neighbour this primitive does not talk to may have died: nothing would ever wake us up, so probe
    // all links whenever nothing moved for a while
    if (!WaitLinks(&ws)) return Status::kPeerClosed;
    if (recvd < size && CanRecv(ws, from)) {
      size_t got; Status s = RecvSome(from, buf + recvd, size - recvd, &got);
      if (!Ok(s)) return s;
      recvd += got;
    }
    if (sent < recvd && CanSend(ws, to)) {
      size_t put; Status s = SendSome(to, buf + sent, recvd - sent, &put);
      if (!Ok(s)) return s;
      sent += put;
    }
  }
  return Status::kOk;
}

Status HostComm::RingRelay(void *buf, size_t size, bool recv_from_prev, bool send_to_next) {
  if (size == 0 || (!recv_from_prev && !send_to_next)) return Status::kOk;
  utils::Assert(ring_prev_ != nullptr && ring_next_ != nullptr, "RingRelay without a ring");
  return RelayImpl(recv_from_prev ? ring_prev_ : nullptr, send_to_next ? ring_next_ : nullptr, buf, size);
}

Status HostComm::RingRelayBack(void *buf, size_t size, bool recv_from_next, bool send_to_prev) {
  if (size == 0 || (!recv_from_next && !send_to_prev)) return Status::kOk;
  utils::Assert(ring_prev_ != nullptr && ring_next_ != nullptr, "RingRelayBack without a ring");
  return RelayImpl(recv_from_next ? ring_next_ : nullptr, send_to_prev ? ring_prev_ : nullptr, buf, size);
}

Status HostComm::RingShift(const void *send_, size_t nsend, void *recv_, size_t nrecv, bool forward) {
  if (nsend == 0 && nrecv == 0) return Status::kOk;
  utils::Assert(ring_prev_ != nullptr && ring_next_ != nullptr, "RingShift without a ring");
  Link *to = forward ? ring_next_ : ring_prev_;
  Link *from = forward ? ring_prev_ : ring_next_;
  const char *send = static_cast<const char *>(send_);
  char *recv = static_cast<char *>(recv_);
  size_t sent = 0, recvd = 0;
  WaitSet ws;
  while (sent < nsend || recvd < nrecv) {
    ws.Clear();
    if (recvd < nrecv) WantRecv(&ws, from);
    if (sent < nsend) WantSend(&ws, to);
    // a neighbour this primitive does not talk to may have died: nothing would ever wake us up, so probe
    // all links whenever nothing moved for a while
    if (!WaitLinks(&ws)) return Status::kPeerClosed;
    if (recvd < nrecv && CanRecv(ws, from)) {
      size_t got; Status s = RecvSome(from, recv + recvd, nrecv - recvd, &got);
      if (!Ok(s)) return s;
      recvd += got;
    }
    if (sent < nsend && CanSend(ws, to)) {
      size_t put; Status s = SendSome(to, send + sent, nsend - sent, &put);
      if (!Ok(s)) return s;
      sent += put;
    }
  }
  return Status::kOk;
}

}  // namespace engine
}  // namespace rabit
