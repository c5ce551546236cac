/*!
 * rabit-b200: host-memory transport — tracker client, peer links and the fallible
 * collective primitives everything else is built from.
 *
 * Role parity with reference src/allreduce_base.{h,cc} (tracker handshake 221-438, tree
 * allreduce 476-668, flood broadcast 677-765, ring allgather 779-843, ring reduce-scatter
 * 857-946). The structure is different on purpose: every primitive is a small byte-flow
 * state machine over non-blocking sockets (byte counters + one poll loop each, with an idle probe
 * for neighbours that died silently), and recovery payloads move with `RingRelay` along the rank
 * ring instead of being routed over the tree.
 *
 * Every Try-style call returns a Status instead of throwing: a lost peer is a normal
 * event that the robust engine turns into "close everything, re-rendezvous, re-decide".
 */
#ifndef RABIT_SRC_HOST_COMM_H_
#define RABIT_SRC_HOST_COMM_H_

#include <string>
#include <vector>

#include "config.h"
#include "rabit/internal/engine.h"
#include "rabit/internal/socket.h"
#include "shm_ring.h"

namespace rabit {
namespace engine {

/*! \brief outcome of a fallible primitive */
enum class Status {
  kOk = 0,
  kPeerClosed,   // orderly close or reset by peer: the peer died or is itself recovering
  kSockError,    // any other socket level failure
  kTimeout,      // device backend only: a peer did not show up in the bounded spin
  kMismatch      // device backend only: fused consensus saw different requests
};
inline bool Ok(Status s) { return s == Status::kOk; }
const char *StatusName(Status s);
/*! \brief errno behind the most recent kSockError / kPeerClosed of this thread (diagnostics) */
int LastSocketErrno();
const char *LastSocketWhat();

/*! \brief one established peer connection */
struct Link {
  utils::TCPSocket sock;
  int rank = -1;
  /*! \brief the peer has closed with bytes still unread (payload, or its bye): stop watching for the hang-up */
  bool hup_seen = false;
  // ---- same-host data plane (shm_ring.h): payload through shared memory, the socket stays for liveness ----
  bool shm = false;           // both ends agreed on it
  ShmRing tx, rx;             // tx created here, rx opened from the peer
  sockaddr_un bell;           // the peer's doorbell
  socklen_t bell_len = 0;
};

/*! \brief what a primitive waits for in one turn of its loop: sockets, and conditions on shared-memory rings */
struct WaitSet {
  struct Item { Link *l; bool recv; size_t need; };   // recv: at least `need` unread bytes; send: room for a piece
  utils::PollHelper poll;
  std::vector<Item> shm;
  bool ready_now = false;     // one of the ring conditions already holds
  bool tcp = false;           // a socket is watched as well
  int left = -1;              // a ring is waited on whose other end has closed its connection: nothing will move
  void Clear() { poll.Clear(); shm.clear(); ready_now = false; tcp = false; left = -1; }
};

class HostComm {
 public:
  typedef IEngine::ReduceFunction ReduceFunction;

  HostComm();
  ~HostComm();

  // ---- identity ----
  Config cfg;
  int rank = 0;
  int world = 1;
  int parent_rank = -1;
  int ring_prev_rank = -1;
  int ring_next_rank = -1;
  std::string host;
  /*! \brief bumped on every successful (re)connect; device backend keys its mappings on it */
  int link_generation = 0;
  /*! \brief set when the tracker answers a "recover" with "the job is over" (peers have shut down) */
  bool job_over = false;

  bool distributed() const { return cfg.tracker_uri != "NULL"; }
  /*! \brief every link of THIS rank carries its payload through shared memory */
  bool AllLinksSharedMemory() const;
  /*!
   * \brief agreed by all ranks after a rendezvous (BaseEngine::AttachDevice): every link of the job is shared memory.
   *  The ring then beats the tree from 32 KiB (measured, profiles/r2_host_engine_cpu.md) instead of from
   *  rabit_reduce_ring_mincount elements; the choice must be the same on every rank, hence the agreement.
   */
  bool shm_everywhere = false;
  /*! \brief tree or ring for an allreduce of count items of elem_bytes */
  bool UseRing(size_t elem_bytes, size_t count) const {
    if (shm_everywhere && !cfg.reduce_ring_mincount_set) {
      return elem_bytes * count >= (32u << 10) && count >= static_cast<size_t>(world);
    }
    return count > cfg.reduce_ring_mincount;
  }

  // ---- rendezvous ----
  /*! \brief talk to the tracker with cmd "start" or "recover" and (re)build every link */
  bool Connect(const char *cmd);
  /*! \brief drop every peer connection (peers see it as our failure and follow suit) */
  void CloseLinks(bool graceful = false);
  /*!
   * \brief non-blocking health probe of every peer link: true if any peer closed or reset its
   *  end. The device backend calls this while a kernel is spinning on a peer's flag, so that a
   *  dead process becomes an abort of the kernel within milliseconds.
   */
  bool LinksBroken();
  /*!
   * \brief poll the sockets a primitive watches PLUS the hang-up of every link, so that the death of a neighbour
   *  the primitive does not talk to ends the wait at once instead of at the next idle probe.
   * \return false when a link is dead
   */
  bool WaitLinks(WaitSet *ws);
  /*! \brief ask the tracker (our tracker's "status" command) whether some rank has already left the job for good;
   *  *parked: whether the tracker still lists this worker as waiting for diallers */
  bool JobIsOver(bool *parked = nullptr);
  /*! \brief tell the tracker this rank is done */
  void SendShutdown();
  /*! \brief log line on the tracker, or stdout without one */
  void TrackerPrint(const std::string &msg);

  // ---- collectives on host memory ----
  Status Allreduce(void *buf, size_t elem_bytes, size_t count, ReduceFunction *reducer);
  Status AllreduceTree(void *buf, size_t elem_bytes, size_t count, ReduceFunction *reducer);
  Status AllreduceRing(void *buf, size_t elem_bytes, size_t count, ReduceFunction *reducer, const void *src = nullptr);
  /*!
   * \brief ring reduce-scatter. With src == NULL it works in place on buf; otherwise src holds this rank's
   *  contribution, stays untouched, and buf only receives: own data leaves straight from src and every
   *  segment is copied over piecewise right before it is folded (no whole-message copy up front).
   */
  Status ReduceScatterRing(void *buf, size_t elem_bytes, size_t count, ReduceFunction *reducer,
                           const void *src = nullptr);
  /*! \brief out-of-place allreduce: reads src, leaves the result in dst, src is never written (src may equal dst) */
  Status AllreduceFrom(const void *src, void *dst, size_t elem_bytes, size_t count, ReduceFunction *reducer);
  Status AllgatherRing(void *buf, size_t total, size_t slice_begin, size_t slice_end,
                       size_t size_prev_slice);
  Status Broadcast(void *buf, size_t size, int root);

  // ---- ring movers used by checkpoint replication and recovery ----
  /*!
   * \brief pipelined pass-along: optionally receive `size` bytes from the ring predecessor
   *  into buf and optionally forward buf to the ring successor (forwarding never runs
   *  ahead of what has been received).
   */
  Status RingRelay(void *buf, size_t size, bool recv_from_prev, bool send_to_next);
  /*! \brief the same in the opposite direction (from successor, to predecessor) */
  Status RingRelayBack(void *buf, size_t size, bool recv_from_next, bool send_to_prev);
  /*!
   * \brief independent send and receive with the two ring neighbours. forward: send to the
   *  successor / receive from the predecessor; otherwise the other way round. Either size
   *  may be zero.
   */
  Status RingShift(const void *send, size_t nsend, void *recv, size_t nrecv, bool forward);

 private:
  utils::TCPSocket DialTracker();
  Link *FindLink(int r);
  Status RelayImpl(Link *from, Link *to, void *buf, size_t size);
  // ---- link I/O: socket or shared-memory ring, same contract (non-blocking, "nothing yet" = 0 bytes and kOk) ----
  void WantRecv(WaitSet *ws, Link *l, size_t need = 1);
  void WantSend(WaitSet *ws, Link *l);
  bool CanRecv(const WaitSet &ws, Link *l) const;
  bool CanSend(const WaitSet &ws, Link *l) const;
  Status RecvSome(Link *l, char *p, size_t max, size_t *got);
  Status SendSome(Link *l, const char *p, size_t max, size_t *put);
  void RingBell(Link *l);
  void DropShm(Link *l);
  void ReplaceLink(Link *l, const Link &fresh);
  void OpenDoorbell();
  bool NegotiateSharedMemory(Link *l, bool dialler);
  int bell_fd_ = -1;                 // this process' doorbell (abstract unix datagram socket)
int spin_us_ = 0;                  // effective rabit_shm_spin_us (0 when the workers outnumber the cores)
    unsigned fast_turns_ = 0;          // consecutive waits answered from shared memory alone (no look at the sockets)
  std::string bell_name_;

  std::vector<Link> links_;          // every live peer connection
  std::vector<Link *> tree_;         // subset: tree neighbours
  int parent_slot_ = -1;             // index into tree_ of the parent, -1 on the root
  Link *ring_prev_ = nullptr;
  Link *ring_next_ = nullptr;
  std::vector<char> scratch_;        // staging for reductions
  std::vector<char> item_;           // one item that straddled the wrap of a shared-memory ring
  bool rank_assigned_ = false;       // the tracker has told us who we are
  bool world_known_ = false;
};

}  // namespace engine
}  // namespace rabit
#endif  // RABIT_SRC_HOST_COMM_H_
