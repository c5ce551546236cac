/*!
 * rabit-b200: the recovery consensus, stated as data + one pure decision function.
 *
 * Before every collective / checkpoint step each rank publishes a Request; an allreduce
 * with SummaryReduce folds them into a Summary that is identical on all ranks, and
 * Decide() maps (Summary, my Request) to what THIS rank does next. Because every rank
 * evaluates the same pure function on the same Summary, all ranks take matching branches.
 *
 * This plays the role of the reference's ActionSummary + RecoverExec pair
 * (allreduce_robust.h:224-322, allreduce_robust.cc:1158-1311). It is re-derived, not
 * transcribed: priorities are explicit (load-ack > commit-ack > bootstrap-cache > load > sequence order),
 * the bootstrap cache is compared by entry count, and "who holds the bytes" is NOT part
 * of the summary — holders are discovered afterwards with a per-rank table gather and the
 * bytes move along the rank ring (see RobustEngine::Serve*).
 *
 * The same 32-bit code (seq << 5 | flags) is what the sm_100a kernels exchange in their
 * entry barrier ("fused consensus"): if all codes are equal and flag-free the collective
 * proceeds without any host round trip.
 */
#ifndef RABIT_SRC_HOST_PROTOCOL_H_
#define RABIT_SRC_HOST_PROTOCOL_H_

#include <algorithm>
#include <cstdint>

namespace MPI { class Datatype; }

namespace rabit {
namespace engine {
namespace proto {

// ---- request flags ----
const uint32_t kLoadCheck = 1;    // wants the latest checkpoint (LoadCheckPoint after a restart)
const uint32_t kCheckPoint = 2;   // waiting at commit barrier 1
const uint32_t kCheckAck = 4;     // waiting at commit barrier 2
const uint32_t kLoadCache = 8;    // bootstrap phase: wants/offers the pre-checkpoint op cache
const uint32_t kLoadAck = 16;     // a restarted rank has finished loading; parks everybody else for one round
const uint32_t kFlagBits = 5;
const uint32_t kFlagMask = (1u << kFlagBits) - 1;

// ---- sequence numbers ----
// plain collectives count 0,1,2,... within a version; the phases of a commit sort after them
const uint32_t kSeqLocalCheckPoint = (1u << 26) - 2;  // replicating per-rank models
const uint32_t kSeqLocalAck = (1u << 26) - 1;         // replication complete everywhere
const uint32_t kSeqSpecial = 1u << 26;                // load / checkpoint / ack

/*! \brief what one rank is about to do */
struct Request {
  uint32_t seq;
  uint32_t flags;
  uint32_t cache_count;  // entries in this rank's bootstrap cache
  /*! \brief packed form used by the device kernels' fused entry barrier */
  uint32_t Code() const { return (seq << kFlagBits) | (flags & kFlagMask); }
};

/*! \brief fold of all ranks' requests */
struct Summary {
  uint32_t min_seq, max_seq;
  uint32_t any_flags, all_flags;
  uint32_t min_cache, max_cache;
  static Summary Of(const Request &r) {
    return Summary{r.seq, r.seq, r.flags, r.flags, r.cache_count, r.cache_count};
  }
  void Merge(const Summary &o) {
    min_seq = std::min(min_seq, o.min_seq); max_seq = std::max(max_seq, o.max_seq);
    any_flags |= o.any_flags; all_flags &= o.all_flags;
    min_cache = std::min(min_cache, o.min_cache); max_cache = std::max(max_cache, o.max_cache);
  }
  bool any(uint32_t f) const { return (any_flags & f) != 0; }
  bool all(uint32_t f) const { return (all_flags & f) != 0; }
  bool uniform_seq() const { return min_seq == max_seq; }
};

/*! \brief engine ReduceFunction over Summary items */
inline void SummaryReduce(const void *src_, void *dst_, int count, const MPI::Datatype &) {
  const Summary *src = static_cast<const Summary *>(src_);
  Summary *dst = static_cast<Summary *>(dst_);
  for (int i = 0; i < count; ++i) dst[i].Merge(src[i]);
}

/*!
 * \brief what actually travels in a consensus round: the summary plus, for tiny allreduces, the payload
 *  itself. If the verdict is "execute now" and every rank shipped the same operation, the reduced payload
 *  is already the result and the separate payload collective is skipped — one network round trip per
 *  small allreduce instead of the reference's two (allreduce_robust.cc:267 then :282).
 *  Every rank sends an Envelope in every round (fixed size), whatever it is doing.
 */
const uint32_t kInlineBytes = 1024;
struct Envelope {
  Summary sum;
  uint32_t state;      // 0 no payload, 1 allreduce payload valid so far, 2 poisoned (ranks disagree on the
                       // operation), 3 broadcast receiver, 4 broadcast payload (from the root)
  uint32_t signature;  // identifies the operation: hash of (seq, item size, item count)
  uint32_t bytes;
  // op counters of the device kernels' flag protocols, folded with max: after a kernel was abandoned on some
  // ranks (consensus mismatch, failed chunk of a multi-kernel op) the counters differ, and every rank adopts
  // the largest before the next fused kernel (DeviceBackend::AdoptEpoch)
  uint32_t dev_epoch;
  uint64_t dev_rounds;
  unsigned char payload[kInlineBytes];
};
/*! \brief reducer context of the rank running the envelope allreduce (thread local, set by Consensus) */
struct InlineOp {
  void *data;                // in: this rank's contribution, out: the result when done
  size_t item_bytes, count;  // broadcast: item_bytes = 1, count = size
  int bcast_root;            // -1: allreduce; otherwise broadcast from this rank
  void (*reducer)(const void *src, void *dst, int count, const MPI::Datatype &dtype);
  bool done;
};

enum Action {
  kExecute = 0,     // nobody is behind: run the requested step now
  kDone = 1,        // the request is satisfied (barrier passed / data recovered)
  kRetry = 2,       // helped (or waited for) someone else; ask again
  kServeLoad = 3,   // move checkpoint state to the ranks asking for it; they are then done
  kServeCache = 4,  // move the bootstrap cache to ranks with fewer entries
  kServeResult = 5  // move the result of op `min_seq` to the ranks that are at min_seq
};

/*! \brief role of this rank inside a Serve* action */
struct Verdict {
  Action action;
  bool requester;  // for kServe*: this rank receives; it is done afterwards
};

inline Verdict Decide(const Summary &s, const Request &me) {
  const bool i_load = me.flags & kLoadCheck, i_ckpt = me.flags & kCheckPoint;
  const bool i_ack = me.flags & kCheckAck, i_cache = me.flags & kLoadCache;
  // 1. Acknowledgements. kCheckAck: somebody already passed commit barrier 1. kLoadAck: a restarted rank holds the
  //    checkpoint now (survivors are parked until then, so that they cannot run ahead and drop state it is still
  //    copying). The reference uses ONE flag for both (allreduce_robust.cc:438-441 and :549-552); that lets a rank
  //    waiting at commit barrier 1 slip through on a loader's acknowledgement while a third rank is still inside the
  //    version's last collective, and the rank that committed then drops results the others still need.
  //    Ranks still knocking at barrier 1 of the SAME commit only got a failed consensus round: they go first,
  //    then loaders are served (they will see the NEW version), then everybody who acknowledged leaves together.
  const bool i_loadack = me.flags & kLoadAck;
  if (s.any(kCheckAck | kLoadAck)) {
    if (s.any(kCheckPoint)) {
      if (s.any(kCheckAck)) return Verdict{i_ckpt ? kDone : kRetry, false};
      // only a loader acknowledged: the knockers belong to a commit that has not begun
      return Verdict{i_loadack ? kDone : kRetry, false};
    }
    if (s.any(kLoadCheck)) return Verdict{kServeLoad, i_load};
    return Verdict{(i_ack || i_loadack) ? kDone : kRetry, false};
  }
  // 2. Bootstrap phase (collectives issued before LoadCheckPoint).
  if (s.any(kLoadCache)) {
    if (s.all(kLoadCache) && s.min_cache == s.max_cache) return Verdict{kExecute, false};
    return Verdict{kServeCache, i_cache && me.cache_count < s.max_cache};
  }
  // 3. A restarted rank wants the checkpoint.
  if (s.any(kLoadCheck)) {
    if (s.all(kLoadCheck)) return Verdict{kExecute, false};  // everybody is new: fresh start
    // every survivor is already at the commit barrier: finish the commit, then serve the new version
    if (s.any(kCheckPoint) && s.min_seq == kSeqSpecial) return Verdict{i_ckpt ? kDone : kRetry, false};
    return Verdict{kServeLoad, i_load};
  }
  // 4. Plain sequence order (ops < local-checkpoint < local-ack < commit barrier 1).
  if (s.uniform_seq()) {
    if (i_ckpt || me.seq == kSeqLocalAck) return Verdict{kDone, false};
    return Verdict{kExecute, false};
  }
  return Verdict{kServeResult, me.seq == s.min_seq};
}

}  // namespace proto
}  // namespace engine
}  // namespace rabit
#endif  // RABIT_SRC_HOST_PROTOCOL_H_
