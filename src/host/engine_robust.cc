/*!
 * rabit-b200: RobustEngine — sequence-numbered result cache, consensus-driven recovery,
 * two-phase in-memory checkpoints, ring-replicated local models, bootstrap cache, watchdog.
 *
 * Semantics follow reference src/allreduce_robust.cc (collective wrappers 172-360,
 * LoadCheckPoint 383-446, CheckPoint_ 484-557, CheckAndRecover 687-725, RecoverExec
 * 1158-1311, local-state ring recovery 1328-1587); the mechanism is this build's own:
 *   - consensus = proto::Decide over an allreduced Summary (protocol.h);
 *   - holders/requesters are discovered with GatherTable, so ANY rank that still has the
 *     bytes can serve, and every rank computes the same transfer plan locally;
 *   - payloads travel along the rank ring (RingDeliver) — on an NVSwitch box "routing" is
 *     meaningless, and on TCP the ring links already exist.
 */
#include <chrono>
#include <cstdlib>
#include <thread>

#include "../device/device_backend.h"
#include "engine_core.h"

namespace rabit {
namespace engine {

using proto::Request;
using proto::Summary;

RobustEngine::RobustEngine() {}

RobustEngine::~RobustEngine() {
  watchdog_cancel_ = true;
  if (watchdog_.joinable()) watchdog_.join();
}

bool RobustEngine::Init(int argc, char *argv[]) {
  if (!BaseEngine::Init(argc, argv)) return false;
  if (comm_.cfg.bootstrap_cache) utils::HandleLogInfo("[rabit] bootstrap cache enabled\n");
  num_local_replica_ = comm_.cfg.local_replica;
  checkpoint_loaded_ = false;
  results_.SetDevice(device_.get());
  return true;
}

int RobustEngine::ReplicaRound() const {
  // a result with sequence number s is retained on ranks r with r % round == s % round, i.e.
  // on about `global_replica` ranks (reference allreduce_robust.cc:47-51); -1 = keep nothing old
  if (comm_.cfg.global_replica == 0) return -1;
  return std::max(comm_.world / comm_.cfg.global_replica, 1);
}

bool RobustEngine::Shutdown(void) {
  try {
    if (comm_.world > 1) {
      // Leave through a pseudo commit so nobody exits while a peer still needs its cache.
      // A lazy checkpoint is serialised now: the model it points to is still intact (no collective followed the
      // last LazyCheckPoint) but may be destroyed by the caller right after Finalize.
      MaterializeLazy();
      utils::Assert(Consensus(nullptr, 0, proto::kCheckPoint, proto::kSeqSpecial), "shutdown barrier 1");
      results_.Clear(); seq_ = 0; boot_cache_.clear();
      utils::Assert(Consensus(nullptr, 0, proto::kCheckAck, proto::kSeqSpecial), "shutdown barrier 2");
    }
    StopWatchdog();
    results_.Clear();
    results_.SetDevice(nullptr);
    return BaseEngine::Shutdown();
  } catch (const std::exception &e) {
    if (comm_.job_over) {
      // the final barrier round reached the others but not us; they are gone and so is any need for our cache
      StopWatchdog();
      results_.Clear();
      results_.SetDevice(nullptr);
      return BaseEngine::Shutdown();
    }
    fprintf(stderr, "[rabit] shutdown: %s\n", e.what());
    // This rank leaves all the same (the caller is past its last collective): say so, or the ranks that are still
    // trying to re-join would wait for it until their timeouts.
    StopWatchdog();
    if (device_) { device_->Detach(); device_.reset(); }
    comm_.CloseLinks();
    comm_.SendShutdown();
    initialized_ = false;
    return false;
  }
}

void RobustEngine::AttachAfterStart() {
  if (this->AttachDevice() != AttachResult::kLinkError) return;
  ++recover_counter_;
  in_sync_ = false;
  if (comm_.cfg.debug) {
    utils::HandleLogInfo("[%d] a peer vanished during the device bootstrap of Init, re-joining\n", comm_.rank);
  }
  StartWatchdog();
  this->Rejoin();
}

void RobustEngine::InitAfterException(void) {
  // the caller swallowed an engine exception and wants to carry on in this process
  this->Rejoin();
}

/*!
 * \brief drop every link, meet the others at the tracker's "recover" rendezvous, rebuild the device arena.
 *  A second peer may die while the survivors are in here (a link breaks during the arena exchange): that is
 *  one more round of the same loop, not a fatal error.
 */
void RobustEngine::Rejoin() {
  const bool had_device = device_ != nullptr;
  static const bool profile = getenv("RABIT_RECOVER_PROFILE") != nullptr;
  const double t0 = utils::GetTime();
  while (true) {
    if (device_) device_->Detach();
    comm_.CloseLinks();
    const double t1 = utils::GetTime();
    std::this_thread::sleep_for(std::chrono::milliseconds(2 * comm_.rank));
    while (!comm_.Connect("recover")) {
      utils::Check(!comm_.job_over, "[%d] cannot recover: the other ranks have already shut down", comm_.rank);
      std::this_thread::sleep_for(std::chrono::milliseconds(100));
    }
    const double t2 = utils::GetTime();
    if (!had_device && comm_.world <= 1) return;
    const AttachResult r = this->AttachDevice();
    if (r == AttachResult::kLinkError) {
      if (comm_.cfg.debug) {
        utils::HandleLogInfo("[%d] a link broke while the device arena was being re-attached (%s), re-joining again\n",
                             comm_.rank, LastSocketWhat());
      }
      continue;
    }
    // cached results and checkpoints of this rank live in the backend's memory: it cannot simply go away
    utils::Check(!(had_device && r == AttachResult::kNoDevice),
                 "[%d] device arena could not be re-established after recovery", comm_.rank);
    if (profile || comm_.cfg.debug) {
      utils::HandleLogInfo("[%d] re-join: detach+close %.1f ms, rendezvous (incl. waiting for the replacement) %.1f ms, "
                           "device attach %.1f ms\n", comm_.rank, (t1 - t0) * 1e3, (t2 - t1) * 1e3,
                           (utils::GetTime() - t2) * 1e3);
    }
    return;
  }
}

// ------------------------------------------------------------------ failures
void RobustEngine::StartWatchdog() {
  if (!comm_.cfg.timeout_enabled || watchdog_.joinable()) return;
  utils::Printf("[rabit] recovery watchdog armed: %d s\n", comm_.cfg.timeout_sec);
  watchdog_cancel_ = false; watchdog_done_ = false; watchdog_fired_ = false;
  const int secs = comm_.cfg.timeout_sec, me = comm_.rank;
  void (*hook)(int, int) = on_timeout;
  watchdog_ = std::thread([this, secs, me, hook]() {
    for (int tick = 0; tick < secs * 10; ++tick) {
      std::this_thread::sleep_for(std::chrono::milliseconds(100));
      if (watchdog_cancel_.load()) { watchdog_done_ = true; return; }
    }
    watchdog_fired_ = true; watchdog_done_ = true;
    fprintf(stderr, "[%d] exit due to time out %d s\n", me, secs);
    fflush(stderr);
    if (hook != nullptr) { hook(me, secs); return; }
    // The reference only throws inside the async task, which nobody observes;
    // the documented intent is to stop the worker, so do that.
    _exit(10);
  });
}

void RobustEngine::StopWatchdog() {
  if (!watchdog_.joinable()) return;
  watchdog_cancel_ = true;
  watchdog_.join();
}

bool RobustEngine::Healthy(Status s) {
  if (Ok(s)) { StopWatchdog(); return true; }
  ++recover_counter_;
  in_sync_ = false;
  if (comm_.cfg.debug) {
    utils::HandleLogInfo("[%d] link failure (%s, errno %d, %s) at version %d seq %d, re-joining\n", comm_.rank,
                         StatusName(s), LastSocketErrno(), LastSocketWhat(), version_number_, seq_);
  }
  StartWatchdog();
  // Drop EVERY link: neighbours observe it as a failure too and the whole job funnels into the
  // tracker's "recover" rendezvous, where it waits for the replacement process.
  this->Rejoin();
  return false;
}

// ----------------------------------------------------------------- consensus
namespace {
thread_local proto::InlineOp *tl_inline = nullptr;  // reducer context while the envelope allreduce runs

void EnvelopeReduce(const void *src_, void *dst_, int count, const MPI::Datatype &) {
  const proto::Envelope *src = static_cast<const proto::Envelope *>(src_);
  proto::Envelope *dst = static_cast<proto::Envelope *>(dst_);
  for (int i = 0; i < count; ++i) {
    dst[i].sum.Merge(src[i].sum);
    dst[i].dev_epoch = std::max(dst[i].dev_epoch, src[i].dev_epoch);
    dst[i].dev_rounds = std::max(dst[i].dev_rounds, src[i].dev_rounds);
    const bool same_sig = src[i].signature == dst[i].signature && src[i].bytes == dst[i].bytes;
    const uint32_t a = src[i].state, b = dst[i].state;
    proto::InlineOp *op = tl_inline;
    if (a == 1 && b == 1 && same_sig && op != nullptr && op->bcast_root < 0 &&
        op->item_bytes * op->count == dst[i].bytes) {
      op->reducer(src[i].payload, dst[i].payload, static_cast<int>(op->count), MPI::Datatype(op->item_bytes));
    } else if (same_sig && ((a == 4 && b == 3) || (a == 3 && b == 4) || (a == 3 && b == 3))) {
      // broadcast: the root's bytes win over the receivers' empty payloads
      if (a == 4) { std::memcpy(dst[i].payload, src[i].payload, src[i].bytes); dst[i].state = 4; }
    } else {
      dst[i].state = 2;  // somebody is elsewhere in the program: whatever is in the payload is meaningless
    }
  }
}
}  // namespace

bool RobustEngine::Consensus(void *buf, size_t size, uint32_t flags, uint32_t seq,
                             const std::function<void *()> *lazy_buf, proto::InlineOp *inline_op) {
  Request me{seq, flags, static_cast<uint32_t>(boot_cache_.size())};
  std::unique_ptr<proto::Envelope> env(new proto::Envelope());
  while (true) {
    this->ReportStatus();
    me.cache_count = static_cast<uint32_t>(boot_cache_.size());
    std::memset(env.get(), 0, sizeof(proto::Envelope));
    env->sum = Summary::Of(me);
    const bool dev_up = device_ != nullptr && device_->attached();
    if (dev_up) { env->dev_epoch = device_->OpEpoch(); env->dev_rounds = device_->OpRounds(); }
    if (inline_op != nullptr) {
      const bool bcast = inline_op->bcast_root >= 0;
      env->state = !bcast ? 1 : (inline_op->bcast_root == comm_.rank ? 4 : 3);
      env->bytes = static_cast<uint32_t>(inline_op->item_bytes * inline_op->count);
      env->signature = (seq * 2654435761u) ^ (static_cast<uint32_t>(inline_op->item_bytes) << 20) ^
                       static_cast<uint32_t>(inline_op->count) ^ (bcast ? 0x80000000u + inline_op->bcast_root : 0u);
      if (env->state != 3) std::memcpy(env->payload, inline_op->data, env->bytes);
    }
    tl_inline = inline_op;
    const Status st = comm_.AllreduceTree(env.get(), sizeof(proto::Envelope), 1, EnvelopeReduce);
    tl_inline = nullptr;
    if (!Healthy(st)) continue;
    if (dev_up) device_->AdoptEpoch(env->dev_epoch, env->dev_rounds);
    const Summary sum = env->sum;
    const proto::Verdict v = proto::Decide(sum, me);
    switch (v.action) {
      case proto::kExecute:
        if (inline_op != nullptr && env->state == (inline_op->bcast_root >= 0 ? 4u : 1u)) {
          // every rank shipped this very operation: the reduced payload IS the result
          std::memcpy(inline_op->data, env->payload, env->bytes);
          inline_op->done = true;
        }
        return false;
      case proto::kDone: return true;
      case proto::kRetry: break;
      case proto::kServeLoad:
        if (Healthy(ServeLoad(v.requester)) && v.requester) return true;
        break;
      case proto::kServeCache:
        if (comm_.cfg.debug) {
          utils::HandleLogInfo("[%d] bootstrap cache sync: mine %u, max %u, requester %d\n", comm_.rank,
                               me.cache_count, sum.max_cache, static_cast<int>(v.requester));
        }
        utils::Check(!((flags & proto::kLoadCache) && !v.requester && !sum.all(proto::kLoadCache)),
                     "[%d] bootstrap collective is not in the peers' cache: the calling sequence before "
                     "LoadCheckPoint differs between runs", comm_.rank);
        if (Healthy(ServeCache(v.requester, sum.max_cache)) && v.requester) return true;
        break;
      case proto::kServeResult:
        // the landing buffer of a requester may be expensive (host mirror of a device tensor):
        // it is only materialised when this rank really is about to receive
        if (v.requester && buf == nullptr && lazy_buf != nullptr) buf = (*lazy_buf)();
        if (Healthy(ServeResult(buf, size, sum.min_seq, v.requester)) && v.requester) return true;
        break;
    }
  }
}

Status RobustEngine::GatherTable(const uint64_t *mine, int nwords, std::vector<uint64_t> *table) {
  table->assign(static_cast<size_t>(comm_.world) * nwords, 0);
  for (int i = 0; i < nwords; ++i) (*table)[static_cast<size_t>(comm_.rank) * nwords + i] = mine[i];
  // everyone contributes zeros outside its own row, so a bitwise OR assembles the table
  return comm_.AllreduceTree(table->data(), sizeof(uint64_t), table->size(),
                             [](const void *s, void *d, int n, const MPI::Datatype &) {
                               const uint64_t *src = static_cast<const uint64_t *>(s);
                               uint64_t *dst = static_cast<uint64_t *>(d);
                               for (int i = 0; i < n; ++i) dst[i] |= src[i];
                             });
}

Status RobustEngine::RingDeliver(const std::vector<char> &has, const std::vector<char> &need,
                                 void *buf, size_t size) {
  // Bytes flow forward (r -> r+1). A rank without the bytes must take them from its predecessor
  // if some requester sits between it (inclusive) and the next holder downstream.
  const int n = comm_.world, me = comm_.rank;
  auto must_recv = [&](int r) {
    if (has[r]) return false;
    for (int q = r, hops = 0; hops < n && !has[q]; q = (q + 1) % n, ++hops) {
      if (need[q]) return true;
    }
    return false;
  };
  const bool recv = must_recv(me);
  const bool send = must_recv((me + 1) % n);
  if (!recv && !send) return Status::kOk;
  std::string pass;  // pure relays need a landing zone
  if (buf == nullptr) { pass.resize(size); buf = &pass[0]; }
  return comm_.RingRelay(buf, size, recv, send);
}

}  // namespace engine
}  // namespace rabit
