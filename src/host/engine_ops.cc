/*!
 * rabit-b200: RobustEngine public operations — the collective wrapper shared by
 * Allreduce/Broadcast/Allgather, LoadCheckPoint and the two-phase Commit.
 * Reference: allreduce_robust.cc:172-360 (wrappers), 383-446 (load), 484-557 (checkpoint).
 */
#include "../device/device_backend.h"
#include "engine_core.h"

namespace rabit {
namespace engine {

void RobustEngine::DropStaleResult() {
  // the newest result is always kept until the NEXT op has passed consensus (a peer may have
  // failed at the tail of that op); after that only designated replicas retain it
  const int last = results_.LastSeq();
  if (last < 0) return;
  const int round = ReplicaRound();
  if (round == -1 || last % round != comm_.rank % round) results_.DropLast();
}

/*!
 * One attempt of the transfer, out of place: `input` (the caller's bytes) is only read, the result is built
 * in the cache slot. A failure half-way therefore leaves the caller's buffer intact for the retry, and no
 * whole-message copy is needed up front: only the bytes this rank actually contributes are taken over.
 */
Status RobustEngine::TryOp(const OpArgs &a, char *slot, const char *input) {
  switch (a.kind) {
    case OpKind::kAllreduce: return comm_.AllreduceFrom(input, slot, a.elem_bytes, a.count, a.reducer);
    case OpKind::kBroadcast:
      if (comm_.rank == a.root) std::memcpy(slot, input, a.bytes);
      return comm_.Broadcast(slot, a.bytes, a.root);
    case OpKind::kAllgather:
      std::memcpy(slot + a.slice_begin, input + a.slice_begin, a.slice_end - a.slice_begin);
      return comm_.AllgatherRing(slot, a.bytes, a.slice_begin, a.slice_end, a.size_prev_slice);
  }
  return Status::kSockError;
}

void RobustEngine::RunOp(OpArgs *a) {
  const bool on_device = Classify(a->buf) == MemKind::kDevice;
  // a large typed allreduce on a HOST buffer rides PCIe + NVLink through the same kernels (HostOverDevice)
  const bool host_fused = !on_device && HostOverDevice(*a);
  if (comm_.world <= 1 && !((on_device || host_fused) && device_ && comm_.cfg.device_force)) {
    if (a->prepare_fun != nullptr) a->prepare_fun(a->prepare_arg);
    return;
  }
  // A device pointer without an attached arena (ranks on several hosts, rabit_device=off, no P2P) still
  // works: the bytes are staged through the host collectives below, recovery semantics unchanged.
  auto to_device = [&](void *dst, const void *src, size_t n) {
    if (device_) device_->CopyH2D(dst, src, n); else DeviceBackend::RawCopy(dst, src, n, true);
  };
  auto to_host = [&](void *dst, const void *src, size_t n) {
    if (device_) device_->CopyD2H(dst, src, n); else DeviceBackend::RawCopy(dst, src, n, false);
  };
  const double t0 = utils::GetTime();
  const bool boot = comm_.cfg.bootstrap_cache && !checkpoint_loaded_;
  const uint32_t flags = boot ? proto::kLoadCache : 0;

  // device buffers whose op the kernels cannot express go through a host mirror
  const bool fused = (on_device || host_fused) && device_ != nullptr && device_->attached() &&
                     (a->kind != OpKind::kAllreduce || (a->dtype >= 0 && a->op >= 0));
  // (allocated only when the host protocol actually has to look at the bytes)
  std::string mirror;
  char *view = static_cast<char *>(a->buf);
  auto need_view = [&]() {
    if (on_device && mirror.size() != a->bytes) { mirror.resize(a->bytes); view = &mirror[0]; }
  };
  if (on_device) view = nullptr;

  bool prepared = false, recovered = false;
  auto prepare_once = [&]() {
    if (!prepared && a->prepare_fun != nullptr) a->prepare_fun(a->prepare_arg);
    prepared = true;
  };

  if (boot) need_view();
  if (boot && BootLookup(a->key, view, a->bytes)) {
    if (on_device) to_device(a->buf, view, a->bytes);
    return;
  }

  char *slot = nullptr;
  bool dropped = false;
  const bool use_kernel = fused && !boot;
  // set once a kernel has reported a mismatch although the host protocol says that every rank is at this very
  // op: the ranks chose different kernels (differently aligned / placed buffers), so take the variant that
  // depends on nothing but the op itself
  int dev_flags = 0;
  bool mismatched = false;
  while (true) {
    if (use_kernel && in_sync_) {
      // Fast path: ONE kernel does consensus + transfer + reduce + the cache copy. It is only
      // attempted while this rank has reason to believe everybody is at the same op; it falls
      // back to the host protocol when the ranks disagree or a peer is gone.
      const uint32_t code = proto::Request{static_cast<uint32_t>(seq_), 0, 0}.Code();
      prepare_once();
      void *keep = results_.AllocTempDevice(a->bytes);
      Status st = Status::kOk;
      switch (a->kind) {
        case OpKind::kAllreduce:
          st = device_->Allreduce(a->buf, a->bytes / DTypeBytes(a->dtype), a->dtype, a->op, 1, code, keep, dev_flags);
          break;
        case OpKind::kBroadcast: st = device_->Broadcast(a->buf, a->bytes, a->root, code, keep, dev_flags); break;
        case OpKind::kAllgather:
          st = device_->Allgather(a->buf, a->bytes, a->slice_begin, a->slice_end, code, keep);
          break;
      }
      if (Ok(st)) {
        // consensus passed inside the kernel: the previous result may go
        if (!dropped) { DropStaleResult(); dropped = true; }
        break;
      }
      results_.DiscardTemp();
      in_sync_ = false;
      // kMismatch: somebody is elsewhere in the protocol. Anything else: a peer is gone, tear down
      // and re-join. Either way the host protocol decides what happens next.
      if (st != Status::kMismatch) Healthy(st);
      else mismatched = true;
    }
    const std::function<void *()> lazy = [&]() -> void * { need_view(); return view; };
    // tiny host allreduces without a lazy prepare ride along with the consensus round itself
    proto::InlineOp inl{view, a->elem_bytes, a->count, -1, a->reducer, false};
    if (a->kind == OpKind::kBroadcast) { inl.item_bytes = 1; inl.count = a->bytes; inl.bcast_root = a->root; }
    const bool can_inline = (a->kind == OpKind::kAllreduce || a->kind == OpKind::kBroadcast) && !on_device &&
                            a->prepare_fun == nullptr && a->bytes != 0 && a->bytes <= proto::kInlineBytes;
    recovered = Consensus(view, a->bytes, flags, static_cast<uint32_t>(seq_), &lazy, can_inline ? &inl : nullptr);
    if (boot && recovered) {
      // the cache was just restored from a peer: this op must be in it
      utils::Check(BootLookup(a->key, view, a->bytes),
                   "[%d] bootstrap collective %s is not in the restored cache", comm_.rank, a->key.c_str());
      if (on_device) to_device(a->buf, view, a->bytes);
      return;
    }
    if (!dropped) { DropStaleResult(); dropped = true; }
    if (recovered) {
      slot = results_.AllocTemp(a->bytes);
      std::memcpy(slot, view, a->bytes);
      if (on_device) to_device(a->buf, view, a->bytes);
      break;
    }
    in_sync_ = true;  // the consensus round just said: every rank is about to run this very op
    if (mismatched) dev_flags |= DeviceBackend::kOpConservative;
    if (inl.done) {
      // the result came back with the consensus round: cache it like any other
      slot = results_.AllocTemp(a->bytes);
      std::memcpy(slot, view, a->bytes);
      break;
    }
    if (use_kernel) continue;
    need_view();
    prepare_once();
    slot = results_.AllocTemp(a->bytes);
    if (on_device) to_host(view, a->buf, a->bytes);
    // the transfer builds its result in the cache slot so that a failure half-way leaves the caller's
    // buffer intact for the retry
    if (Healthy(TryOp(*a, slot, view))) {
      // (the root of a broadcast already holds the result)
      if (!(a->kind == OpKind::kBroadcast && comm_.rank == a->root)) {
        std::memcpy(view, slot, a->bytes);
        if (on_device) to_device(a->buf, view, a->bytes);
      }
      break;
    }
  }

  if (comm_.cfg.debug) {
    utils::HandleLogInfo("[%d] %s (%s) finished version %d, seq %d, take %f seconds%s\n", comm_.rank,
                         a->kind == OpKind::kAllreduce ? "allreduce"
                         : a->kind == OpKind::kBroadcast ? "broadcast" : "allgather",
                         a->key.c_str(), version_number_, seq_, utils::GetTime() - t0,
                         recovered ? " [recovered]" : "");
  }
  if (boot) {
    boot_cache_.emplace_back(a->key, std::string(slot, a->bytes));
    results_.DiscardTemp();
  } else {
    results_.Commit(seq_);
    ++seq_;
    if (device_) results_.SpillTo(comm_.cfg.device_cache_bytes);
  }
}

/*!
 * A large HOST buffer rides PCIe + NVLink instead of the TCP links (no counterpart in the reference, whose
 * only transport is the socket ring of allreduce_base.cc:857-977). It is ONE sequence-numbered op: the device
 * backend cuts it into chunks and overlaps H2D(i+1) | kernel(i) | D2H(i-1) on copy engines; the stage copy
 * of the input is the backup that makes the op all-or-nothing (device_backend_stream.inc).
 * The decision uses only values that are the same on every rank.
 */
bool RobustEngine::HostOverDevice(const OpArgs &a) const {
  const Config &cfg = comm_.cfg;
  if (a.kind != OpKind::kAllreduce || device_ == nullptr || !device_->attached()) return false;
  if (cfg.device_host_min == 0 || a.bytes < cfg.device_host_min) return false;
  if (a.dtype < 0 || a.op < 0 || a.elem_bytes == 0) return false;
  if (cfg.bootstrap_cache && !checkpoint_loaded_) return false;  // bootstrap results live in the host cache
  return true;
}

void RobustEngine::AllreduceTyped(void *buf, size_t type_nbytes, size_t count, ReduceFunction reducer,
                                  int dtype, int op, PreprocFunction prepare_fun, void *prepare_arg,
                                  const char *_file, const int _line, const char *_caller) {
  OpArgs a{};
  a.kind = OpKind::kAllreduce; a.buf = buf; a.bytes = type_nbytes * count;
  a.elem_bytes = type_nbytes; a.count = count; a.reducer = reducer; a.dtype = dtype; a.op = op;
  a.prepare_fun = prepare_fun; a.prepare_arg = prepare_arg;
  a.key = std::string(_file) + "::" + std::to_string(_line) + "::" + _caller + "#" +
          std::to_string(type_nbytes) + "x" + std::to_string(count);
  // An op that goes through the kernels must fit the arena as ONE recoverable unit. What does not (symmetric
  // buffers pin the arena, or more than 2 GiB) is cut into pieces that are sequence-numbered ops of their own:
  // each has its own cache entry, commits or fails alone and is replayed alone. Same cut on every rank.
  const bool kernels = device_ != nullptr && device_->attached() && dtype >= 0 && op >= 0 && type_nbytes != 0 &&
                       (comm_.world > 1 || comm_.cfg.device_force) &&
                       !(comm_.cfg.bootstrap_cache && !checkpoint_loaded_) &&
                       (Classify(buf) == MemKind::kDevice || HostOverDevice(a));
  if (kernels) {
    const size_t cap = device_->MaxOpBytes(a.bytes) / type_nbytes * type_nbytes;
    if (cap != 0 && a.bytes > cap) {
      bool first = true;
      for (size_t off = 0; off < a.bytes; off += cap) {
        OpArgs piece = a;
        piece.buf = static_cast<char *>(buf) + off;
        piece.bytes = std::min(cap, a.bytes - off);
        piece.count = piece.bytes / type_nbytes;
        piece.key = a.key + "/" + std::to_string(off / cap);
        if (!first) piece.prepare_fun = nullptr;   // the lazy preparation fills the whole buffer, once
        RunOp(&piece);
        first = false;
      }
      return;
    }
  }
  RunOp(&a);
}

void RobustEngine::Broadcast(void *buf, size_t size, int root, const char *_file, const int _line,
                             const char *_caller) {
  OpArgs a{};
  a.kind = OpKind::kBroadcast; a.buf = buf; a.bytes = size; a.root = root; a.dtype = a.op = -1;
  a.key = std::string(_file) + "::" + std::to_string(_line) + "::" + _caller + "#" +
          std::to_string(size) + "@" + std::to_string(root);
  RunOp(&a);
}

void RobustEngine::Allgather(void *buf, size_t total_size, size_t slice_begin, size_t slice_end,
                             size_t size_prev_slice, const char *_file, const int _line,
                             const char *_caller) {
  OpArgs a{};
  a.kind = OpKind::kAllgather; a.buf = buf; a.bytes = total_size; a.dtype = a.op = -1;
  a.slice_begin = slice_begin; a.slice_end = slice_end; a.size_prev_slice = size_prev_slice;
  a.key = std::string(_file) + "::" + std::to_string(_line) + "::" + _caller + "#" +
          std::to_string(total_size);
  RunOp(&a);
}

// ---------------------------------------------------------------- checkpoints
void RobustEngine::LocalModelCheck(bool with_local) {
  if (use_local_model_ == -1) {
    use_local_model_ = with_local ? 1 : 0;
    if (with_local) {
      if (num_local_replica_ == 0) num_local_replica_ = 2;  // default ring replication
    } else {
      num_local_replica_ = 0;
    }
  } else {
    utils::Check(use_local_model_ == static_cast<int>(with_local),
                 "Can only call Checkpoint/LoadCheckPoint always with or without local_model, "
                 "but not mixed case");
  }
}

int RobustEngine::LoadCheckPoint(Serializable *global_model, Serializable *local_model) {
  checkpoint_loaded_ = true;
  if (comm_.world <= 1) return 0;
  LocalModelCheck(local_model != nullptr);
  const double t0 = utils::GetTime();
  if (!Consensus(nullptr, 0, proto::kLoadCheck, proto::kSeqSpecial)) {
    // every rank is asking: nothing exists yet
    if (comm_.cfg.debug) utils::HandleLogInfo("[%d] loadcheckpoint reset\n", comm_.rank);
    results_.Clear(); seq_ = 0; version_number_ = 0;
    // nothing of an earlier attempt may survive a fresh start (e.g. a checkpoint transfer that broke half-way)
    global_blob_.clear(); lazy_model_ = nullptr;
    for (int g = 0; g < 2; ++g) { local_[g].clear(); local_has_[g].clear(); }
    return 0;
  }
  std::vector<std::string> &slot = local_[local_gen_];
  const bool have_local = !local_has_[local_gen_].empty() && local_has_[local_gen_][0];
  if (local_model != nullptr && have_local) {
    utils::MemoryFixSizeBuffer fs(BeginPtr(slot[0]), slot[0].size());
    local_model->Load(&fs);
  }
  results_.Clear(); seq_ = 0;
  if (global_blob_.empty()) {
    version_number_ = 0;
  } else {
    utils::MemoryBufferStream fs(&global_blob_);
    utils::Assert(fs.Read(&version_number_, sizeof(version_number_)) == sizeof(version_number_),
                  "checkpoint blob too short");
    global_model->Load(&fs);
    utils::Assert(local_model == nullptr || have_local, "local model missing from a version %d checkpoint",
                  version_number_);
  }
  // survivors stay parked until we acknowledge, so that they cannot run ahead and drop state
  utils::Assert(Consensus(nullptr, 0, proto::kLoadAck, proto::kSeqSpecial), "load ack");
  if (comm_.cfg.debug) {
    utils::HandleLogInfo("[%d] loadcheckpoint size %lu finished version %d, seq %d, take %f seconds\n",
                         comm_.rank, static_cast<unsigned long>(global_blob_.size()),  // NOLINT(*)
                         version_number_, seq_, utils::GetTime() - t0);
  }
  return version_number_;
}

void RobustEngine::Commit(const Serializable *global_model, const Serializable *local_model, bool lazy) {
  if (comm_.world <= 1) {
    // nothing to replicate; with rabit_device_force the result cache still fills up, so empty it
    ++version_number_; results_.Clear(); seq_ = 0;
    return;
  }
  checkpoint_loaded_ = true;  // a commit also ends the bootstrap phase
  const double t0 = utils::GetTime();
  LocalModelCheck(local_model != nullptr);
  const int next_gen = !local_gen_;
  if (num_local_replica_ != 0) {
    local_[next_gen].assign(1, std::string());
    local_has_[next_gen].assign(1, 1);
    if (local_model != nullptr) {
      utils::MemoryBufferStream fs(&local_[next_gen][0]);
      local_model->Save(&fs);
    }
    // replicate to the next k ranks; a peer that is already further along completes us instead
    while (true) {
      if (Consensus(nullptr, 0, 0, proto::kSeqLocalCheckPoint)) break;
      if (Healthy(CompleteLocal(next_gen))) break;
    }
    Consensus(nullptr, 0, 0, proto::kSeqLocalAck);
  }
  // barrier 1: everybody has finished the version's collectives
  at_barrier1_ = true;
  utils::Assert(Consensus(nullptr, 0, proto::kCheckPoint, proto::kSeqSpecial), "commit barrier 1");
  at_barrier1_ = false;
  // ---- the commit itself: purely local, identical on all ranks ----
  ++version_number_;
  if (num_local_replica_ != 0) local_gen_ = next_gen;
  if (lazy) {
    lazy_model_ = global_model;
  } else {
    lazy_model_ = nullptr;
    global_blob_.resize(0);
    utils::MemoryBufferStream fs(&global_blob_);
    fs.Write(&version_number_, sizeof(version_number_));
    global_model->Save(&fs);
  }
  results_.Clear(); seq_ = 0;
  if (comm_.cfg.debug) {
    utils::HandleLogInfo("[%d] checkpoint finished version %d, take %f seconds\n", comm_.rank,
                         version_number_, utils::GetTime() - t0);
  }
  // barrier 2: nobody starts the next version before everyone holds the new state
  utils::Assert(Consensus(nullptr, 0, proto::kCheckAck, proto::kSeqSpecial), "commit barrier 2");
}

}  // namespace engine
}  // namespace rabit
