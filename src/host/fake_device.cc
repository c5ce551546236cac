/*!
 * rabit-b200: a DeviceBackend made of host memory and the TCP collectives (rabit_device=fake).
 *
 * It exists for the CPU test tier: everything the engines do AROUND the device kernels - routing large host
 * buffers through the backend, the fused-consensus fast path and its mismatch fall-back, the result cache with
 * backend-owned entries (keep copy, spill, host view), cutting an op that does not fit the arena into sequence
 * numbered pieces, recovery payloads and local-model replicas handed to the backend's transport, the op counters in
 * the consensus envelope - is exercised by the multi-process tests without a GPU, kills included. The contract is
 * the one of the CUDA backend (src/device/device_backend.h): an op either leaves the result in the caller's buffer
 * and in `keep`, or fails with a Status and leaves the caller's buffer as it was.
 * No counterpart in the reference.
 */
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../device/device_backend.h"
#include "engine_core.h"

namespace rabit {
namespace engine {
namespace {

template <typename T>
void ReduceTyped(const void *src_, void *dst_, int n, int op) {
  const T *src = static_cast<const T *>(src_);
  T *dst = static_cast<T *>(dst_);
  for (int i = 0; i < n; ++i) {
    switch (op) {
      case 0: if (dst[i] < src[i]) dst[i] = src[i]; break;
      case 1: if (src[i] < dst[i]) dst[i] = src[i]; break;
      case 2: dst[i] = static_cast<T>(dst[i] + src[i]); break;
      default: break;
    }
  }
}
template <typename T>
void ReduceOr(const void *src_, void *dst_, int n) {
  const T *src = static_cast<const T *>(src_);
  T *dst = static_cast<T *>(dst_);
  for (int i = 0; i < n; ++i) dst[i] = static_cast<T>(dst[i] | src[i]);
}

thread_local int tl_dtype = 0, tl_op = 0;   // the reduce function of the TCP collectives carries no context

void TypedReducer(const void *src, void *dst, int n, const MPI::Datatype &) {
  const int op = tl_op;
  switch (tl_dtype) {
    case 0: if (op == 3) ReduceOr<signed char>(src, dst, n); else ReduceTyped<signed char>(src, dst, n, op); break;
    case 1: if (op == 3) ReduceOr<unsigned char>(src, dst, n); else ReduceTyped<unsigned char>(src, dst, n, op); break;
    case 2: if (op == 3) ReduceOr<int>(src, dst, n); else ReduceTyped<int>(src, dst, n, op); break;
    case 3: if (op == 3) ReduceOr<unsigned>(src, dst, n); else ReduceTyped<unsigned>(src, dst, n, op); break;
    case 4: case 8: if (op == 3) ReduceOr<long long>(src, dst, n); else ReduceTyped<long long>(src, dst, n, op); break;  // NOLINT(*)
    case 5: case 9:
      if (op == 3) ReduceOr<unsigned long long>(src, dst, n); else ReduceTyped<unsigned long long>(src, dst, n, op);  // NOLINT(*)
      break;
    case 6: ReduceTyped<float>(src, dst, n, op); break;
    case 7: ReduceTyped<double>(src, dst, n, op); break;
    default: break;
  }
}

/*! \brief {min, max} of the ranks' consensus codes in one tiny allreduce */
void CodeRange(const void *src_, void *dst_, int n, const MPI::Datatype &) {
  const uint32_t *src = static_cast<const uint32_t *>(src_);
  uint32_t *dst = static_cast<uint32_t *>(dst_);
  for (int i = 0; i + 1 < n; i += 2) {
    dst[i] = std::min(dst[i], src[i]);
    dst[i + 1] = std::max(dst[i + 1], src[i + 1]);
  }
}

class FakeBackend : public DeviceBackend {
 public:
  explicit FakeBackend(const Config &cfg) : cfg_(cfg) {}
  ~FakeBackend() override { for (void *p : live_) free(p); }
  bool Attach(HostComm *comm) override {
    comm_ = comm;
    epoch_ = 0;
    attached_ = true;
    return true;
  }
  void Detach() override { attached_ = false; }
  bool attached() const override { return attached_; }

  Status Allreduce(void *ptr, size_t count, int dtype, int op, int, uint32_t code, void *keep, int flags) override {
    const size_t eb = DTypeBytes(dtype), bytes = eb * count;
    if (bytes == 0) return Status::kOk;
    utils::Check(bytes <= MaxOpBytes(bytes), "fake device: recoverable allreduce of %lu bytes exceeds the arena",
                 static_cast<unsigned long>(bytes));  // NOLINT(*)
    if ((flags & kOpConservative) != 0) ++conservative_ops_;
    Status st = Agree(code);
    if (!Ok(st)) return st;
    // out of place: the caller's buffer is only touched once the collective has succeeded
    std::vector<char> tmp(bytes);
    tl_dtype = dtype; tl_op = op;
    st = comm_->AllreduceFrom(ptr, tmp.data(), eb, count, TypedReducer);
    ++epoch_;
    ++stats_.kernel_launches;
    if (!Ok(st)) return st;
    std::memcpy(ptr, tmp.data(), bytes);
    if (keep != nullptr) std::memcpy(keep, tmp.data(), bytes);
    ++stats_.fused_consensus_ok;
    return Status::kOk;
  }
  Status Broadcast(void *ptr, size_t size, int root, uint32_t code, void *keep, int) override {
    if (size == 0) return Status::kOk;
    Status st = Agree(code);
    if (!Ok(st)) return st;
    std::vector<char> tmp(size);
    if (comm_->rank == root) std::memcpy(tmp.data(), ptr, size);
    st = comm_->Broadcast(tmp.data(), size, root);
    ++epoch_;
    ++stats_.kernel_launches;
    if (!Ok(st)) return st;
    std::memcpy(ptr, tmp.data(), size);
    if (keep != nullptr) std::memcpy(keep, tmp.data(), size);
    return Status::kOk;
  }
  Status Allgather(void *ptr, size_t total, size_t begin, size_t end, uint32_t code, void *keep) override {
    if (total == 0) return Status::kOk;
    Status st = Agree(code);
    if (!Ok(st)) return st;
    std::vector<char> tmp(total);
    std::memcpy(tmp.data() + begin, static_cast<char *>(ptr) + begin, end - begin);
    // the size of the predecessor's slice is not part of the device interface: ask for it
    unsigned long long mine = end - begin, prev = 0;  // NOLINT(*)
    st = comm_->RingShift(&mine, sizeof(mine), &prev, sizeof(prev), true);
    if (Ok(st)) st = comm_->AllgatherRing(tmp.data(), total, begin, end, prev);
    ++epoch_;
    ++stats_.kernel_launches;
    if (!Ok(st)) return st;
    std::memcpy(ptr, tmp.data(), total);
    if (keep != nullptr) std::memcpy(keep, tmp.data(), total);
    return Status::kOk;
  }
  size_t MaxOpBytes(size_t) override { return cfg_.device_arena != 0 ? cfg_.device_arena / 2 / 256 * 256 : (1UL << 40); }
  uint32_t OpEpoch() override { return epoch_; }
  uint64_t OpRounds() const override { return rounds_; }
  void AdoptEpoch(uint32_t epoch, uint64_t rounds) override { epoch_ = epoch; rounds_ = rounds; }

  Status BlobDeliver(int holder, const std::vector<char> &need, const void *src_host, const void *src_dev, void *dst_host,
                     size_t size) override {
    if (size == 0) return Status::kOk;
    // everybody calls this, so a broadcast from the holder will do; ranks that neither hold nor need drop the bytes
    std::vector<char> tmp(size);
    if (comm_->rank == holder) std::memcpy(tmp.data(), src_dev != nullptr ? src_dev : src_host, size);
    const Status st = comm_->Broadcast(tmp.data(), size, holder);
    if (Ok(st) && need[comm_->rank] && comm_->rank != holder && dst_host != nullptr) std::memcpy(dst_host, tmp.data(), size);
    stats_.blob_bytes += size;
    return st;
  }
  Status BlobShift(const void *out, size_t out_len, void *in, size_t in_len, bool forward) override {
    stats_.blob_bytes += out_len;
    return comm_->RingShift(out, out_len, in, in_len, forward);
  }

  void *Alloc(size_t nbytes) override {
    void *p = malloc(std::max<size_t>(nbytes, 1));
    utils::Check(p != nullptr, "fake device: out of memory");
    live_.push_back(p);
    return p;
  }
  void Free(void *p) override {
    auto it = std::find(live_.begin(), live_.end(), p);
    if (it == live_.end()) return;
    live_.erase(it);
    free(p);
  }
  void CopyD2H(void *dst, const void *src, size_t n) override { std::memcpy(dst, src, n); }
  void CopyH2D(void *dst, const void *src, size_t n) override { std::memcpy(dst, src, n); }
  void CopyD2D(void *dst, const void *src, size_t n) override { std::memcpy(dst, src, n); }
  void PrefetchH2D(void *dst, const void *src, size_t n, int) override { std::memcpy(dst, src, n); }
  void AwaitPrefetch(int) override {}
  void DrainD2H(void *dst, const void *src, size_t n) override { std::memcpy(dst, src, n); }
  void DrainSync() override {}
  const DeviceStats &stats() const override { return stats_; }
  std::string Describe() const override {
    return "device backend: fake (host memory + TCP collectives, tests only), conservative retries " +
           std::to_string(conservative_ops_);
  }

 private:
  /*!
   * \brief the fused consensus of the kernels' entry barrier: all codes equal, or everybody sees a mismatch.
   *  The op counters travel along: the flag protocols of the real kernels only work while they are equal on all
   *  ranks, so a difference here is an engine bug (it must re-align them through the consensus envelope).
   *  RABIT_FAKE_MISMATCH=rank,nth: that rank presents a wrong code in its nth op (once per process) and, like a
   *  rank that abandoned a multi-kernel op, comes out of it with its counter three ahead.
   */
  Status Agree(uint32_t code) {
    if (comm_->world <= 1) return Status::kOk;
    static const char *spec = getenv("RABIT_FAKE_MISMATCH");
    int bad_rank = -1, bad_op = -1;
    if (spec != nullptr && !injected_) sscanf(spec, "%d,%d", &bad_rank, &bad_op);
    const bool inject = bad_rank == comm_->rank && bad_op == static_cast<int>(ops_);
    ++ops_;
    uint32_t range[4] = {code ^ (inject ? 1u : 0u), code ^ (inject ? 1u : 0u), epoch_, epoch_};
    const Status st = comm_->Allreduce(range, sizeof(uint32_t), 4, CodeRange);
    if (!Ok(st)) return st;
    utils::Check(range[2] == range[3], "[%d] fake device: op counters differ between ranks (%u..%u): the engine did not "
                 "re-align them after an abandoned op", comm_->rank, range[2], range[3]);
    if (inject) { injected_ = true; epoch_ += 3; }
    if (range[0] != range[1]) { ++epoch_; ++stats_.fused_consensus_mismatch; return Status::kMismatch; }
    return Status::kOk;
  }
  Config cfg_;
  HostComm *comm_ = nullptr;
  bool attached_ = false;
  uint32_t epoch_ = 0;
  uint64_t rounds_ = 0;
  uint64_t ops_ = 0, conservative_ops_ = 0;
  bool injected_ = false;
  DeviceStats stats_;
  std::vector<void *> live_;
};

}  // namespace

DeviceBackend *CreateFakeBackend(const Config &cfg) { return new FakeBackend(cfg); }

}  // namespace engine
}  // namespace rabit
