/*!
 * rabit-b200: RobustEngine data movers — what every rank runs once proto::Decide has said
 * "somebody needs bytes". Each begins with GatherTable so that all ranks share one view of
 * who holds / who needs, then executes the same locally computed ring plan.
 *
 * Reference counterparts: TryGetResult (allreduce_robust.cc:1103-1140), TryLoadCheckPoint
 * (1037-1088), TryRestoreCache (980-1024), TryRecoverLocalState / TryCheckinLocalState
 * (1328-1511). The reference negotiates routes with two tree message-passing rounds
 * (TryDecideRouting 798-845); here the table makes the plan explicit instead.
 */
#include "../device/device_backend.h"
#include "engine_core.h"

namespace rabit {
namespace engine {

Status RobustEngine::ServeResult(void *buf, size_t size, uint32_t seq, bool requester) {
  if (seq == proto::kSeqLocalAck) return Status::kOk;  // nothing to hand over, just let them pass
  if (seq == proto::kSeqLocalCheckPoint) return CompleteLocal(!local_gen_);
  ResultStore::Entry *held = requester ? nullptr : results_.Find(static_cast<int>(seq));
  uint64_t row[3] = {held != nullptr ? 1u : 0u, requester ? 1u : 0u,
                     held != nullptr ? held->bytes : (requester ? size : 0)};
  std::vector<uint64_t> tab;
  Status st = GatherTable(row, 3, &tab);
  if (!Ok(st)) return st;
  const int n = comm_.world;
  std::vector<char> has(n), need(n);
  size_t held_size = 0; bool any = false;
  for (int r = 0; r < n; ++r) {
    has[r] = tab[r * 3] != 0; need[r] = tab[r * 3 + 1] != 0;
    if (has[r]) { any = true; held_size = tab[r * 3 + 2]; }
  }
  utils::Check(any, "[%d] too many nodes went down: no surviving copy of the result of op %u "
               "(version %d)", comm_.rank, seq, version_number_);
  for (int r = 0; r < n; ++r) {
    utils::Check(!(need[r] || has[r]) || tab[r * 3 + 2] == held_size,
                 "Allreduce Recovered data size do not match the specification of function call.\n"
                 "Please check if calling sequence of recovered program is the same the original one "
                 "in current VersionNumber (rank %d expects %lu bytes, cached %lu)", r,
                 static_cast<unsigned long>(tab[r * 3 + 2]), static_cast<unsigned long>(held_size));  // NOLINT(*)
  }
  // A requester lands the bytes next to its buffer and takes them over only when all of them are there. Its buffer
  // still holds its own INPUT for this op: if the last holder dies half-way and comes back without the result, the
  // op is executed again by everybody, from that input. (The reference receives straight into the caller's buffer,
  // allreduce_robust.cc:1103-1140, and would then reduce a half-overwritten buffer.)
  HostBytes landing;
  if (requester) landing.resize(held_size);
  Status dst = Status::kOk;
  {
    const void *src_dev = held != nullptr ? held->dev : nullptr;
    const void *src_host = (held != nullptr && held->dev == nullptr) ? held->host.data() : nullptr;
    if (DeviceDeliver(has, need, src_host, src_dev, requester ? landing.data() : nullptr, held_size, &dst)) {
      if (Ok(dst) && requester && held_size != 0) std::memcpy(buf, landing.data(), held_size);
      return dst;
    }
  }
  // only a holder whose successor actually pulls pays for a device->host copy of the entry
  void *p = requester ? landing.data() : nullptr;
  if (held != nullptr && !has[(comm_.rank + 1) % n]) p = results_.HostView(held);
  if (held != nullptr && p == nullptr) return Status::kOk;
  dst = RingDeliver(has, need, p, held_size);
  tcp_payload_bytes_ += p != nullptr ? held_size : 0;
  if (Ok(dst) && requester && held_size != 0) std::memcpy(buf, landing.data(), held_size);
  return dst;
}

bool RobustEngine::DeviceDeliver(const std::vector<char> &has, const std::vector<char> &need,
                                 const void *src_host, const void *src_dev, void *dst_host, size_t size,
                                 Status *st) {
  // below rabit_device_blob_min the TCP ring is as fast and needs no GPU round trip
  if (device_ == nullptr || !device_->attached() || size < comm_.cfg.device_blob_min || size == 0) return false;
  int holder = -1;
  for (int r = 0; r < comm_.world && holder < 0; ++r) if (has[r]) holder = r;
  if (holder < 0) return false;
  // ONE holder serves: its copy engine puts the bytes into its arena window, every requester pulls them over
  // NVLink through its mapping of that arena. Ranks that neither hold nor need the bytes only keep count.
  const double t0 = utils::GetTime();
  *st = device_->BlobDeliver(holder, need, src_host, src_dev, dst_host, size);
  nvlink_payload_bytes_ += (comm_.rank == holder || need[comm_.rank]) ? size : 0;
  if (comm_.cfg.debug && (comm_.rank == holder || need[comm_.rank])) {
    const double dt = utils::GetTime() - t0;
    utils::HandleLogInfo("[%d] recovery payload of %lu bytes moved over NVLink from rank %d (%s, %.3f ms, %.1f GB/s)\n",
                         comm_.rank, static_cast<unsigned long>(size), holder, StatusName(*st), dt * 1e3,  // NOLINT(*)
                         dt > 0 ? size / dt / 1e9 : 0.0);
  }
  return true;
}

void RobustEngine::MaterializeLazy() {
  if (lazy_model_ == nullptr) return;
  global_blob_.resize(0);
  utils::MemoryBufferStream fs(&global_blob_);
  fs.Write(&version_number_, sizeof(version_number_));
  lazy_model_->Save(&fs);
  lazy_model_ = nullptr;
}

Status RobustEngine::ServeLoad(bool requester) {
  if (num_local_replica_ != 0) {
    if (requester) { local_[local_gen_].clear(); local_has_[local_gen_].clear(); }
    Status st = CompleteLocal(local_gen_);
    if (!Ok(st)) return st;
  }
  // A rank waiting at commit barrier 1 has finished the version: under the LazyCheckPoint contract its model object
  // already holds the NEXT version, so it cannot serialise the current one any more. Ranks that are still inside
  // the version can (there is at least one, otherwise the commit would have gone first, see proto::Decide).
  const bool stale_lazy = lazy_model_ != nullptr && at_barrier1_;
  const bool holder = !requester && checkpoint_loaded_ && !stale_lazy;
  if (holder) MaterializeLazy();  // a lazy checkpoint is serialised the first time anyone asks
  uint64_t row[3] = {holder ? 1u : 0u, requester ? 1u : 0u, holder ? global_blob_.size() : 0};
  std::vector<uint64_t> tab;
  Status st = GatherTable(row, 3, &tab);
  if (!Ok(st)) return st;
  const int n = comm_.world;
  std::vector<char> has(n), need(n);
  size_t size = 0; bool any = false;
  for (int r = 0; r < n; ++r) {
    has[r] = tab[r * 3] != 0; need[r] = tab[r * 3 + 1] != 0;
    if (!has[r]) continue;
    // every holder sends `size` bytes out of its own copy: they had better be the same blob
    utils::Check(!any || size == tab[r * 3 + 2], "[%d] LoadCheckPoint: holders disagree on the checkpoint size "
                 "(%lu vs %lu bytes)", comm_.rank, static_cast<unsigned long>(size),  // NOLINT(*)
                 static_cast<unsigned long>(tab[r * 3 + 2]));                          // NOLINT(*)
    any = true; size = tab[r * 3 + 2];
  }
  utils::Check(any, "[%d] LoadCheckPoint: no surviving rank holds the global checkpoint", comm_.rank);
  if (size == 0) {  // nothing committed yet: version 0
    if (requester) global_blob_.clear();
    return Status::kOk;
  }
  // A requester installs the blob only once all of it has arrived: a transfer that breaks half-way must not
  // leave a torn checkpoint behind (this rank may be asked to serve "its" checkpoint a moment later).
  std::string incoming;
  if (requester) incoming.resize(size);
  Status dst = Status::kOk;
  if (!DeviceDeliver(has, need, holder ? global_blob_.data() : nullptr, nullptr, requester ? &incoming[0] : nullptr,
                     size, &dst)) {
    void *p = requester ? static_cast<void *>(&incoming[0]) : (holder ? static_cast<void *>(&global_blob_[0]) : nullptr);
    dst = RingDeliver(has, need, p, size);
    tcp_payload_bytes_ += p != nullptr ? size : 0;
  }
  if (Ok(dst) && requester) global_blob_.swap(incoming);
  return dst;
}

// bootstrap cache wire form: u64 n, then n x (u64 klen, key, u64 vlen, value)
std::string RobustEngine::BootSerialize() const {
  std::string out;
  utils::MemoryBufferStream fs(&out);
  uint64_t n = boot_cache_.size();
  fs.Write(&n, sizeof(n));
  for (const auto &kv : boot_cache_) {
    fs.Write(kv.first);
    fs.Write(kv.second);
  }
  return out;
}

void RobustEngine::BootDeserialize(const std::string &blob) {
  boot_cache_.clear();
  std::string copy(blob);
  utils::MemoryBufferStream fs(&copy);
  uint64_t n = 0;
  utils::Check(fs.Read(&n, sizeof(n)) == sizeof(n), "corrupt bootstrap cache");
  for (uint64_t i = 0; i < n; ++i) {
    std::string k, v;
    utils::Check(fs.Read(&k) && fs.Read(&v), "corrupt bootstrap cache entry");
    boot_cache_.emplace_back(std::move(k), std::move(v));
  }
}

bool RobustEngine::BootLookup(const std::string &key, void *buf, size_t size) const {
  for (const auto &kv : boot_cache_) {
    if (kv.first != key) continue;
    utils::Assert(kv.second.size() == size, "bootstrap cache entry %s has %lu bytes, caller expects %lu",
                  key.c_str(), static_cast<unsigned long>(kv.second.size()),  // NOLINT(*)
                  static_cast<unsigned long>(size));                          // NOLINT(*)
    std::memcpy(buf, kv.second.data(), size);
    return true;
  }
  return false;
}

Status RobustEngine::ServeCache(bool requester, uint32_t max_count) {
  const bool holder = !requester && boot_cache_.size() == max_count;
  std::string blob = holder ? BootSerialize() : std::string();
  uint64_t row[3] = {holder ? 1u : 0u, requester ? 1u : 0u, blob.size()};
  std::vector<uint64_t> tab;
  Status st = GatherTable(row, 3, &tab);
  if (!Ok(st)) return st;
  const int n = comm_.world;
  std::vector<char> has(n), need(n);
  size_t size = 0; bool any = false;
  for (int r = 0; r < n; ++r) {
    has[r] = tab[r * 3] != 0; need[r] = tab[r * 3 + 1] != 0;
    if (has[r]) { any = true; size = tab[r * 3 + 2]; }
  }
  utils::Check(any && size != 0, "[%d] bootstrap cache: nobody holds %u entries", comm_.rank, max_count);
  if (requester) blob.resize(size);
  if (!DeviceDeliver(has, need, holder ? blob.data() : nullptr, nullptr, requester ? &blob[0] : nullptr, size, &st)) {
    st = RingDeliver(has, need, (requester || holder) ? &blob[0] : nullptr, size);
    tcp_payload_bytes_ += (requester || holder) ? size : 0;
  }
  if (!Ok(st)) return st;
  if (requester) BootDeserialize(blob);
  return Status::kOk;
}

Status RobustEngine::CompleteLocal(int g) {
  // Goal: slot j of every rank r holds the blob of rank r-j for j = 0..k. A blob only ever
  // moves between ring neighbours: forward sweeps push it from its earliest holder towards
  // r+k, backward sweeps pull it back towards its owner.
  const int n = comm_.world;
  const int k = std::min(num_local_replica_, n - 1);
  std::vector<std::string> &slot = local_[g];
  std::vector<char> &mine = local_has_[g];
  slot.resize(k + 1); mine.resize(k + 1, 0);
  uint64_t mask = 0;
  for (int j = 0; j <= k; ++j) if (mine[j]) mask |= 1ull << j;
  std::vector<uint64_t> tab;
  Status st = GatherTable(&mask, 1, &tab);
  if (!Ok(st)) return st;
  auto bit = [&](int r, int j) { return (tab[((r % n) + n) % n] >> j) & 1ull; };
  auto set = [&](int r, int j) { tab[((r % n) + n) % n] |= 1ull << j; };
  const int me = comm_.rank;
  // one sweep step: every rank may hand slot `sj` to a neighbour where it becomes slot `dj`
  auto step = [&](int sj, int dj, bool forward) -> Status {
    const int to = forward ? me + 1 : me - 1, from = forward ? me - 1 : me + 1;
    const bool send = bit(me, sj) && !bit(to, dj);
    const bool recv = !bit(me, dj) && bit(from, sj);
    uint64_t out_len = send ? slot[sj].size() : 0, in_len = 0;
    Status s = comm_.RingShift(&out_len, send ? sizeof(out_len) : 0, &in_len, recv ? sizeof(in_len) : 0, forward);
    if (!Ok(s)) return s;
    if (recv) slot[dj].resize(in_len);
    // the blobs themselves go neighbour to neighbour over NVLink (copy engines on the side stream) when the arena
    // is up; each end decides from the length of ITS transfer, which both ends of that transfer know
    const bool dev_up = device_ != nullptr && device_->attached();
    const size_t blob_min = std::max<size_t>(comm_.cfg.device_blob_min, 1);
    const bool out_dev = dev_up && out_len >= blob_min, in_dev = dev_up && in_len >= blob_min;
    if (out_dev || in_dev) {
      s = device_->BlobShift(out_dev ? slot[sj].data() : nullptr, out_dev ? out_len : 0,
                             in_dev ? &slot[dj][0] : nullptr, in_dev ? in_len : 0, forward);
      if (!Ok(s)) return s;
      nvlink_payload_bytes_ += (out_dev ? out_len : 0) + (in_dev ? in_len : 0);
    }
    const size_t tcp_out = out_dev ? 0 : out_len, tcp_in = in_dev ? 0 : in_len;
    s = comm_.RingShift(tcp_out != 0 ? slot[sj].data() : nullptr, tcp_out, tcp_in != 0 ? &slot[dj][0] : nullptr, tcp_in, forward);
    if (!Ok(s)) return s;
    tcp_payload_bytes_ += tcp_out + tcp_in;
    if (recv) mine[dj] = 1;
    // replay the step for the whole table so that all ranks keep the same picture
    std::vector<int> gained;
    for (int r = 0; r < n; ++r) {
      const int f = forward ? r - 1 : r + 1;
      if (!bit(r, dj) && bit(f, sj)) gained.push_back(r);
    }
    for (int r : gained) set(r, dj);
    return Status::kOk;
  };
  for (int j = 1; j <= k; ++j) {
    st = step(j - 1, j, true);
    if (!Ok(st)) return st;
  }
  for (int j = k; j >= 1; --j) {
    st = step(j, j - 1, false);
    if (!Ok(st)) return st;
  }
  // all-or-nothing: either every rank is complete or nobody ever stored a local model
  const uint64_t full = (k + 1 >= 64) ? ~0ull : ((1ull << (k + 1)) - 1);
  bool all_full = true, all_empty = true;
  for (int r = 0; r < n; ++r) {
    all_full = all_full && (tab[r] & full) == full;
    all_empty = all_empty && (tab[r] & full) == 0;
  }
  utils::Check(all_full || all_empty,
               "LoadCheckPoint: too many nodes fails, cannot recover local state");
  if (comm_.cfg.debug) {
    utils::HandleLogInfo("[%d] local state complete (generation %d): payload bytes so far over NVLink %lu, over TCP %lu\n",
                         comm_.rank, g, static_cast<unsigned long>(nvlink_payload_bytes_),  // NOLINT(*)
                         static_cast<unsigned long>(tcp_payload_bytes_));                   // NOLINT(*)
  }
  return Status::kOk;
}

}  // namespace engine
}  // namespace rabit
