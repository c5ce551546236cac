/*!
 * rabit-b200: ResultStore — the result cache that makes replay after a restart possible
 * (role of reference ResultBuffer, allreduce_robust.h:324-388, extended with GPU-resident
 * entries so that caching a device collective costs one fused D2D write, not a PCIe trip).
 */
#include "../device/device_backend.h"
#include "engine_core.h"

#include <sys/mman.h>

#include <cstdlib>
#include <map>
#include <vector>

namespace rabit {
namespace engine {

namespace {
const size_t kPoolMinBlock = 64UL << 10;    // smaller blocks: plain malloc is fine
const size_t kPoolMaxBytes = 1UL << 30;     // idle bytes kept per thread
const size_t kHuge = 2UL << 20;              // transparent huge page
const size_t kSlabBytes = 32UL << 20;       // medium blocks are carved from slabs of this size
struct BlockPool {
  std::multimap<size_t, char *> idle;       // capacity -> block
  size_t idle_bytes = 0;
  // Results that stay cached (no checkpoint yet) cannot be recycled, so every op touches fresh pages.
  // Medium blocks come out of 2 MiB-aligned slabs advised as huge pages: one fault per 2 MiB, not per 4 KiB.
  // Slab memory is never returned to the system, only to `idle`.
  std::vector<char *> slabs;
  char *slab_cur = nullptr;
  size_t slab_left = 0;
  bool FromSlab(const char *p) const {
    for (const char *s : slabs) if (p >= s && p < s + kSlabBytes) return true;
    return false;
  }
  char *Carve(size_t n) {
    n = (n + 255) & ~static_cast<size_t>(255);
    if (n > slab_left) {
      void *fresh = nullptr;
      if (posix_memalign(&fresh, kHuge, kSlabBytes) != 0) return nullptr;
      madvise(fresh, kSlabBytes, MADV_HUGEPAGE);
      slabs.push_back(static_cast<char *>(fresh));
      slab_cur = static_cast<char *>(fresh);
      slab_left = kSlabBytes;
    }
    char *p = slab_cur;
    slab_cur += n; slab_left -= n;
    return p;
  }
  ~BlockPool() {
    for (auto &kv : idle) if (!FromSlab(kv.second)) std::free(kv.second);
    for (char *s : slabs) std::free(s);
  }
};
BlockPool &Pool() {
  // never destroyed: cache entries of a thread-local engine may be released after any thread_local of this
  // file would already be gone; the blocks go back to the system with the process
  static thread_local BlockPool *pool = new BlockPool();
  return *pool;
}
}  // namespace

void HostBytes::resize(size_t n) {
  if (n <= cap_) { size_ = n; return; }
  clear();
  if (n == 0) return;
  if (n >= kPoolMinBlock) {
    BlockPool &pool = Pool();
    auto it = pool.idle.lower_bound(n);
    if (it != pool.idle.end() && it->first <= n + n / 2) {   // close enough in size: reuse, still mapped
      p_ = it->second; cap_ = it->first; size_ = n;
      pool.idle_bytes -= cap_;
      pool.idle.erase(it);
      return;
    }
  }
  void *fresh = nullptr;
  // (slab memory is never returned to the system: once a lot of it sits idle in sizes nobody asks for any more,
  // stop carving and let malloc / free handle the odd sizes)
  if (n >= kPoolMinBlock && n < 2 * kHuge && Pool().idle_bytes <= kPoolMaxBytes && (p_ = Pool().Carve(n)) != nullptr) {
    size_ = n; cap_ = (n + 255) & ~static_cast<size_t>(255);
    return;
  }
  if (n >= 2 * kHuge && posix_memalign(&fresh, kHuge, n) == 0) {
    // a block this size is touched end to end right away: 2 MiB pages cut the page-fault bill of a fresh
    // block roughly in half where transparent huge pages are available (a no-op elsewhere)
    madvise(fresh, n, MADV_HUGEPAGE);
    p_ = static_cast<char *>(fresh);
  } else {
    p_ = static_cast<char *>(std::malloc(n));
  }
  utils::Check(p_ != nullptr, "out of memory allocating %lu bytes for the result cache", static_cast<unsigned long>(n));  // NOLINT(*)
  size_ = cap_ = n;
}

void HostBytes::clear() {
  if (p_ == nullptr) return;
  BlockPool &pool = Pool();
  if (cap_ >= kPoolMinBlock && (pool.idle_bytes + cap_ <= kPoolMaxBytes || pool.FromSlab(p_))) {
    pool.idle.emplace(cap_, p_);
    pool.idle_bytes += cap_;
  } else {
    std::free(p_);
  }
  p_ = nullptr; size_ = cap_ = 0;
}

void ResultStore::Release(Entry *e) {
  if (e->dev != nullptr && device_ != nullptr) {
    device_->Free(e->dev);
    device_bytes_ -= e->bytes;
  }
  e->dev = nullptr;
  e->host.clear();
}

char *ResultStore::AllocTemp(size_t nbytes) {
  DiscardTemp();
  temp_.bytes = nbytes;
  temp_.host.resize(nbytes);   // uninitialised: every caller overwrites all of it
  return nbytes != 0 ? &temp_.host[0] : nullptr;
}

void *ResultStore::AllocTempDevice(size_t nbytes) {
  DiscardTemp();
  utils::Assert(device_ != nullptr, "ResultStore: device entry without a device backend");
  temp_.bytes = nbytes;
  temp_.dev = device_->Alloc(nbytes);
  device_bytes_ += nbytes;
  return temp_.dev;
}

void ResultStore::DiscardTemp() {
  Release(&temp_);
  temp_ = Entry();
}

void ResultStore::Commit(int seq) {
  temp_.seq = seq;
  entries_.push_back(std::move(temp_));
  temp_ = Entry();
}

void ResultStore::Clear() {
  for (Entry &e : entries_) Release(&e);
  entries_.clear();
  DiscardTemp();
}

void ResultStore::DropLast() {
  if (entries_.empty()) return;
  Release(&entries_.back());
  entries_.pop_back();
}

ResultStore::Entry *ResultStore::Find(int seq) {
  for (Entry &e : entries_) if (e.seq == seq) return &e;
  return nullptr;
}

char *ResultStore::HostView(Entry *e) {
  if (e->dev != nullptr && e->host.size() != e->bytes) {
    e->host.resize(e->bytes);
    device_->CopyD2H(&e->host[0], e->dev, e->bytes);
  }
  return e->bytes != 0 ? &e->host[0] : nullptr;
}

void ResultStore::SpillTo(size_t budget) {
  for (Entry &e : entries_) {
    if (device_bytes_ <= budget) break;
    if (e.dev == nullptr) continue;
    HostView(&e);
    device_->Free(e.dev);
    e.dev = nullptr;
    device_bytes_ -= e.bytes;
  }
}

}  // namespace engine
}  // namespace rabit
