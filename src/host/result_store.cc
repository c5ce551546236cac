/*!
 * rabit-b200: ResultStore — the result cache that makes replay after a restart possible
 * (role of reference ResultBuffer, allreduce_robust.h:324-388, extended with GPU-resident
 * entries so that caching a device collective costs one fused D2D write, not a PCIe trip).
 */
#include "../device/device_backend.h"
#include "engine_core.h"

namespace rabit {
namespace engine {

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
  temp_.host.assign(nbytes, '\0');
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
