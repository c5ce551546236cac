/*!
 * rabit-b200: shared-memory data plane for links between workers of the same host.
 *
 * The reference moves every byte through TCP sockets, also between two workers of one machine
 * (src/allreduce_base.cc:476-946 all sit on TCPSocket::Send/Recv). Here a link whose two ends
 * share a host keeps its TCP connection for rendezvous, failure detection (hang-up) and the `bye`
 * byte, but the payload goes through one single-producer / single-consumer byte ring per direction:
 * an anonymous memfd created by the sender and opened by the receiver through /proc/<pid>/fd/<n>,
 * so nothing has a name that could outlive a killed worker. Blocking is done on a per-process
 * doorbell (an abstract-namespace unix datagram socket) that sits in the same poll set as the
 * sockets, so the byte-flow state machines in comm.cc do not change.
 */
#ifndef RABIT_SRC_HOST_SHM_RING_H_
#define RABIT_SRC_HOST_SHM_RING_H_

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <sys/syscall.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

namespace rabit {
namespace engine {

/*! \brief one direction of a same-host link */
class ShmRing {
 public:
  struct Header {
    alignas(64) std::atomic<uint64_t> head;              // bytes ever produced
    alignas(64) std::atomic<uint64_t> tail;              // bytes ever consumed
    alignas(64) std::atomic<uint32_t> consumer_waiting;  // the consumer sleeps (or is about to) on its doorbell
    std::atomic<uint32_t> producer_waiting;              // the producer does
    uint64_t capacity;
    uint32_t magic;
  };
  static const uint32_t kMagic = 0x52534d31;  // "RSM1"
  static const size_t kDataOffset = 4096;

  bool Active() const { return hdr_ != nullptr; }
  int fd() const { return fd_; }
  size_t capacity() const { return cap_; }

  /*! \brief producer side: a fresh anonymous segment */
  bool Create(size_t capacity) {
    Close();
    fd_ = static_cast<int>(syscall(SYS_memfd_create, "rabit-ring", 1u /* MFD_CLOEXEC */));
    if (fd_ < 0) return false;
    if (ftruncate(fd_, static_cast<off_t>(kDataOffset + capacity)) != 0 || !Map(kDataOffset + capacity)) { Close(); return false; }
    hdr_->capacity = capacity;
    hdr_->magic = kMagic;
    cap_ = capacity;
    return true;
  }
  /*! \brief consumer side: the segment behind file descriptor remote_fd of process pid */
  bool Open(int pid, int remote_fd, size_t capacity) {
    Close();
    char path[64];
    snprintf(path, sizeof(path), "/proc/%d/fd/%d", pid, remote_fd);
    fd_ = open(path, O_RDWR | O_CLOEXEC);
    if (fd_ < 0) return false;
    struct stat st;
    if (fstat(fd_, &st) != 0 || static_cast<size_t>(st.st_size) != kDataOffset + capacity || !Map(kDataOffset + capacity) ||
        hdr_->magic != kMagic || hdr_->capacity != capacity) {
      Close();
      return false;
    }
    cap_ = capacity;
    return true;
  }
  void Close() {
    if (base_ != nullptr) munmap(base_, map_bytes_);
    if (fd_ >= 0) ::close(fd_);
    base_ = nullptr; hdr_ = nullptr; data_ = nullptr; fd_ = -1; cap_ = 0; map_bytes_ = 0;
  }

  /*! \brief bytes the consumer can take now */
  size_t Avail() const {
    return static_cast<size_t>(hdr_->head.load(std::memory_order_seq_cst) - hdr_->tail.load(std::memory_order_relaxed));
  }
  /*! \brief bytes the producer can put now */
  size_t Space() const {
    return cap_ - static_cast<size_t>(hdr_->head.load(std::memory_order_relaxed) - hdr_->tail.load(std::memory_order_seq_cst));
  }
  /*! \brief a producer only starts a piece when this much is free: keeps the copies large when the consumer is the slower side */
  size_t SendThreshold() const { return cap_ / 4; }

  /*! \brief producer: copy in as much of [p, p + n) as fits; returns the bytes taken */
  size_t Write(const char *p, size_t n) {
    const uint64_t head = hdr_->head.load(std::memory_order_relaxed);
    const size_t space = cap_ - static_cast<size_t>(head - hdr_->tail.load(std::memory_order_acquire));
    n = std::min(n, space);
    if (n == 0) return 0;
    const size_t off = static_cast<size_t>(head % cap_);
    const size_t first = std::min(n, cap_ - off);
    std::memcpy(data_ + off, p, first);
    std::memcpy(data_, p + first, n - first);
    hdr_->head.store(head + n, std::memory_order_seq_cst);
    return n;
  }
  /*! \brief consumer: copy out up to n bytes; returns the bytes delivered */
  size_t Read(char *p, size_t n) {
    const uint64_t tail = hdr_->tail.load(std::memory_order_relaxed);
    const size_t avail = static_cast<size_t>(hdr_->head.load(std::memory_order_acquire) - tail);
    n = std::min(n, avail);
    if (n == 0) return 0;
    const size_t off = static_cast<size_t>(tail % cap_);
    const size_t first = std::min(n, cap_ - off);
    std::memcpy(p, data_ + off, first);
    std::memcpy(p + first, data_, n - first);
    hdr_->tail.store(tail + n, std::memory_order_seq_cst);
    return n;
  }
  /*! \brief consumer: the unread bytes that lie contiguously in the mapping (zero-copy read); 0 = nothing there */
  size_t Peek(const char **p) const {
    const uint64_t tail = hdr_->tail.load(std::memory_order_relaxed);
    const size_t avail = static_cast<size_t>(hdr_->head.load(std::memory_order_acquire) - tail);
    const size_t off = static_cast<size_t>(tail % cap_);
    *p = data_ + off;
    return std::min(avail, cap_ - off);
  }
  /*! \brief consumer: done with the first n bytes handed out by Peek */
  void Consume(size_t n) {
    hdr_->tail.store(hdr_->tail.load(std::memory_order_relaxed) + n, std::memory_order_seq_cst);
  }
  /*! \brief the other side said it is (about to be) asleep: take the note, the caller rings its doorbell */
  bool TakeConsumerWaiting() {
    return hdr_->consumer_waiting.load(std::memory_order_seq_cst) != 0 && hdr_->consumer_waiting.exchange(0) != 0;
  }
  bool TakeProducerWaiting() {
    return hdr_->producer_waiting.load(std::memory_order_seq_cst) != 0 && hdr_->producer_waiting.exchange(0) != 0;
  }
  /*! \brief "I am about to sleep": the caller MUST look at Avail() / Space() once more afterwards. Note, look (here) and
   *  publish, read the note (other side) are all seq_cst: either that look sees the new bytes, or the other side sees
   *  the note and rings the doorbell. */
  void SetConsumerWaiting(uint32_t v) { hdr_->consumer_waiting.store(v, std::memory_order_seq_cst); }
  void SetProducerWaiting(uint32_t v) { hdr_->producer_waiting.store(v, std::memory_order_seq_cst); }

 private:
  bool Map(size_t bytes) {
    void *m = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd_, 0);
    if (m == MAP_FAILED) return false;
    base_ = m;
    map_bytes_ = bytes;
    hdr_ = static_cast<Header *>(m);
    data_ = static_cast<char *>(m) + kDataOffset;
    return true;
  }
  void *base_ = nullptr;
  Header *hdr_ = nullptr;
  char *data_ = nullptr;
  size_t cap_ = 0, map_bytes_ = 0;
  int fd_ = -1;
};

/*! \brief identity of "this host as seen by this process": two workers may share memory and abstract unix sockets only
 *  if boot, pid / net / ipc namespaces and user agree */
inline void HostToken(uint64_t out[2]) {
  std::string s;
  char buf[256];
  if (FILE *f = fopen("/proc/sys/kernel/random/boot_id", "r")) {
    if (fgets(buf, sizeof(buf), f) != nullptr) s += buf;
    fclose(f);
  }
  for (const char *ns : {"/proc/self/ns/pid", "/proc/self/ns/net", "/proc/self/ns/ipc", "/proc/self/ns/mnt"}) {
    const ssize_t n = readlink(ns, buf, sizeof(buf) - 1);
    if (n > 0) { buf[n] = '\0'; s += buf; }
    s += '|';
  }
  s += std::to_string(static_cast<unsigned>(geteuid()));
  uint64_t h0 = 1469598103934665603ull, h1 = 0x9e3779b97f4a7c15ull;
  for (unsigned char c : s) {
    h0 = (h0 ^ c) * 1099511628211ull;
    h1 = (h1 + c) * 0xff51afd7ed558ccdull; h1 ^= h1 >> 29;
  }
  out[0] = h0; out[1] = h1;
}

}  // namespace engine
}  // namespace rabit
#endif  // RABIT_SRC_HOST_SHM_RING_H_
