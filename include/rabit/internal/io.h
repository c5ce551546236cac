/*!
 * rabit-b200: in-memory streams used for checkpoints and (de)serialising reducers.
 * Parity: reference include/rabit/internal/io.h:24-111 (MemoryFixSizeBuffer over a fixed
 * region, MemoryBufferStream over a growing std::string). kSeekEnd positions at the end
 * of the region, which is what reference test/cpp/test_io.cc:10-17 exercises.
 */
#ifndef RABIT_INTERNAL_IO_H_
#define RABIT_INTERNAL_IO_H_

#include <algorithm>
#include <cstring>
#include <limits>
#include <string>

#include "rabit/internal/utils.h"
#include "rabit/serializable.h"

namespace rabit {
namespace utils {
/*! \brief sentinel accepted by Seek(): jump to the end of the buffer */
static const size_t kSeekEnd = std::numeric_limits<size_t>::max();
typedef ::rabit::SeekStream SeekStream;

/*! \brief read/write view over caller-owned memory of fixed size */
class MemoryFixSizeBuffer : public SeekStream {
 public:
  MemoryFixSizeBuffer(void *base, size_t nbytes)
      : base_(static_cast<char *>(base)), cap_(nbytes), pos_(0) {}
  size_t Read(void *ptr, size_t size) override {
    const size_t n = std::min(size, cap_ - pos_);
    if (n != 0) std::memcpy(ptr, base_ + pos_, n);
    pos_ += n;
    return n;
  }
  void Write(const void *ptr, size_t size) override {
    if (size == 0) return;
    Assert(pos_ + size <= cap_, "MemoryFixSizeBuffer: write of %lu bytes overflows region of %lu",
           static_cast<unsigned long>(size), static_cast<unsigned long>(cap_));  // NOLINT(*)
    std::memcpy(base_ + pos_, ptr, size);
    pos_ += size;
  }
  void Seek(size_t pos) override { pos_ = (pos == kSeekEnd) ? cap_ : pos; }
  size_t Tell(void) override { return pos_; }
  bool AtEnd(void) const { return pos_ == cap_; }
  using Stream::Read;
  using Stream::Write;

 private:
  char *base_;
  size_t cap_;
  size_t pos_;
};

/*! \brief stream over a std::string that grows on write */
class MemoryBufferStream : public SeekStream {
 public:
  explicit MemoryBufferStream(std::string *store) : store_(store), pos_(0) {}
  size_t Read(void *ptr, size_t size) override {
    Assert(pos_ <= store_->length(), "MemoryBufferStream: cursor past end");
    const size_t n = std::min(size, store_->length() - pos_);
    if (n != 0) std::memcpy(ptr, store_->data() + pos_, n);
    pos_ += n;
    return n;
  }
  void Write(const void *ptr, size_t size) override {
    if (size == 0) return;
    if (pos_ + size > store_->length()) store_->resize(pos_ + size);
    std::memcpy(&(*store_)[pos_], ptr, size);
    pos_ += size;
  }
  void Seek(size_t pos) override { pos_ = (pos == kSeekEnd) ? store_->length() : pos; }
  size_t Tell(void) override { return pos_; }
  bool AtEnd(void) const { return pos_ == store_->length(); }
  using Stream::Read;
  using Stream::Write;

 private:
  std::string *store_;
  size_t pos_;
};
}  // namespace utils
}  // namespace rabit
#endif  // RABIT_INTERNAL_IO_H_
