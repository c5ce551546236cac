/*!
 * rabit-b200: stream + serializable contracts for checkpointed models.
 *
 * The reference borrows these from dmlc-core (include/rabit/serializable.h:12-23 typedefs
 * dmlc::Stream / dmlc::Serializable). This build is standalone, so the contracts are
 * defined here with the same call shapes and the same wire encoding for the helpers
 * (std::vector<POD> and std::string are written as `uint64 count` + raw bytes, the
 * "checkpoint format" described in SURVEY.md section 5.4).
 */
#ifndef RABIT_SERIALIZABLE_H_
#define RABIT_SERIALIZABLE_H_

#include <cstdint>
#include <cstring>
#include <string>
#include <type_traits>
#include <vector>

#include "rabit/base.h"

namespace rabit {

/*! \brief byte stream; Read returns the number of bytes actually delivered */
class Stream {
 public:
  virtual ~Stream() {}
  virtual size_t Read(void *ptr, size_t size) = 0;
  virtual void Write(const void *ptr, size_t size) = 0;

  // ---- typed helpers (encoding compatible with dmlc's POD handlers) ----
  template <typename T>
  typename std::enable_if<std::is_trivially_copyable<T>::value>::type
  Write(const T &v) { this->Write(static_cast<const void *>(&v), sizeof(T)); }

  template <typename T>
  typename std::enable_if<std::is_trivially_copyable<T>::value, bool>::type
  Read(T *v) { return this->Read(static_cast<void *>(v), sizeof(T)) == sizeof(T); }

  template <typename T>
  void Write(const std::vector<T> &vec) {
    static_assert(std::is_trivially_copyable<T>::value, "vector payload must be POD");
    uint64_t n = vec.size();
    this->Write(static_cast<const void *>(&n), sizeof(n));
    if (n != 0) this->Write(static_cast<const void *>(vec.data()), sizeof(T) * n);
  }
  template <typename T>
  bool Read(std::vector<T> *vec) {
    static_assert(std::is_trivially_copyable<T>::value, "vector payload must be POD");
    uint64_t n = 0;
    if (this->Read(static_cast<void *>(&n), sizeof(n)) != sizeof(n)) return false;
    vec->resize(n);
    if (n == 0) return true;
    return this->Read(static_cast<void *>(vec->data()), sizeof(T) * n) == sizeof(T) * n;
  }
  void Write(const std::string &s) {
    uint64_t n = s.length();
    this->Write(static_cast<const void *>(&n), sizeof(n));
    if (n != 0) this->Write(static_cast<const void *>(s.data()), n);
  }
  bool Read(std::string *s) {
    uint64_t n = 0;
    if (this->Read(static_cast<void *>(&n), sizeof(n)) != sizeof(n)) return false;
    s->resize(n);
    if (n == 0) return true;
    return this->Read(static_cast<void *>(&(*s)[0]), n) == n;
  }
};

/*! \brief stream with random access */
class SeekStream : public Stream {
 public:
  virtual void Seek(size_t pos) = 0;
  virtual size_t Tell(void) = 0;
};

/*! \brief anything that can be checkpointed: Load must accept what Save wrote */
class Serializable {
 public:
  virtual ~Serializable() {}
  virtual void Load(Stream *fi) = 0;
  virtual void Save(Stream *fo) const = 0;
};

}  // namespace rabit
#endif  // RABIT_SERIALIZABLE_H_
