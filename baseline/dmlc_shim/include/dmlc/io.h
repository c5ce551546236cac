/*!
 * Minimal dmlc/io.h stand-in for the baseline build. It has the API generation the reference
 * snapshot was written against (Stream::Write returns void; current dmlc-core returns size_t,
 * SURVEY.md 2.7 #11), which is why no line of the reference needs to be touched.
 */
#ifndef DMLC_IO_H_
#define DMLC_IO_H_
#include <cstring>
#include <string>
#include <vector>
#include "./base.h"
#include "./logging.h"
namespace dmlc {
class Stream {
 public:
  virtual size_t Read(void *ptr, size_t size) = 0;
  virtual void Write(const void *ptr, size_t size) = 0;
  virtual ~Stream(void) {}
  template <typename T>
  inline void Write(const std::vector<T> &vec) {
    uint64_t n = vec.size();
    this->Write(static_cast<const void *>(&n), sizeof(n));
    if (n != 0) this->Write(static_cast<const void *>(&vec[0]), sizeof(T) * n);
  }
  template <typename T>
  inline bool Read(std::vector<T> *vec) {
    uint64_t n;
    if (this->Read(static_cast<void *>(&n), sizeof(n)) != sizeof(n)) return false;
    vec->resize(n);
    if (n == 0) return true;
    return this->Read(static_cast<void *>(&(*vec)[0]), sizeof(T) * n) == sizeof(T) * n;
  }
  inline void Write(const std::string &s) {
    uint64_t n = s.length();
    this->Write(static_cast<const void *>(&n), sizeof(n));
    if (n != 0) this->Write(static_cast<const void *>(s.data()), n);
  }
  inline bool Read(std::string *s) {
    uint64_t n;
    if (this->Read(static_cast<void *>(&n), sizeof(n)) != sizeof(n)) return false;
    s->resize(n);
    if (n == 0) return true;
    return this->Read(static_cast<void *>(&(*s)[0]), n) == n;
  }
};
class SeekStream : public Stream {
 public:
  virtual ~SeekStream(void) {}
  virtual void Seek(size_t pos) = 0;
  virtual size_t Tell(void) = 0;
};
class Serializable {
 public:
  virtual ~Serializable() {}
  virtual void Load(Stream *fi) = 0;
  virtual void Save(Stream *fo) const = 0;
};
}  // namespace dmlc
#endif  // DMLC_IO_H_
