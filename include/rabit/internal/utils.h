/*!
 * rabit-b200: error handling, logging and small helpers.
 *
 * Behavioural parity with reference include/rabit/internal/utils.h:
 *   Assert/Check/Error format a message and throw (utils.h:148-181); Printf and
 *   HandleLogInfo print to stdout (104-112); StringToBool accepts "true" or a non-zero
 *   integer (78-80); BeginPtr returns NULL for empty containers (194-217).
 * The exception type is rabit::Error (the reference throws dmlc::Error; an alias is kept
 * in namespace dmlc so `catch (dmlc::Error&)` in client code still compiles).
 */
#ifndef RABIT_INTERNAL_UTILS_H_
#define RABIT_INTERNAL_UTILS_H_

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <strings.h>
#include <vector>

#include "rabit/base.h"

namespace rabit {
/*! \brief every failed Check/Assert/Error in the library surfaces as this type */
struct Error : public std::runtime_error {
  explicit Error(const std::string &s) : std::runtime_error(s) {}
};
}  // namespace rabit

#ifndef RABIT_NO_DMLC_COMPAT
#ifndef DMLC_LOGGING_H_
namespace dmlc {
typedef ::rabit::Error Error;
}  // namespace dmlc
#endif
#endif

namespace rabit {
namespace utils {

/*! \brief longest formatted message */
static const int kMsgBufSize = 1 << 12;

/*! \brief "true"/"TRUE"/non-zero number -> true */
inline bool StringToBool(const char *s) {
  return strcasecmp(s, "true") == 0 || atoi(s) != 0;
}

namespace detail {
inline std::string VFormat(const char *fmt, va_list ap) {
  std::string out(kMsgBufSize, '\0');
  int n = vsnprintf(&out[0], out.size(), fmt, ap);
  if (n < 0) n = 0;
  if (static_cast<size_t>(n) >= out.size()) n = static_cast<int>(out.size()) - 1;
  out.resize(n);
  return out;
}
/*! \brief print to stderr and throw; the single exit for all fatal conditions */
[[noreturn]] inline void Raise(const char *kind, const std::string &msg) {
  fprintf(stderr, "%s%s\n", kind, msg.c_str());
  fflush(stderr);
  throw ::rabit::Error(std::string(kind) + msg);
}
}  // namespace detail

/*! \brief printf to stdout */
inline void Printf(const char *fmt, ...) {
  va_list ap; va_start(ap, fmt);
  std::string m = detail::VFormat(fmt, ap);
  va_end(ap);
  fputs(m.c_str(), stdout);
}
/*! \brief log line to stdout, flushed so launchers that pipe output see it promptly */
inline void HandleLogInfo(const char *fmt, ...) {
  va_list ap; va_start(ap, fmt);
  std::string m = detail::VFormat(fmt, ap);
  va_end(ap);
  fputs(m.c_str(), stdout);
  fflush(stdout);
}
/*! \brief internal invariant; failure means a bug in the library */
inline void Assert(bool ok, const char *fmt, ...) {
  if (ok) return;
  va_list ap; va_start(ap, fmt);
  std::string m = detail::VFormat(fmt, ap);
  va_end(ap);
  detail::Raise("AssertError:", m);
}
/*! \brief user-facing precondition */
inline void Check(bool ok, const char *fmt, ...) {
  if (ok) return;
  va_list ap; va_start(ap, fmt);
  std::string m = detail::VFormat(fmt, ap);
  va_end(ap);
  detail::Raise("", m);
}
/*! \brief unconditional failure */
[[noreturn]] inline void Error(const char *fmt, ...) {
  va_list ap; va_start(ap, fmt);
  std::string m = detail::VFormat(fmt, ap);
  va_end(ap);
  detail::Raise("", m);
}
}  // namespace utils

/*! \brief pointer to first element, NULL when empty */
template <typename T>
inline T *BeginPtr(std::vector<T> &vec) {  // NOLINT(*)
  return vec.empty() ? nullptr : vec.data();
}
template <typename T>
inline const T *BeginPtr(const std::vector<T> &vec) {
  return vec.empty() ? nullptr : vec.data();
}
inline char *BeginPtr(std::string &str) {  // NOLINT(*)
  return str.empty() ? nullptr : &str[0];
}
inline const char *BeginPtr(const std::string &str) {
  return str.empty() ? nullptr : str.data();
}
}  // namespace rabit
#endif  // RABIT_INTERNAL_UTILS_H_
