/*!
 * Wall-clock helper. Parity: reference include/rabit/internal/timer.h:21-38
 * (seconds as double); here on std::chrono instead of clock_gettime/mach.
 */
#ifndef RABIT_INTERNAL_TIMER_H_
#define RABIT_INTERNAL_TIMER_H_

#include <chrono>

namespace rabit {
namespace utils {
/*! \return seconds since the unix epoch, sub-microsecond resolution */
inline double GetTime(void) {
  using namespace std::chrono;  // NOLINT
  return duration_cast<duration<double> >(system_clock::now().time_since_epoch()).count();
}
/*! \return monotonic seconds, for intervals that must not jump with NTP */
inline double MonoTime(void) {
  using namespace std::chrono;  // NOLINT
  return duration_cast<duration<double> >(steady_clock::now().time_since_epoch()).count();
}
}  // namespace utils
}  // namespace rabit
#endif  // RABIT_INTERNAL_TIMER_H_
