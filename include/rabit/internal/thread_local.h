/*!
 * Per-thread singleton store. Parity: reference include/rabit/internal/thread_local.h:38-85.
 * C++11 `thread_local` does all the work the reference hand-rolled for old compilers.
 */
#ifndef RABIT_INTERNAL_THREAD_LOCAL_H_
#define RABIT_INTERNAL_THREAD_LOCAL_H_

namespace rabit {
/*! \brief one lazily constructed T per thread; destroyed at thread exit */
template <typename T>
class ThreadLocalStore {
 public:
  static T *Get() {
    static thread_local T inst;
    return &inst;
  }
 private:
  ThreadLocalStore() = delete;
};
}  // namespace rabit
#endif  // RABIT_INTERNAL_THREAD_LOCAL_H_
