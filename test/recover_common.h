/*!
 * rabit-b200: self-checking workload shared by the recovery test programs.
 *
 * One "iteration" = Max-allreduce, a few string broadcasts from different roots, a
 * Sum-allreduce, optionally an allgather, then a checkpoint. Every collective's result
 * is compared with a closed-form expectation, so a rank that was killed (mock=...) and
 * replayed from the checkpoint + result cache must reproduce exactly what the survivors
 * saw. Plays the role of reference test/model_recover.cc, local_recover.cc and
 * lazy_recover.cc; the formulas are this build's own.
 */
#ifndef RABIT_TEST_RECOVER_COMMON_H_
#define RABIT_TEST_RECOVER_COMMON_H_

#include <execinfo.h>
#include <rabit/rabit.h>
#include <signal.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

namespace rtest {

/*! \brief post-mortem aid for the kill campaigns: RABIT_TEST_STACKDUMP=1 makes SIGUSR1 print the stack of the worker */
inline void StackDumpHandler(int) {
  void *frames[48];
  const int n = backtrace(frames, 48);
  backtrace_symbols_fd(frames, n, 2);
}
struct StackDumpInstaller {
  StackDumpInstaller() { if (getenv("RABIT_TEST_STACKDUMP") != nullptr) signal(SIGUSR1, StackDumpHandler); }
};
static StackDumpInstaller g_stack_dump_installer;

/*! \brief a vector of floats as a checkpointable model */
struct VecModel : public rabit::Serializable {
  std::vector<float> w;
  void Load(rabit::Stream *fi) override { fi->Read(&w); }
  void Save(rabit::Stream *fo) const override { fo->Write(w); }
  void Reset(size_t n, float v) { w.assign(n, v); }
};

/*! \brief what rank r contributes at position i in iteration it (small exact integers) */
inline float Contribution(size_t i, int r, int it, int modulus) {
  return static_cast<float>((i * static_cast<size_t>(r + 1) + static_cast<size_t>(it)) % modulus);
}

/*!
 * \brief Max step. `bias[r]` is what rank r adds on top of its contribution (0 for the
 *  global-model tests, the rank for the local-model test).
 */
template <typename Fill>
inline void StepMax(std::vector<float> *out, const std::vector<float> &base, int it, bool lazy,
                    Fill per_rank_bias) {
  const int me = rabit::GetRank(), n = rabit::GetWorldSize();
  const int mod = 97 + it;
  std::vector<float> v(base.size());
  auto fill = [&]() {
    for (size_t i = 0; i < v.size(); ++i) v[i] = Contribution(i, me, it, mod) + base[i] + per_rank_bias(me);
  };
  if (lazy) {
    rabit::Allreduce<rabit::op::Max>(v.data(), v.size(), fill);
  } else {
    fill();
    rabit::Allreduce<rabit::op::Max>(v.data(), v.size());
  }
  for (size_t i = 0; i < v.size(); ++i) {
    float want = Contribution(i, 0, it, mod) + base[i] + per_rank_bias(0);
    for (int r = 1; r < n; ++r) want = std::max(want, Contribution(i, r, it, mod) + base[i] + per_rank_bias(r));
    rabit::utils::Check(want == v[i], "[%d] max mismatch at %lu: want %g got %g (iter %d)", me,
                        static_cast<unsigned long>(i), want, v[i], it);  // NOLINT(*)
  }
  *out = v;
}

template <typename Fill>
inline void StepSum(std::vector<float> *out, const std::vector<float> &base, int it, Fill per_rank_bias) {
  const int me = rabit::GetRank(), n = rabit::GetWorldSize();
  const int mod = 113 + it;
  std::vector<float> v(base.size());
  for (size_t i = 0; i < v.size(); ++i) v[i] = Contribution(i, me, it, mod) + base[i] + per_rank_bias(me);
  rabit::Allreduce<rabit::op::Sum>(v.data(), v.size());
  for (size_t i = 0; i < v.size(); ++i) {
    double want = 0.0;
    for (int r = 0; r < n; ++r) want += Contribution(i, r, it, mod) + base[i] + per_rank_bias(r);
    rabit::utils::Check(std::fabs(want - v[i]) <= 1e-5 * std::max(1.0, std::fabs(want)),
                        "[%d] sum mismatch at %lu: want %g got %g (iter %d)", me,
                        static_cast<unsigned long>(i), want, v[i], it);  // NOLINT(*)
  }
  *out = v;
}

inline void StepBroadcast(size_t len, int root, int it) {
  std::string want(len, ' ');
  for (size_t i = 0; i < len; ++i) want[i] = static_cast<char>('!' + (i * 7 + root + it) % 90);
  std::string got;
  if (rabit::GetRank() == root) got = want;
  rabit::Broadcast(&got, root);
  rabit::utils::Check(got == want, "[%d] broadcast from %d mismatch (iter %d)", rabit::GetRank(), root, it);
}

/*! \brief every rank contributes `per` floats; checks the concatenation */
inline void StepAllgather(size_t per, int it) {
  const int me = rabit::GetRank(), n = rabit::GetWorldSize();
  std::vector<float> all(per * n, -1.0f);
  for (size_t i = 0; i < per; ++i) all[me * per + i] = Contribution(i, me, it, 89);
  rabit::Allgather(all.data(), all.size(), me * per, per, per);
  for (int r = 0; r < n; ++r) {
    for (size_t i = 0; i < per; ++i) {
      rabit::utils::Check(all[r * per + i] == Contribution(i, r, it, 89),
                          "[%d] allgather mismatch in slice %d (iter %d)", me, r, it);
    }
  }
}

inline int ArgInt(int argc, char *argv[], const char *name, int dflt) {
  const size_t n = strlen(name);
  for (int i = 1; i < argc; ++i) {
    if (strncmp(argv[i], name, n) == 0 && argv[i][n] == '=') return atoi(argv[i] + n + 1);
  }
  return dflt;
}

}  // namespace rtest
#endif  // RABIT_TEST_RECOVER_COMMON_H_
