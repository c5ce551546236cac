/*!
 * Baseline harness: the bench.py step (Allreduce<Sum> over `count` floats) written against
 * the REFERENCE's public API (rabit/rabit.h of bobjiang82/rabit) and linked with the
 * reference's own, unmodified engine sources. Metric definition follows the reference's
 * test/speed_test.cc: wall time around rabit::Allreduce, here reported as the max over ranks
 * of the per-step mean so that it is comparable with the device-timed numbers.
 * Usage: ref_speed <count> <steps> <warmup> [rabit params...]; rank 0 prints "REF_RESULT <ms>".
 */
#include <rabit/rabit.h>
#include <rabit/internal/timer.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

int main(int argc, char *argv[]) {
  if (argc < 4) { std::printf("usage: ref_speed <count> <steps> <warmup> [params]\n"); return 1; }
  const size_t count = static_cast<size_t>(atoll(argv[1]));
  const int steps = atoi(argv[2]), warmup = atoi(argv[3]);
  rabit::Init(argc, argv);
  const int rank = rabit::GetRank(), world = rabit::GetWorldSize();
  std::vector<float> buf(count), src(count);
  for (size_t i = 0; i < count; ++i) src[i] = static_cast<float>((i % 131) * (rank + 1));
  for (int i = 0; i < warmup; ++i) { buf = src; rabit::Allreduce<rabit::op::Sum>(&buf[0], count); }
  const float want = static_cast<float>((5 % 131) * (world * (world + 1) / 2));
  rabit::utils::Check(count < 6 || buf[5] == want, "reference allreduce result is wrong");
  double total = 0.0;
  for (int i = 0; i < steps; ++i) {
    buf = src;
    int sync = 0;
    rabit::Allreduce<rabit::op::Sum>(&sync, 1);  // line the ranks up, like a barrier
    const double t0 = rabit::utils::GetTime();
    rabit::Allreduce<rabit::op::Sum>(&buf[0], count);
    total += rabit::utils::GetTime() - t0;
  }
  double ms = total / steps * 1e3;
  rabit::Allreduce<rabit::op::Max>(&ms, 1);
  if (rank == 0) { std::printf("REF_RESULT %f\n", ms); std::fflush(stdout); }
  rabit::Finalize();
  return 0;
}
