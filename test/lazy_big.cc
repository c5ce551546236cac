/*!
 * rabit-b200 benchmark/test: LazyCheckPoint of a LARGE global model fetched by a restarted rank
 * (BASELINE config 5: "LazyCheckPoint of a 1 GB model"; reference lazy path allreduce_robust.cc:527-528,
 * 1070-1076 and test/lazy_recover.cc:86-124). Every iteration is one tiny allreduce plus a lazy checkpoint,
 * so the run time of a life that starts with a recovery is the time to move the model: the holder
 * serialises it on demand, then it travels holder -> requester (NVLink peer copies when the device arena
 * is up, the TCP ring otherwise). The restarted rank prints the seconds its LoadCheckPoint took.
 * usage: lazy_big <nfloats> [iters=N] [mock=r,v,s,t ...] [rabit_*=...]
 */
#include "recover_common.h"

#include "rabit/internal/timer.h"

int main(int argc, char *argv[]) {
  if (argc < 2) { std::printf("Usage: lazy_big <nfloats> [name=value ...]\n"); return 0; }
  const size_t n = static_cast<size_t>(atol(argv[1]));
  const int iters = rtest::ArgInt(argc, argv, "iters", 3);
  rabit::Init(argc, argv);
  const int me = rabit::GetRank(), world = rabit::GetWorldSize();
  rtest::VecModel model;
  const double t0 = rabit::utils::GetTime();
  int it = rabit::LoadCheckPoint(&model);
  const double load_s = rabit::utils::GetTime() - t0;
  if (it == 0) {
    model.Reset(n, 1.0f);
  } else {
    rabit::utils::Check(model.w.size() == n, "[%d] recovered model has %lu values, expected %lu", me,
                        static_cast<unsigned long>(model.w.size()), static_cast<unsigned long>(n));  // NOLINT(*)
    for (size_t i = 0; i < n; i += 4097) {
      rabit::utils::Check(model.w[i] == static_cast<float>(it) + static_cast<float>(i % 7), "[%d] recovered model is corrupt at %lu",
                          me, static_cast<unsigned long>(i));  // NOLINT(*)
    }
    const double gb = (n * sizeof(float)) / 1e9;
    std::printf("[%d] LAZY_LOAD version %d: %.3f GB in %.4f s (%.2f GB/s; NVLink at 900 GB/s would need %.4f s)\n", me, it, gb,
                load_s, gb / load_s, gb / 900.0);
  }
  for (; it < iters; ++it) {
    float v = static_cast<float>(me + it);
    rabit::Allreduce<rabit::op::Sum>(&v, 1);
    rabit::utils::Check(v == static_cast<float>(world * (world - 1) / 2 + world * it), "[%d] sum mismatch (iter %d)", me, it);
    // the new version of the model: value = version + (i % 7); no collective between here and the checkpoint
    for (size_t i = 0; i < n; ++i) model.w[i] = static_cast<float>(it + 1) + static_cast<float>(i % 7);
    rabit::LazyCheckPoint(&model);
  }
  rabit::Finalize();
  return 0;
}
