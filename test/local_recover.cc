/*!
 * rabit-b200 test: global + per-rank (local) model recovery with the lazy-prepare lambda.
 * The local model differs per rank (base + rank), is ring-replicated at every checkpoint
 * and must come back intact on a restarted rank. Scenario parity: reference
 * test/local_recover.cc + test/test.mk:24-25.
 */
#include "recover_common.h"

int main(int argc, char *argv[]) {
  if (argc < 2) { std::printf("Usage: local_recover <ndata> [name=value ...]\n"); return 0; }
  const size_t n = static_cast<size_t>(atol(argv[1]));
  const int iters = rtest::ArgInt(argc, argv, "iters", 3);
  rabit::Init(argc, argv);
  const int me = rabit::GetRank(), world = rabit::GetWorldSize();
  rtest::VecModel model, local;
  int it = rabit::LoadCheckPoint(&model, &local);
  if (it == 0) { model.Reset(n, 1.0f); local.Reset(n, 1.0f + me); }
  std::printf("[%d] start at iter=%d\n", me, it);
  // invariant: local.w[i] == model.w[i] + rank, so the expectation only needs the global model
  for (size_t i = 0; i < n; ++i) {
    rabit::utils::Check(local.w[i] == model.w[i] + me, "[%d] local model corrupt after load (iter %d)", me, it);
  }
  auto by_rank = [](int r) { return static_cast<float>(r); };
  for (; it < iters; ++it) {
    std::vector<float> out;
    rtest::StepMax(&out, model.w, it, true, by_rank);
    for (int root = 0; root < world; root += std::max(world / 3, 1)) rtest::StepBroadcast(n, root, it);
    rtest::StepSum(&out, model.w, it, by_rank);
    for (size_t i = 0; i < n; ++i) {
      model.w[i] = std::fmod(out[i], 1000.0f);
      local.w[i] = model.w[i] + me;
    }
    rabit::CheckPoint(&model, &local);
    std::printf("[%d] iter %d checkpointed (version %d)\n", me, it, rabit::VersionNumber());
  }
  rabit::Finalize();
  return 0;
}
