/*!
 * rabit-b200 test: global-model recovery. Usage: model_recover <ndata> [name=value ...]
 * Issues one collective BEFORE LoadCheckPoint (needs rabit_bootstrap_cache=1 to survive a
 * restart), then 3 checkpointed iterations of max / broadcasts / sum /
 * allgather. Scenario parity: reference test/model_recover.cc + test/test.mk:15-22.
 */
#include "recover_common.h"

int main(int argc, char *argv[]) {
  if (argc < 2) { std::printf("Usage: model_recover <ndata> [name=value ...]\n"); return 0; }
  const size_t n = static_cast<size_t>(atol(argv[1]));
  const int iters = rtest::ArgInt(argc, argv, "iters", 3);
  rabit::Init(argc, argv);
  const int me = rabit::GetRank(), world = rabit::GetWorldSize();
  int top = me;
  rabit::Allreduce<rabit::op::Max>(&top, 1);
  rabit::utils::Check(top == world - 1, "[%d] bootstrap allreduce: max rank %d != %d", me, top, world - 1);

  rtest::VecModel model;
  int it = rabit::LoadCheckPoint(&model);
  if (it == 0) model.Reset(n, 1.0f);
  std::printf("[%d] start at iter=%d trial=%d\n", me, it, rtest::ArgInt(argc, argv, "rabit_num_trial", 0));
  auto none = [](int) { return 0.0f; };
  for (; it < iters; ++it) {
    rtest::StepMax(&model.w, model.w, it, false, none);
    for (int root = 0; root < world; root += std::max(world / 3, 1)) rtest::StepBroadcast(n, root, it);
    rtest::StepSum(&model.w, model.w, it, none);
    rtest::StepAllgather(std::min<size_t>(n, 64), it);
    // keep values small so float sums stay exact across iterations
    for (float &x : model.w) x = std::fmod(x, 1000.0f);
    rabit::CheckPoint(&model);
    std::printf("[%d] iter %d checkpointed (version %d)\n", me, it, rabit::VersionNumber());
  }
  rabit::Finalize();
  return 0;
}
