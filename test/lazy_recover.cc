/*!
 * rabit-b200 test: LazyCheckPoint. The model is only serialised if a peer needs it, so it
 * must stay untouched between the lazy checkpoint and the last collective of the next
 * version: results are staged in `next` and swapped in right before the checkpoint.
 * Scenario parity: reference test/lazy_recover.cc + test/test.mk:30-34.
 */
#include "recover_common.h"

int main(int argc, char *argv[]) {
  if (argc < 2) { std::printf("Usage: lazy_recover <ndata> [name=value ...]\n"); return 0; }
  const size_t n = static_cast<size_t>(atol(argv[1]));
  const int iters = rtest::ArgInt(argc, argv, "iters", 3);
  rabit::Init(argc, argv);
  const int me = rabit::GetRank(), world = rabit::GetWorldSize();
  rtest::VecModel model;
  int it = rabit::LoadCheckPoint(&model);
  if (it == 0) model.Reset(n, 1.0f);
  std::printf("[%d] start at iter=%d\n", me, it);
  auto none = [](int) { return 0.0f; };
  for (; it < iters; ++it) {
    std::vector<float> next;
    rtest::StepMax(&next, model.w, it, true, none);
    for (int root = 0; root < world; root += std::max(world / 3, 1)) rtest::StepBroadcast(n, root, it);
    std::vector<float> next2;
    rtest::StepSum(&next2, next, it, none);
    for (float &x : next2) x = std::fmod(x, 1000.0f);
    model.w.swap(next2);  // no collective between here and the checkpoint
    rabit::LazyCheckPoint(&model);
    std::printf("[%d] iter %d lazily checkpointed (version %d)\n", me, it, rabit::VersionNumber());
  }
  rabit::Finalize();
  return 0;
}
