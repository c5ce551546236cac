/*!
 * rabit-b200 guide: lazy initialisation. The lambda that fills the buffer only runs when the
 * allreduce is really executed; a restarted rank that recovers the result from a peer's
 * cache skips it. Try: launcher -n 2 build/bin/lazy_allreduce mock=0,0,1,0
 * (cf. reference guide/lazy_allreduce.cc:15-33)
 */
#include <rabit/rabit.h>

#include <cstdio>

int main(int argc, char *argv[]) {
  const int kLen = 3;
  int a[kLen] = {0, 0, 0};
  rabit::Init(argc, argv);
  const int me = rabit::GetRank();
  auto fill = [&]() {
    std::printf("@node[%d] run prepare function\n", me);
    for (int i = 0; i < kLen; ++i) a[i] = me + i;
  };
  std::printf("@node[%d] before-allreduce: a={%d, %d, %d}\n", me, a[0], a[1], a[2]);
  rabit::Allreduce<rabit::op::Max>(&a[0], kLen, fill);
  std::printf("@node[%d] after-allreduce-max: a={%d, %d, %d}\n", me, a[0], a[1], a[2]);
  rabit::Allreduce<rabit::op::Sum>(&a[0], kLen);
  std::printf("@node[%d] after-allreduce-sum: a={%d, %d, %d}\n", me, a[0], a[1], a[2]);
  rabit::Finalize();
  return 0;
}
