/*!
 * rabit-b200 guide: the two-line allreduce. Each rank fills a[i] = rank + i; after Max
 * every rank holds world-1+i, after Sum it holds the sum over ranks of the maxima.
 * World of 4 prints {3,4,5} then {12,16,20} (same oracle as reference guide/basic.cc:14-32).
 *   python -m rabit_b200.parallel.launcher -n 4 build/bin/basic
 */
#include <rabit/rabit.h>

#include <cstdio>
#include <vector>

int main(int argc, char *argv[]) {
  const int kLen = 3;
  rabit::Init(argc, argv);
  const int me = rabit::GetRank();
  std::vector<int> a(kLen);
  for (int i = 0; i < kLen; ++i) a[i] = me + i;
  std::printf("@node[%d] before-allreduce: a={%d, %d, %d}\n", me, a[0], a[1], a[2]);
  rabit::Allreduce<rabit::op::Max>(a.data(), a.size());
  std::printf("@node[%d] after-allreduce-max: a={%d, %d, %d}\n", me, a[0], a[1], a[2]);
  rabit::Allreduce<rabit::op::Sum>(a.data(), a.size());
  std::printf("@node[%d] after-allreduce-sum: a={%d, %d, %d}\n", me, a[0], a[1], a[2]);
  rabit::Finalize();
  return 0;
}
