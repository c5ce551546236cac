/*!
 * rabit-b200: the throughput program (role of reference test/speed_test.cc + speed_runner.py):
 * nrep x {Allreduce Max, Allreduce Sum over n floats, Broadcast of n bytes from a rotating root} on HOST
 * buffers through the selected engine; prints mean/std seconds per op class across ranks and the
 * algorithm bandwidth in MB/s (n * nrep * sizeof / mean_time / 2^20), the reference's metric.
 * The device-buffer version of this measurement is bench.py / benchmarks/algo_sweep.py.
 * usage: speed_test <ndata> <nrepeat> [name=value ...]
 */
#include <rabit/rabit.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

namespace {
double t_max = 0, t_sum = 0, t_bcast = 0;

void Report(const char *name, double mine, double unit_bytes, int nrep) {
  const int world = rabit::GetWorldSize();
  double sum = mine;
  rabit::Allreduce<rabit::op::Sum>(&sum, 1);
  const double mean = sum / world;
  double sq = (mine - mean) * (mine - mean);
  rabit::Allreduce<rabit::op::Sum>(&sq, 1);
  if (rabit::GetRank() == 0) {
    rabit::TrackerPrintf("%s: mean=%g, std=%g sec\n", name, mean, std::sqrt(sq / world));
    if (unit_bytes > 0) {
      rabit::TrackerPrintf("%s-speed: %g MB/sec\n", name, unit_bytes * nrep / mean / 1048576.0);
    }
  }
}
}  // namespace

int main(int argc, char *argv[]) {
  if (argc < 3) { std::printf("Usage: speed_test <ndata> <nrepeat> [name=value ...]\n"); return 0; }
  const size_t n = static_cast<size_t>(atol(argv[1]));
  const int nrep = atoi(argv[2]);
  rabit::utils::Check(nrep >= 1, "need to at least repeat running once");
  rabit::Init(argc, argv);
  const int me = rabit::GetRank(), world = rabit::GetWorldSize();
  std::vector<float> v(n);
  std::string s(n, 'x');
  const double t0 = rabit::utils::GetTime();
  for (int rep = 0; rep < nrep; ++rep) {
    for (size_t i = 0; i < n; ++i) v[i] = static_cast<float>((i * (me + 1)) % 111);
    double t = rabit::utils::GetTime();
    rabit::Allreduce<rabit::op::Max>(v.data(), n);
    t_max += rabit::utils::GetTime() - t;
    for (size_t i = 0; i < n; ++i) v[i] = static_cast<float>((i * (me + 1)) % 131);
    t = rabit::utils::GetTime();
    rabit::Allreduce<rabit::op::Sum>(v.data(), n);
    t_sum += rabit::utils::GetTime() - t;
    const int root = rep % world;
    t = rabit::utils::GetTime();
    rabit::Broadcast(&s[0], s.length(), root);
    t_bcast += rabit::utils::GetTime() - t;
  }
  const double total = rabit::utils::GetTime() - t0;
  Report("max_tdiff", t_max, static_cast<double>(n) * sizeof(float), nrep);
  Report("sum_tdiff", t_sum, static_cast<double>(n) * sizeof(float), nrep);
  Report("bcast_tdiff", t_bcast, static_cast<double>(n), nrep);
  Report("tot_tdiff", total, 0, nrep);
  rabit::Finalize();
  return 0;
}
