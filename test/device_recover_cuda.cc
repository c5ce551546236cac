/*!
 * rabit-b200 test: the recovery workload on CUDA device buffers, through the ordinary C++ API.
 * Same closed-form checks as model_recover (recover_common.h), but every collective is called
 * with DEVICE pointers, so it runs the fused sm_100a kernels (consensus + transfer + reduce +
 * cache copy in one launch) and, after a mock kill, replays from GPU-resident cache entries.
 * A native worker restarts in a fraction of the time a python/torch one needs, which is what
 * benchmarks/recovery_time.py --native measures.
 * usage: device_recover_cuda <ndata> [iters=N] [mock=r,v,s,t ...] [rabit_*=...]
 */
#include <cuda_runtime.h>

#include "recover_common.h"

#define CK(x) do { cudaError_t e_ = (x); rabit::utils::Check(e_ == cudaSuccess, "CUDA: %s (%s:%d)", \
                   cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

namespace {
struct GradPair { double g, h; };
void PairSum(GradPair &d, const GradPair &s) { d.g += s.g; d.h += s.h; }  // NOLINT(*)
struct DevBuf {
  float *d = nullptr; size_t n = 0;
  explicit DevBuf(size_t n_) : n(n_) { CK(cudaMalloc(&d, n * sizeof(float))); }
  ~DevBuf() { cudaFree(d); }
  void Up(const std::vector<float> &h) { CK(cudaMemcpy(d, h.data(), n * sizeof(float), cudaMemcpyHostToDevice)); }
  std::vector<float> Down() const {
    std::vector<float> h(n);
    CK(cudaMemcpy(h.data(), d, n * sizeof(float), cudaMemcpyDeviceToHost));
    return h;
  }
};
}  // namespace

int main(int argc, char *argv[]) {
  if (argc < 2) { std::printf("Usage: device_recover_cuda <ndata> [name=value ...]\n"); return 0; }
  const size_t n = static_cast<size_t>(atol(argv[1]));
  const int iters = rtest::ArgInt(argc, argv, "iters", 3);
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  const char *ord = getenv("RABIT_DEVICE_ORDINAL");
  CK(cudaSetDevice((ord != nullptr ? atoi(ord) : 0) % ndev));
  rabit::Init(argc, argv);
  const int me = rabit::GetRank(), world = rabit::GetWorldSize();
  rtest::VecModel model;
  int it = rabit::LoadCheckPoint(&model);
  if (it == 0) model.Reset(n, 1.0f);
  std::printf("[%d] start at iter=%d\n", me, it);
  DevBuf buf(n);
  for (; it < iters; ++it) {
    // Max with a lazy prepare that fills the DEVICE buffer
    std::vector<float> h(n);
    const int mod = 97 + it;
    auto fill = [&]() {
      for (size_t i = 0; i < n; ++i) h[i] = rtest::Contribution(i, me, it, mod) + model.w[i];
      buf.Up(h);
    };
    rabit::Allreduce<rabit::op::Max>(buf.d, n, fill);
    std::vector<float> got = buf.Down();
    for (size_t i = 0; i < n; ++i) {
      float want = rtest::Contribution(i, 0, it, mod) + model.w[i];
      for (int r = 1; r < world; ++r) want = std::max(want, rtest::Contribution(i, r, it, mod) + model.w[i]);
      rabit::utils::Check(want == got[i], "[%d] device max mismatch at %lu (iter %d)", me, (unsigned long)i, it);  // NOLINT(*)
    }
    // Broadcast of the device buffer from a rotating root
    const int root = it % world;
    if (me == root) { for (size_t i = 0; i < n; ++i) h[i] = static_cast<float>((i * 3 + it) % 1000); buf.Up(h); }
    rabit::Broadcast(buf.d, n * sizeof(float), root);
    std::vector<float> b = buf.Down();
    for (size_t i = 0; i < n; ++i) {
      rabit::utils::Check(b[i] == static_cast<float>((i * 3 + it) % 1000), "[%d] device broadcast mismatch (iter %d)", me, it);
    }
    // Sum
    const int mod2 = 113 + it;
    for (size_t i = 0; i < n; ++i) h[i] = rtest::Contribution(i, me, it, mod2) + got[i];
    buf.Up(h);
    rabit::Allreduce<rabit::op::Sum>(buf.d, n);
    std::vector<float> s = buf.Down();
    for (size_t i = 0; i < n; ++i) {
      double want = 0.0;
      for (int r = 0; r < world; ++r) want += rtest::Contribution(i, r, it, mod2) + got[i];
      rabit::utils::Check(std::fabs(want - s[i]) <= 1e-5 * std::max(1.0, std::fabs(want)),
                          "[%d] device sum mismatch at %lu (iter %d)", me, (unsigned long)i, it);  // NOLINT(*)
    }
    // custom struct reducer on DEVICE memory: {double g, h} pairs. First as an opaque reducer (staged
    // through the host engine), then declared as a packed sum (fused float64 kernel); same answer.
    for (int declared = 0; declared < 2; ++declared) {
      const size_t np = 1000;
      std::vector<GradPair> hp(np);
      for (size_t i = 0; i < np; ++i) { hp[i].g = static_cast<double>((i + me + it) % 17); hp[i].h = 0.5 * me; }
      GradPair *dp = nullptr;
      CK(cudaMalloc(&dp, np * sizeof(GradPair)));
      CK(cudaMemcpy(dp, hp.data(), np * sizeof(GradPair), cudaMemcpyHostToDevice));
      rabit::Reducer<GradPair, PairSum> red;
      if (declared) red.DeclarePackedSum<double>(2);
      red.Allreduce(dp, np);
      CK(cudaMemcpy(hp.data(), dp, np * sizeof(GradPair), cudaMemcpyDeviceToHost));
      cudaFree(dp);
      for (size_t i = 0; i < np; ++i) {
        double g = 0, h = 0;
        for (int r = 0; r < world; ++r) { g += static_cast<double>((i + r + it) % 17); h += 0.5 * r; }
        rabit::utils::Check(hp[i].g == g && hp[i].h == h, "[%d] pair reducer (declared=%d) mismatch (iter %d)", me, declared, it);
      }
    }
    for (size_t i = 0; i < n; ++i) model.w[i] = std::fmod(s[i], 1000.0f);
    rabit::CheckPoint(&model);
    std::printf("[%d] iter %d checkpointed (version %d)\n", me, it, rabit::VersionNumber());
  }
  rabit::Finalize();
  return 0;
}
