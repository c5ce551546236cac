/*!
 * rabit-b200: round-interleaved staged allreduce (the default large-message path for ordinary
 * device pointers).
 *
 * TwoShotAllreduce / NvlsAllreduce run "stage everything -> transfer everything -> copy out
 * everything": while HBM is busy NVLink idles and vice versa, on all 148 SMs at once. Here every
 * block walks its channel in ROUNDS of `world` tiles (one owned by each rank) and always works one
 * round ahead with its local traffic:
 *
 *     stage(0)  signal A0
 *     for j:    stage(j+1), signal A(j+1)      HBM
 *               wait A(j) of all ranks
 *               reduce my tile of round j       NVLink (P2P pulls, or multimem.ld_reduce + multimem.st)
 *               signal B(j)
 *               wait B(j-1), copy round j-1 out HBM (+ NVLink pulls in the P2P flavour)
 *
 * All 512 threads take part in every phase (unlike the warp-specialised PipelinedAllreduce, whose
 * 256-thread groups kept too little in flight); the overlap of HBM and NVLink comes from the blocks
 * of the 8 x 148 SMs being in different phases at any moment. A/B are the per-(block, rank) round
 * counters in the control block. Consensus: one code-carrying entry handshake before round 0.
 * No counterpart in the reference (closest: allreduce_base.cc:857-946, the streaming ring pipeline).
 */
#ifndef RABIT_SRC_DEVICE_COLL_ROUNDS_CUH_
#define RABIT_SRC_DEVICE_COLL_ROUNDS_CUH_

#include "coll_kernels.cuh"
#include "coll_nvls.cuh"

namespace rabit {
namespace device {

template <typename T, typename OP, typename MM, bool kNvls>
__device__ __forceinline__ void RoundsAllreduceBody(const Params &p, unsigned long long round_base) {
  __shared__ int s_verdict;
  __shared__ int s_abort;
  const int me = p.rank, n = p.world;
  const int b = blockIdx.x, G = gridDim.x;
  char *user = static_cast<char *>(p.user[0]);
  char *keep = static_cast<char *>(p.keep[0]);
  const size_t nvec = p.bytes / kVec;  // dispatch guarantees bytes % 16 == 0 and aligned pointers
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const size_t chan_tiles = ntiles > static_cast<size_t>(b) ? (ntiles - b + G - 1) / G : 0;
  const size_t rounds = (chan_tiles + n - 1) / n;
  if (threadIdx.x == 0) s_abort = 0;
  Stamp(p, 0);
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[0], true, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 1);
  char *my_stage = StageOf(p, me);
  Control *my_ctrl = CtrlOf(p, me);

  auto stage_round = [&](size_t j) {
    for (int i = 0; i < n; ++i) {
      const size_t k = j * n + i;
      if (k >= chan_tiles) break;
      const size_t v0 = (b + k * G) * tile_vecs + threadIdx.x;
      uint4 r[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) r[u] = LoadVec(user + v * kVec);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) StoreVec(my_stage + v * kVec, r[u]);
      }
    }
    __syncthreads();
    if (threadIdx.x < n) StoreReleaseSys(&CtrlOf(p, threadIdx.x)->pipe_a[b][me], round_base + j + 1);
  };
  // thread i < n polls rank i's counter; the block follows. false = abandon the op.
  auto wait_all = [&](const unsigned long long *row, unsigned long long target) -> bool {
    if (threadIdx.x < n) {
      const unsigned long long deadline = GlobalTimer() + p.timeout_ns;
      unsigned int spins = 0;
      while (LoadAcquireSys(row + threadIdx.x) < target) {
        if ((++spins & 0x3ff) == 0) {
          if (*reinterpret_cast<volatile int *>(&s_abort) != 0) break;
          if (*p.abort_flag != 0) { atomicMax(&s_abort, kStAborted); break; }
          if (GlobalTimer() > deadline) { atomicMax(&s_abort, kStTimeout); break; }
        }
      }
    }
    __syncthreads();
    return *reinterpret_cast<volatile int *>(&s_abort) == 0;
  };
  auto copy_out = [&](size_t j) {
    for (int i = 0; i < n; ++i) {
      const size_t k = j * n + i;
      if (k >= chan_tiles) break;
      const char *src = kNvls ? my_stage : StageOf(p, i);  // NVLS: the owner multicast it into our stage
      const size_t v0 = (b + k * G) * tile_vecs + threadIdx.x;
      uint4 r[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) r[u] = LoadVec(src + v * kVec);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) {
          StoreVec(user + v * kVec, r[u]);
          if (keep != nullptr) StoreVec(keep + v * kVec, r[u]);
        }
      }
    }
  };

  if (rounds > 0) stage_round(0);
  bool ok = true;
  for (size_t j = 0; j < rounds && ok; ++j) {
    if (j + 1 < rounds) stage_round(j + 1);
    ok = wait_all(&my_ctrl->pipe_a[b][0], round_base + j + 1);
    if (!ok) break;
    const size_t k = j * n + me;
    if (k < chan_tiles) {
      const size_t v0 = (b + k * G) * tile_vecs + threadIdx.x;
      uint4 acc[kUnroll];
      if (kNvls) {
        char *mc = McStage(p);
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
          if (v < nvec) acc[u] = MM::LdReduce(mc + v * kVec);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
          if (v < nvec) MultimemStore(mc + v * kVec, acc[u]);
        }
      } else {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
          if (v < nvec) acc[u] = LoadVec(StageOf(p, 0) + v * kVec);
        }
        for (int r = 1; r < n; ++r) {
          uint4 in[kUnroll];
          const char *src = StageOf(p, r);
#pragma unroll
          for (int u = 0; u < kUnroll; ++u) {
            const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
            if (v < nvec) in[u] = LoadVec(src + v * kVec);
          }
#pragma unroll
          for (int u = 0; u < kUnroll; ++u) {
            const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
            if (v < nvec) ReduceVec<T, OP>(&acc[u], in[u]);
          }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
          if (v < nvec) StoreVec(my_stage + v * kVec, acc[u]);
        }
      }
    }
    __syncthreads();
    if (threadIdx.x < n) StoreReleaseSys(&CtrlOf(p, threadIdx.x)->pipe_b[b][me], round_base + j + 1);
    if (j > 0) {
      ok = wait_all(&my_ctrl->pipe_b[b][0], round_base + j);
      if (ok) copy_out(j - 1);
    }
  }
  if (ok && rounds > 0) {
    ok = wait_all(&my_ctrl->pipe_b[b][0], round_base + rounds);
    if (ok) copy_out(rounds - 1);
  }
  if (s_abort != 0) Report(p, me, s_abort);
  Stamp(p, 5);
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
template <typename T, typename OP, typename MM, bool kNvls>
__global__ void __launch_bounds__(512) RoundsAllreduce(Params p, unsigned long long round_base) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  RoundsAllreduceBody<T, OP, MM, kNvls>(p, round_base);
  EndOp(p, me);
}

/*! \brief rounds RoundsAllreduce consumes (identical on every rank) */
inline unsigned long long AllreduceRounds(size_t bytes, int grid, int world) {
  const size_t tile = 512 * kUnroll * kVec;
  const size_t ntiles = (bytes + tile - 1) / tile;
  const size_t per_chan = (ntiles + grid - 1) / grid;
  return (per_chan + world - 1) / world;
}

}  // namespace device
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_COLL_ROUNDS_CUH_
