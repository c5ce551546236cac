/*!
 * rabit-b200: the bandwidth path — a warp-specialised, software-pipelined two-shot allreduce.
 *
 * The plain two-shot kernel runs stage -> reduce-scatter -> allgather one after the other, so
 * HBM (staging) and NVLink (the two transfer phases) are never busy together. Here one
 * persistent CTA per SM splits into three groups of 256 threads that work on different ROUNDS
 * of the same channel at the same time:
 *
 *   group 0  stager    user buffer -> local stage          (HBM only)        signals A[round]
 *   group 1  reducer   waits A[round] from every rank; pulls its owned tile from all stages
 *                      (or lets the switch do it: multimem.ld_reduce + multimem.st), writes
 *                      the reduced tile in place                              signals B[round]
 *   group 2  gatherer  waits B[round]; pulls every tile of the round from its owner's stage
 *                      (NVLS: from the local stage) into the user buffer and the cache copy
 *
 * A round of channel b is world consecutive tiles of that channel, one owned by each rank.
 * A[.] / B[.] are per-(block, source rank) monotonically increasing counters in the peer's
 * control block, so the three groups of all ranks form one long pipeline whose steady-state
 * cost per round is max(stage, reduce, gather) instead of their sum, and the two flag
 * latencies per round are hidden behind the rounds in flight.
 *
 * Fused consensus: a code-carrying handshake on the entry row happens once, before any group
 * touches data; a mismatch makes every rank leave with kStMismatch.
 * No counterpart in the reference (closest: the streaming ring pipeline of
 * allreduce_base.cc:857-946, which overlaps recv / reduce / send on one socket pair).
 */
#ifndef RABIT_SRC_DEVICE_COLL_PIPE_CUH_
#define RABIT_SRC_DEVICE_COLL_PIPE_CUH_

#include "coll_kernels.cuh"
#include "coll_nvls.cuh"

namespace rabit {
namespace device {

constexpr int kPipeGroup = 256;
constexpr int kPipeThreads = 3 * kPipeGroup;
constexpr int kPipeUnroll = 4;
constexpr size_t kPipeTileVecs = static_cast<size_t>(kPipeGroup) * kPipeUnroll;
constexpr size_t kPipeTileBytes = kPipeTileVecs * kVec;  // 16 KiB

__device__ __forceinline__ void GroupSync(int id) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(kPipeGroup) : "memory");
}

/*! \brief rounds a channel needs: every rank/block computes the same number */
__host__ __device__ inline unsigned long long PipeRounds(size_t bytes, int grid, int world) {
  const size_t ntiles = (bytes + kPipeTileBytes - 1) / kPipeTileBytes;
  const size_t per_chan = (ntiles + grid - 1) / grid;
  return (per_chan + world - 1) / world;
}

/*! \brief one group thread per peer waits until that peer's counter reaches `target` */
__device__ __forceinline__ void GroupWait(const Params &p, const unsigned long long *slot, unsigned long long target,
                                          int lane, int *s_abort) {
  if (lane < p.world) {
    const unsigned long long deadline = GlobalTimer() + p.timeout_ns;
    unsigned int spins = 0;
    while (LoadAcquireSys(slot + lane) < target) {
      if ((++spins & 0x3ff) == 0) {
        if (*reinterpret_cast<volatile int *>(s_abort) != 0) break;
        if (*p.abort_flag != 0) { atomicMax(s_abort, kStAborted); break; }
        if (GlobalTimer() > deadline) { atomicMax(s_abort, kStTimeout); break; }
      }
    }
  }
}

template <typename T, typename OP, typename MM, bool kNvls>
__device__ __forceinline__ void PipelinedAllreduceBody(const Params &p, unsigned long long round_base) {
  __shared__ int s_verdict;
  __shared__ int s_abort;
  const int me = p.emulate ? blockIdx.y : p.rank;
  const int ix = p.emulate ? me : 0;
  char *user = static_cast<char *>(p.user[ix]);
  char *keep = static_cast<char *>(p.keep[ix]);
  if (threadIdx.x == 0) s_abort = 0;
  Stamp(p, 0);
  // consensus + "everybody is here" before any byte moves
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[ix], true, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 1);

  const int n = p.world;
  const int b = blockIdx.x, G = gridDim.x;
  const int group = threadIdx.x / kPipeGroup, lane = threadIdx.x % kPipeGroup;
  const size_t nvec_full = p.bytes / kVec;
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t ntiles = (nvec + kPipeTileVecs - 1) / kPipeTileVecs;
  const size_t chan_tiles = ntiles > static_cast<size_t>(b) ? (ntiles - b + G - 1) / G : 0;
  const size_t rounds = (chan_tiles + n - 1) / n;
  Control *my_ctrl = CtrlOf(p, me);
  char *my_stage = StageOf(p, me);

  if (group == 0) {
    // ---------------------------------------------------------------- stager
    for (size_t j = 0; j < rounds; ++j) {
      for (int i = 0; i < n; ++i) {
        const size_t k = j * n + i;
        if (k >= chan_tiles) break;
        const size_t v0 = (b + k * G) * kPipeTileVecs + lane;
        uint4 r[kPipeUnroll];
#pragma unroll
        for (int u = 0; u < kPipeUnroll; ++u) {
          const size_t v = v0 + static_cast<size_t>(u) * kPipeGroup;
          if (v < nvec_full) r[u] = LoadVec(user + v * kVec);
        }
#pragma unroll
        for (int u = 0; u < kPipeUnroll; ++u) {
          const size_t v = v0 + static_cast<size_t>(u) * kPipeGroup;
          if (v < nvec_full) {
            StoreVec(my_stage + v * kVec, r[u]);
          } else if (v < nvec) {  // ragged last vector
            for (size_t q = v * kVec; q < p.bytes; ++q) my_stage[q] = user[q];
          }
        }
      }
      GroupSync(1);
      if (lane < n) StoreReleaseSys(&CtrlOf(p, lane)->pipe_a[b][me], round_base + j + 1);
    }
  } else if (group == 1) {
    // --------------------------------------------------------------- reducer
    for (size_t j = 0; j < rounds; ++j) {
      GroupWait(p, &my_ctrl->pipe_a[b][0], round_base + j + 1, lane, &s_abort);
      GroupSync(2);
      if (*reinterpret_cast<volatile int *>(&s_abort) != 0) break;
      const size_t k = j * n + me;
      if (k < chan_tiles) {
        const size_t v0 = (b + k * G) * kPipeTileVecs + lane;
        uint4 acc[kPipeUnroll];
        if (kNvls) {
          char *mc = McStage(p);
#pragma unroll
          for (int u = 0; u < kPipeUnroll; ++u) {
            const size_t v = v0 + static_cast<size_t>(u) * kPipeGroup;
            if (v < nvec) acc[u] = MM::LdReduce(mc + v * kVec);
          }
#pragma unroll
          for (int u = 0; u < kPipeUnroll; ++u) {
            const size_t v = v0 + static_cast<size_t>(u) * kPipeGroup;
            if (v < nvec) MultimemStore(mc + v * kVec, acc[u]);
          }
        } else {
#pragma unroll
          for (int u = 0; u < kPipeUnroll; ++u) {
            const size_t v = v0 + static_cast<size_t>(u) * kPipeGroup;
            if (v < nvec) acc[u] = LoadVec(StageOf(p, 0) + v * kVec);
          }
          for (int r = 1; r < n; ++r) {
            uint4 in[kPipeUnroll];
            const char *src = StageOf(p, r);
#pragma unroll
            for (int u = 0; u < kPipeUnroll; ++u) {
              const size_t v = v0 + static_cast<size_t>(u) * kPipeGroup;
              if (v < nvec) in[u] = LoadVec(src + v * kVec);
            }
#pragma unroll
            for (int u = 0; u < kPipeUnroll; ++u) {
              const size_t v = v0 + static_cast<size_t>(u) * kPipeGroup;
              if (v < nvec) ReduceVec<T, OP>(&acc[u], in[u]);
            }
          }
#pragma unroll
          for (int u = 0; u < kPipeUnroll; ++u) {
            const size_t v = v0 + static_cast<size_t>(u) * kPipeGroup;
            if (v < nvec) StoreVec(my_stage + v * kVec, acc[u]);
          }
        }
      }
      GroupSync(2);
      if (lane < n) StoreReleaseSys(&CtrlOf(p, lane)->pipe_b[b][me], round_base + j + 1);
    }
  } else {
    // -------------------------------------------------------------- gatherer
    for (size_t j = 0; j < rounds; ++j) {
      GroupWait(p, &my_ctrl->pipe_b[b][0], round_base + j + 1, lane, &s_abort);
      GroupSync(3);
      if (*reinterpret_cast<volatile int *>(&s_abort) != 0) break;
      for (int i = 0; i < n; ++i) {
        const size_t k = j * n + i;
        if (k >= chan_tiles) break;
        const char *src = kNvls ? my_stage : StageOf(p, i);
        const size_t v0 = (b + k * G) * kPipeTileVecs + lane;
        uint4 r[kPipeUnroll];
#pragma unroll
        for (int u = 0; u < kPipeUnroll; ++u) {
          const size_t v = v0 + static_cast<size_t>(u) * kPipeGroup;
          if (v < nvec) r[u] = LoadVec(src + v * kVec);
        }
#pragma unroll
        for (int u = 0; u < kPipeUnroll; ++u) {
          const size_t v = v0 + static_cast<size_t>(u) * kPipeGroup;
          if (v < nvec) Emit(user, keep, v, nvec_full, p.bytes, r[u], true);
        }
      }
    }
  }
  __syncthreads();
  if (s_abort != 0) Report(p, me, s_abort);
  Stamp(p, 5);
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
template <typename T, typename OP, typename MM, bool kNvls>
__global__ void __launch_bounds__(kPipeThreads, 1) PipelinedAllreduce(Params p, unsigned long long round_base) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  PipelinedAllreduceBody<T, OP, MM, kNvls>(p, round_base);
  EndOp(p, me);
}

}  // namespace device
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_COLL_PIPE_CUH_
