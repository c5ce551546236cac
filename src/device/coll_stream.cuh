/*!
 * rabit-b200: chunk kernels of the STREAMED allreduce (large messages on the fault-tolerant path, and
 * host buffers that ride PCIe + NVLink).
 *
 * A large allreduce is cut into chunks by the host (device_backend_stream.inc). Copy engines move chunk
 * i+1 of the caller's buffer into the stage ("source region": it is both what the peers read and the
 * backup of the caller's input) while the SMs reduce chunk i; the reduced chunk is assembled in a second
 * arena region ("result region") on EVERY rank and leaves it either in this kernel's epilogue (device
 * destinations: the caller's buffer and the result-cache entry, one pass from registers) or through a
 * copy engine (host destinations). The caller's buffer may therefore be overwritten chunk by chunk while
 * later chunks are still in flight; the all-or-nothing guarantee of the fault-tolerant API comes from
 * the backup: after a failed kernel the host copies the source region back over the caller's buffer.
 * (The single-launch kernels keep the CommitGate instead; it costs them a serial copy-out.)
 *
 *   StreamP2P   entry barrier -> owner LDGs its tiles from every rank's source region, reduces, stores
 *               the result into its own result region (+ destinations) -> mid barrier -> everybody
 *               pulls the tiles it does not own from the owners' result regions -> exit barrier
 *   StreamNvls  entry barrier -> owner: multimem.ld_reduce on the source region, multimem.st into every
 *               rank's result region -> mid barrier -> local epilogue result region -> destinations
 *
 * Offsets are arena-relative and identical on all ranks. A kernel that is not the first of its op first
 * looks at the status word: after a failed chunk the peers may have left (consensus mismatch), so the
 * rest of the op must not wait for them.
 *
 * Reference counterpart of the whole op: ring reduce-scatter + allgather, allreduce_base.cc:857-977,
 * wrapped by allreduce_robust.cc:243-303 (cache slot copy in, reduce there, copy out).
 */
#ifndef RABIT_SRC_DEVICE_COLL_STREAM_CUH_
#define RABIT_SRC_DEVICE_COLL_STREAM_CUH_

#include "coll_inplace.cuh"

namespace rabit {
namespace device {

/*! \brief per-launch arguments of a chunk kernel (Params stays as it is for every other kernel) */
struct StreamArgs {
  size_t src_off;   // arena-relative start of this chunk in every rank's source region
  size_t res_off;   // arena-relative start of this chunk in every rank's result region
  int fill_result;  // P2P: also copy the tiles this rank does not own into its own result region
  int res_in_stage; // NVLS: the result region is this op's stage (chosen by the op's parity on the device); res_off unused
  int gate;         // NVLS: single-kernel recoverable op: nothing is written to the destinations unless the whole grid
                    // passed its last barrier (CommitGate), so no backup of the caller's buffer is needed
};

/*! \brief store one result vector to whichever destinations exist (user may be unaligned, keep never is) */
__device__ __forceinline__ void EmitOpt(char *user, char *keep, size_t v, size_t nvec_full, size_t bytes,
                                        const uint4 &val, bool user_aligned) {
  if (v < nvec_full) {
    if (user != nullptr) {
      if (user_aligned) {
        StoreVec(user + v * kVec, val);
      } else {
        const char *s = reinterpret_cast<const char *>(&val);
        for (int k = 0; k < kVec; ++k) user[v * kVec + k] = s[k];
      }
    }
    if (keep != nullptr) StoreVec(keep + v * kVec, val);
  } else {
    const char *s = reinterpret_cast<const char *>(&val);
    for (size_t i = v * kVec, k = 0; i < bytes; ++i, ++k) {
      if (user != nullptr) user[i] = s[k];
      if (keep != nullptr) keep[i] = s[k];
    }
  }
}

template <typename T, typename OP>
__device__ __forceinline__ void StreamP2PBody(const Params &p, const StreamArgs &a) {
  __shared__ int s_verdict;
  const int me = p.emulate ? blockIdx.y : p.rank;
  const int ix = p.emulate ? me : 0;
  if (Poisoned(p, me)) return;
  char *user = static_cast<char *>(p.user[ix]);
  char *keep = static_cast<char *>(p.keep[ix]);
  const bool aligned = (reinterpret_cast<uintptr_t>(user) % kVec) == 0;
  Stamp(p, 0);
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u, p.block_base}, OpEpoch(p) + 1, p.code[ix], true, &s_verdict, false);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 2);
  const size_t nvec_full = p.bytes / kVec;
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const int n = p.world;
  char *my_res = p.arena[me] + a.res_off;
  // reduce-scatter: the k-th tile of this channel is owned by rank TileOwner(k, n)
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    if (TileOwner(k, n) != me) continue;
    const size_t v0 = tile * tile_vecs + threadIdx.x;
    uint4 acc[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
      if (v < nvec) acc[u] = LoadVec(p.arena[0] + a.src_off + v * kVec);
    }
    // fixed rank order on every owner => every rank ends up with bit-identical results
    for (int r = 1; r < n; ++r) {
      uint4 in[kUnroll];
      const char *src = p.arena[r] + a.src_off;
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
      if (v < nvec) {
        StoreVec(my_res + v * kVec, acc[u]);
        EmitOpt(user, keep, v, nvec_full, p.bytes, acc[u], aligned);
      }
    }
  }
  Stamp(p, 3);
  verdict = FlagBarrier(p, me, MidRow{p.block_base}, OpEpoch(p) + 1, 0u, false, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 4);
  // allgather: pull every tile this rank does not own from its owner's result region
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    const int owner = TileOwner(k, n);
    if (owner == me) continue;
    const char *src = p.arena[owner] + a.res_off;
    const size_t v0 = tile * tile_vecs + threadIdx.x;
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
        if (a.fill_result) StoreVec(my_res + v * kVec, r[u]);
        EmitOpt(user, keep, v, nvec_full, p.bytes, r[u], aligned);
      }
    }
  }
  // a peer may still be pulling from OUR result region: nobody leaves (and lets the next op write into it) early
  verdict = FlagBarrier(p, me, ExitRow{p.block_base}, OpEpoch(p) + 1, 0u, false, &s_verdict, false);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 5);
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
template <typename T, typename OP>
__global__ void __launch_bounds__(512) StreamP2P(Params p, StreamArgs a) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  StreamP2PBody<T, OP>(p, a);
  EndOp(p, me);
}

template <typename MM>
__device__ __forceinline__ void StreamNvlsBody(const Params &p, const StreamArgs &a) {
  __shared__ int s_verdict;
  const int me = p.rank;  // multicast needs real peers: no single-GPU emulation
  if (Poisoned(p, me)) return;
  char *user = static_cast<char *>(p.user[0]);
  char *keep = static_cast<char *>(p.keep[0]);
  const bool aligned = (reinterpret_cast<uintptr_t>(user) % kVec) == 0;
  Stamp(p, 0);
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u, p.block_base}, OpEpoch(p) + 1, p.code[0], true, &s_verdict, false);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 2);
  const size_t nvec_full = p.bytes / kVec;
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kNvlsUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const int n = p.world;
  const size_t res_off = a.res_in_stage ? kCtrlBytes + static_cast<size_t>(OpEpoch(p) & 1u) * p.stage_bytes : a.res_off;
  const char *mc_src = p.mc_arena + a.src_off;
  char *mc_res = p.mc_arena + res_off;
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    if (TileOwner(k, n) != me) continue;
    const size_t v0 = tile * tile_vecs + threadIdx.x;
    uint4 acc[kNvlsUnroll];
#pragma unroll
    for (int u = 0; u < kNvlsUnroll; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
      if (v < nvec) acc[u] = MM::LdReduce(mc_src + v * kVec);
    }
#pragma unroll
    for (int u = 0; u < kNvlsUnroll; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
      if (v < nvec) MultimemStore(mc_res + v * kVec, acc[u]);
    }
  }
  Stamp(p, 3);
  verdict = FlagBarrier(p, me, MidRow{p.block_base}, OpEpoch(p) + 1, 0u, false, &s_verdict);
  if (a.gate) verdict = CommitGate(p, me, verdict, &s_verdict);   // all blocks or none
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 4);
  if (user != nullptr || keep != nullptr) {
    // every tile of this channel now sits, fully reduced, in OUR result region
    const char *src = p.arena[me] + res_off;
    constexpr int kHalf = kNvlsUnroll / 2;   // keeps the register count of the whole kernel low
    for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
#pragma unroll 1
      for (int h = 0; h < 2; ++h) {
        const size_t v0 = tile * tile_vecs + static_cast<size_t>(h) * kHalf * blockDim.x + threadIdx.x;
        uint4 r[kHalf];
#pragma unroll
        for (int u = 0; u < kHalf; ++u) {
          const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
          if (v < nvec) r[u] = LoadVec(src + v * kVec);
        }
        if (aligned) {
#pragma unroll
          for (int u = 0; u < kHalf; ++u) {
            const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
            if (v < nvec_full) {
              if (user != nullptr) StoreVec(user + v * kVec, r[u]);
              if (keep != nullptr) StoreVec(keep + v * kVec, r[u]);
            } else if (v < nvec) {
              EmitOpt(user, keep, v, nvec_full, p.bytes, r[u], true);
            }
          }
        } else {
#pragma unroll 1
          for (int u = 0; u < kHalf; ++u) {
            const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
            if (v < nvec) EmitOpt(user, keep, v, nvec_full, p.bytes, r[u], false);
          }
        }
      }
    }
  }
  Stamp(p, 5);
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
template <typename MM>
__global__ void __launch_bounds__(512) StreamNvls(Params p, StreamArgs a) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  StreamNvlsBody<MM>(p, a);
  EndOp(p, me);
}

/*! \brief tile size the host uses to size the grids of the two kernels above */
constexpr size_t kStreamP2PTile = 512 * kUnroll * kVec;
constexpr size_t kStreamNvlsTile = 512 * kNvlsUnroll * kVec;

}  // namespace device
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_COLL_STREAM_CUH_
