/*!
 * rabit-b200: the collective kernels (LDG/STG flavour). See coll_common.cuh for the model.
 *
 * Work mapping shared by all kernels: the payload is cut into tiles of blockDim.x * kUnroll
 * 16-byte vectors; tile t belongs to channel (block) t % gridDim.x on EVERY rank, and inside a
 * channel the k-th tile is "owned" (reduced) by rank k % world. So block b of rank r depends
 * only on what block b of the other ranks wrote.
 *
 *   OneShotAllreduce  stage -> barrier -> every rank reduces every tile from all stages
 *                     (latency path; replaces the reference's tree allreduce,
 *                      allreduce_base.cc:476-668)
 *   TwoShotAllreduce  stage -> barrier -> owner reduces its tiles in place -> barrier ->
 *                     everybody gathers the reduced tiles (bandwidth path; replaces ring
 *                     reduce-scatter + allgather, allreduce_base.cc:857-977)
 *   Broadcast         root stages, peers pull   (allreduce_base.cc:677-765)
 *   Allgather         everybody stages its slice, peers pull (allreduce_base.cc:779-843)
 * Every kernel writes the result to `user` and, when given, also to `keep` (the result
 * cache entry of the robust engine: reference allreduce_robust.cc:279-283 does this with
 * two host memcpys per op).
 */
#ifndef RABIT_SRC_DEVICE_COLL_KERNELS_CUH_
#define RABIT_SRC_DEVICE_COLL_KERNELS_CUH_

#include "coll_common.cuh"

namespace rabit {
namespace device {

// `base` shifts a kernel onto its own set of flag rows: two chunk kernels of one streamed op that run at the same
// time (device_backend_stream.inc, "lanes") must not share rows
struct EntryRow {
  unsigned int par;
  unsigned int base = 0;
  __device__ __forceinline__ unsigned long long *operator()(Control *c) const {
    return &c->entry[par][blockIdx.x + base][0];
  }
};
struct MidRow {
  unsigned int base = 0;
  __device__ __forceinline__ unsigned long long *operator()(Control *c) const {
    return &c->mid[blockIdx.x + base][0];
  }
};

/*! \brief store a result vector to the user buffer and the optional cache copy */
__device__ __forceinline__ void Emit(char *user, char *keep, size_t v, size_t nvec_full, size_t bytes,
                                     const uint4 &val, bool user_aligned) {
  if (v < nvec_full) {
    if (user_aligned) {
      StoreVec(user + v * kVec, val);
    } else {
      const char *s = reinterpret_cast<const char *>(&val);
      for (int k = 0; k < kVec; ++k) user[v * kVec + k] = s[k];
    }
    if (keep != nullptr) StoreVec(keep + v * kVec, val);
  } else {
    // the padded last vector: only the real bytes leave the stage
    const char *s = reinterpret_cast<const char *>(&val);
    for (size_t i = v * kVec, k = 0; i < bytes; ++i, ++k) {
      user[i] = s[k];
      if (keep != nullptr) keep[i] = s[k];
    }
  }
}

template <typename T, typename OP>
__device__ __forceinline__ void OneShotAllreduceBody(const Params &p) {
  __shared__ int s_verdict;
  const int me = p.emulate ? blockIdx.y : p.rank;
  char *user = static_cast<char *>(p.user[p.emulate ? me : 0]);
  char *keep = static_cast<char *>(p.keep[p.emulate ? me : 0]);
  const bool aligned = (reinterpret_cast<uintptr_t>(user) % kVec) == 0;
  Stamp(p, 0);
  CopyTiles<true>(user, StageOf(p, me), p.bytes, aligned);
  Stamp(p, 1);
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[p.emulate ? me : 0], true, &s_verdict);
  verdict = CommitGate(p, me, verdict, &s_verdict);  // fault-tolerant path: all blocks or none
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 2);

  const size_t nvec_full = p.bytes / kVec;
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const int n = p.world;
  for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const size_t v0 = tile * tile_vecs + threadIdx.x;
    uint4 acc[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
      if (v < nvec) acc[u] = LoadVec(StageOf(p, 0) + v * kVec);
    }
    // fixed rank order 0..n-1 on every rank => bit-identical results everywhere
    for (int r = 1; r < n; ++r) {
      uint4 in[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) in[u] = LoadVec(StageOf(p, r) + v * kVec);
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
      if (v < nvec) Emit(user, keep, v, nvec_full, p.bytes, acc[u], aligned);
    }
  }
  Stamp(p, 5);
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
template <typename T, typename OP>
__global__ void __launch_bounds__(512) OneShotAllreduce(Params p) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  OneShotAllreduceBody<T, OP>(p);
  EndOp(p, me);
}

template <typename T, typename OP>
__device__ __forceinline__ void TwoShotAllreduceBody(const Params &p) {
  __shared__ int s_verdict;
  const int me = p.emulate ? blockIdx.y : p.rank;
  char *user = static_cast<char *>(p.user[p.emulate ? me : 0]);
  char *keep = static_cast<char *>(p.keep[p.emulate ? me : 0]);
  const bool aligned = (reinterpret_cast<uintptr_t>(user) % kVec) == 0;
  Stamp(p, 0);
  CopyTiles<true>(user, StageOf(p, me), p.bytes, aligned);
  Stamp(p, 1);
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[p.emulate ? me : 0], true, &s_verdict);
  if (verdict != kStOk) { CommitGate(p, me, verdict, &s_verdict); Report(p, me, verdict); return; }
  Stamp(p, 2);

  const size_t nvec_full = p.bytes / kVec;
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const int n = p.world;
  char *my_stage = StageOf(p, me);
  // reduce-scatter: the k-th tile of this channel is owned by rank TileOwner(k, n)
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    if (TileOwner(k, n) != me) continue;
    const size_t v0 = tile * tile_vecs + threadIdx.x;
    uint4 acc[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
      if (v < nvec) acc[u] = LoadVec(StageOf(p, 0) + v * kVec);
    }
    for (int r = 1; r < n; ++r) {
      uint4 in[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) in[u] = LoadVec(StageOf(p, r) + v * kVec);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) ReduceVec<T, OP>(&acc[u], in[u]);
      }
    }
    // in place: peers never read OUR stage at tiles WE own during this phase
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
      if (v < nvec) StoreVec(my_stage + v * kVec, acc[u]);
    }
  }
  Stamp(p, 3);
  verdict = FlagBarrier(p, me, MidRow{}, OpEpoch(p) + 1, 0u, false, &s_verdict);
  verdict = CommitGate(p, me, verdict, &s_verdict);  // fault-tolerant path: all blocks or none
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 4);
  // allgather: pull every tile from its owner's stage
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    const char *src = StageOf(p, TileOwner(k, n));
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
      if (v < nvec) Emit(user, keep, v, nvec_full, p.bytes, r[u], aligned);
    }
  }
  Stamp(p, 5);
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
template <typename T, typename OP>
__global__ void __launch_bounds__(512) TwoShotAllreduce(Params p) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  TwoShotAllreduceBody<T, OP>(p);
  EndOp(p, me);
}

/*! \brief root stages its buffer; everyone (root included, for the cache copy) pulls it */
__device__ __forceinline__ void BroadcastPullBody(const Params &p) {
  __shared__ int s_verdict;
  const int me = p.emulate ? blockIdx.y : p.rank;
  char *user = static_cast<char *>(p.user[p.emulate ? me : 0]);
  char *keep = static_cast<char *>(p.keep[p.emulate ? me : 0]);
  const bool aligned = (reinterpret_cast<uintptr_t>(user) % kVec) == 0;
  if (me == p.root) CopyTiles<true>(user, StageOf(p, me), p.bytes, aligned);
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[p.emulate ? me : 0], true, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  if (me == p.root && keep == nullptr) return;
  const size_t nvec_full = p.bytes / kVec;
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const char *src = StageOf(p, p.root);
  for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
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
      if (v < nvec) Emit(user, keep, v, nvec_full, p.bytes, r[u], aligned);
    }
  }
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
__global__ void __launch_bounds__(512) BroadcastPull(Params p) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  BroadcastPullBody(p);
  EndOp(p, me);
}

/*!
 * \brief copy the bytes of [lo, hi) that fall into THIS block's tiles from src to dst (and dst2).
 *  Offsets are positions in the conceptual total buffer; all three pointers are bases of that
 *  buffer. Whole 16-byte vectors move as vectors, the ragged edges byte by byte.
 */
__device__ __forceinline__ void CopyRangeTiles(char *dst, char *dst2, const char *src, size_t lo, size_t hi,
                                               size_t total, bool vec_ok) {
  if (lo >= hi) return;
  const size_t nvec = (total + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t first_tile = (lo / kVec) / tile_vecs, last_tile = ((hi - 1) / kVec) / tile_vecs;
  // first tile of this block at or after first_tile
  size_t tile = first_tile + ((blockIdx.x + gridDim.x - first_tile % gridDim.x) % gridDim.x);
  for (; tile <= last_tile; tile += gridDim.x) {
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t v = tile * tile_vecs + threadIdx.x + static_cast<size_t>(u) * blockDim.x;
      if (v >= nvec) continue;
      const size_t b0 = v * kVec, b1 = b0 + kVec;
      if (b1 <= lo || b0 >= hi) continue;
      if (vec_ok && b0 >= lo && b1 <= hi) {
        const uint4 x = LoadVec(src + b0);
        StoreVec(dst + b0, x);
        if (dst2 != nullptr) StoreVec(dst2 + b0, x);
      } else {
        const size_t e0 = b0 > lo ? b0 : lo, e1 = b1 < hi ? b1 : hi;
        for (size_t i = e0; i < e1; ++i) {
          const char c = src[i];
          dst[i] = c;
          if (dst2 != nullptr) dst2[i] = c;
        }
      }
    }
  }
}

/*!
 * \brief every rank stages its own byte range [slice_begin, slice_end) of the total buffer at
 *  the same offsets of its stage, then pulls everybody else's range. Ranges are arbitrary byte
 *  ranges in the API (reference allreduce_base.cc:779-843), and a rank only knows its own, so
 *  each block also writes (begin, end) into a 16-byte stage header before the barrier.
 */
__device__ __forceinline__ void AllgatherPullBody(const Params &p) {
  __shared__ int s_verdict;
  const int me = p.emulate ? blockIdx.y : p.rank;
  const int ix = p.emulate ? me : 0;
  char *user = static_cast<char *>(p.user[ix]);
  char *keep = static_cast<char *>(p.keep[ix]);
  const size_t begin = p.slice_begin[ix], end = p.slice_end[ix];
  const bool vec_ok = (reinterpret_cast<uintptr_t>(user) % kVec) == 0;
  char *stage = StageOf(p, me);
  if (threadIdx.x == 0) {  // every block writes the same header: its own barrier then covers it
    reinterpret_cast<volatile unsigned long long *>(stage)[0] = begin;
    reinterpret_cast<volatile unsigned long long *>(stage)[1] = end;
  }
  CopyRangeTiles(stage + kVec, nullptr, user, begin, end, p.bytes, vec_ok);
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[ix], true, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  const int n = p.world;
  for (int r = 0; r < n; ++r) {
    const char *src = StageOf(p, r);
    const size_t b = reinterpret_cast<const volatile unsigned long long *>(src)[0];
    const size_t e = reinterpret_cast<const volatile unsigned long long *>(src)[1];
    if (r == me) {
      if (keep != nullptr) CopyRangeTiles(keep, nullptr, user, b, e, p.bytes, vec_ok);
    } else {
      CopyRangeTiles(user, keep, src + kVec, b, e, p.bytes, vec_ok);
    }
  }
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
__global__ void __launch_bounds__(512) AllgatherPull(Params p) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  AllgatherPullBody(p);
  EndOp(p, me);
}

}  // namespace device
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_COLL_KERNELS_CUH_
