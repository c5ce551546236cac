/*!
 * rabit-b200: zero-copy collectives on SYMMETRIC buffers.
 *
 * A buffer obtained from RabitDeviceSymAlloc lives inside the arena heap, at the same offset
 * on every rank, so peers can address it directly: no staging copy in, no copy out.
 *
 *   TwoShotInPlace  barrier -> owner pulls its tiles from every rank's buffer and writes the
 *                   reduction back into ITS OWN buffer -> barrier -> everybody pulls the tiles
 *                   it does not own from their owners -> exit barrier (a rank must not start
 *                   refilling its buffer while a peer still reads it)
 *   NvlsInPlace     barrier -> owner: multimem.ld_reduce on the tile, multimem.st of the result
 *                   into every rank's buffer -> barrier. Nothing else: the switch did the
 *                   reduction and the fan-out, each byte crossed NVLink about once each way.
 * The optional `keep` copy (result cache) is written from registers in the same pass.
 * No counterpart in the reference.
 */
#ifndef RABIT_SRC_DEVICE_COLL_INPLACE_CUH_
#define RABIT_SRC_DEVICE_COLL_INPLACE_CUH_

#include "coll_kernels.cuh"
#include "coll_nvls.cuh"

namespace rabit {
namespace device {

struct ExitRow {
  __device__ __forceinline__ unsigned long long *operator()(Control *c) const {
    return &c->exit_[blockIdx.x][0];
  }
};

/*! \brief rank r's copy of the symmetric buffer */
__device__ __forceinline__ char *SymOf(const Params &p, int r) { return p.arena[r] + p.sym_offset; }

template <typename T, typename OP>
__global__ void __launch_bounds__(512) TwoShotInPlace(Params p) {
  __shared__ int s_verdict;
  const int me = p.rank;
  char *mine = SymOf(p, me);
  char *keep = static_cast<char *>(p.keep[0]);
  Stamp(p, 0);
  int verdict = FlagBarrier(p, me, EntryRow{p.epoch & 1u}, p.epoch + 1, p.code[0], true, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 2);
  // symmetric buffers are allocated in 256-byte units, so the padded last vector is addressable
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const int n = p.world;
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    if (static_cast<int>(k % n) != me) continue;
    const size_t v0 = tile * tile_vecs + threadIdx.x;
    uint4 acc[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
      if (v < nvec) acc[u] = LoadVec(SymOf(p, 0) + v * kVec);
    }
    for (int r = 1; r < n; ++r) {
      uint4 in[kUnroll];
      const char *src = SymOf(p, r);
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
        StoreVec(mine + v * kVec, acc[u]);
        if (keep != nullptr) StoreVec(keep + v * kVec, acc[u]);
      }
    }
  }
  Stamp(p, 3);
  verdict = FlagBarrier(p, me, MidRow{}, p.epoch + 1, 0u, false, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 4);
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    const int owner = static_cast<int>(k % n);
    if (owner == me) continue;
    const char *src = SymOf(p, owner);
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
        StoreVec(mine + v * kVec, r[u]);
        if (keep != nullptr) StoreVec(keep + v * kVec, r[u]);
      }
    }
  }
  verdict = FlagBarrier(p, me, ExitRow{}, p.epoch + 1, 0u, false, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 5);
}

/*! \brief switch-side reductions have a longer round trip than peer loads: keep 8 vectors per thread in flight */
constexpr int kNvlsUnroll = 8;

template <typename MM>
__global__ void __launch_bounds__(512) NvlsInPlace(Params p) {
  __shared__ int s_verdict;
  const int me = p.rank;
  char *keep = static_cast<char *>(p.keep[0]);
  char *mc = p.mc_arena + p.sym_offset;
  Stamp(p, 0);
  int verdict = FlagBarrier(p, me, EntryRow{p.epoch & 1u}, p.epoch + 1, p.code[0], true, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 2);
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kNvlsUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const int n = p.world;
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    if (static_cast<int>(k % n) != me) continue;
    const size_t v0 = tile * tile_vecs + threadIdx.x;
    uint4 acc[kNvlsUnroll];
#pragma unroll
    for (int u = 0; u < kNvlsUnroll; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
      if (v < nvec) acc[u] = MM::LdReduce(mc + v * kVec);
    }
#pragma unroll
    for (int u = 0; u < kNvlsUnroll; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
      if (v < nvec) MultimemStore(mc + v * kVec, acc[u]);
    }
  }
  Stamp(p, 3);
  verdict = FlagBarrier(p, me, MidRow{}, p.epoch + 1, 0u, false, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 4);
  if (keep != nullptr) {
    const char *src = SymOf(p, me);
    for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const size_t v0 = tile * tile_vecs + threadIdx.x;
#pragma unroll
      for (int u = 0; u < kNvlsUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) StoreVec(keep + v * kVec, LoadVec(src + v * kVec));
      }
    }
  }
  Stamp(p, 5);
}

}  // namespace device
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_COLL_INPLACE_CUH_
