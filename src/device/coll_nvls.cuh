/*!
 * rabit-b200: NVLS (NVLink SHARP) collectives through a multicast mapping of the arenas.
 *
 *   NvlsAllreduce   owner tiles: multimem.ld_reduce pulls the tile ALREADY REDUCED by the
 *                   switch, multimem.st pushes the result into every rank's stage; then each
 *                   rank copies its own stage out at HBM speed. Per GPU only ~S/N bytes enter
 *                   and S/N leave over NVLink instead of 2*(N-1)/N*S for the P2P two-shot.
 *   NvlsBroadcast   root stores its buffer once into the multicast address; the switch fans
 *                   it out (BASELINE config 5: LazyCheckPoint blob / 1 GB model broadcast).
 *
 * Hardware reduction exists for: add on f32/f64/u32/s32/u64(s64), min/max on u32/s32/u64/s64,
 * or on b32/b64. Max/Min on float/double and 1-byte types are NOT expressible and stay on
 * the P2P kernels (dispatch in device_backend_ops.inc). A {double,double} grad/hess pair sum
 * is an f64 add over 2x the elements, i.e. it IS expressible (BASELINE config 3).
 * No counterpart in the reference.
 */
#ifndef RABIT_SRC_DEVICE_COLL_NVLS_CUH_
#define RABIT_SRC_DEVICE_COLL_NVLS_CUH_

#include "coll_kernels.cuh"

namespace rabit {
namespace device {

// ---- one 16-byte vector: switch-side reduce, then multicast store -----------------------
struct MmAddF32 {
  __device__ __forceinline__ static uint4 LdReduce(const void *mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
    return v;
  }
};
#define RABIT_MM_SCALAR64(NAME, INSTR)                                                          \
  struct NAME {                                                                                 \
    __device__ __forceinline__ static uint4 LdReduce(const void *mc) {                          \
      unsigned long long a, b;                                                                  \
      asm volatile("multimem.ld_reduce.relaxed.sys.global." INSTR " %0, [%1];"                 \
                   : "=l"(a) : "l"(mc) : "memory");                                             \
      asm volatile("multimem.ld_reduce.relaxed.sys.global." INSTR " %0, [%1];"                 \
                   : "=l"(b) : "l"(static_cast<const char *>(mc) + 8) : "memory");              \
      uint4 v;                                                                                  \
      v.x = static_cast<unsigned>(a); v.y = static_cast<unsigned>(a >> 32);                     \
      v.z = static_cast<unsigned>(b); v.w = static_cast<unsigned>(b >> 32);                     \
      return v;                                                                                 \
    }                                                                                           \
  }
#define RABIT_MM_SCALAR32(NAME, INSTR)                                                          \
  struct NAME {                                                                                 \
    __device__ __forceinline__ static uint4 LdReduce(const void *mc) {                          \
      uint4 v;                                                                                  \
      const char *c = static_cast<const char *>(mc);                                            \
      asm volatile("multimem.ld_reduce.relaxed.sys.global." INSTR " %0, [%1];" : "=r"(v.x) : "l"(c) : "memory");      \
      asm volatile("multimem.ld_reduce.relaxed.sys.global." INSTR " %0, [%1];" : "=r"(v.y) : "l"(c + 4) : "memory");  \
      asm volatile("multimem.ld_reduce.relaxed.sys.global." INSTR " %0, [%1];" : "=r"(v.z) : "l"(c + 8) : "memory");  \
      asm volatile("multimem.ld_reduce.relaxed.sys.global." INSTR " %0, [%1];" : "=r"(v.w) : "l"(c + 12) : "memory"); \
      return v;                                                                                 \
    }                                                                                           \
  }
RABIT_MM_SCALAR64(MmAddF64, "add.f64");
RABIT_MM_SCALAR64(MmAddU64, "add.u64");
RABIT_MM_SCALAR64(MmMaxU64, "max.u64");
RABIT_MM_SCALAR64(MmMinU64, "min.u64");
RABIT_MM_SCALAR64(MmMaxS64, "max.s64");
RABIT_MM_SCALAR64(MmMinS64, "min.s64");
RABIT_MM_SCALAR64(MmOrB64, "or.b64");
RABIT_MM_SCALAR32(MmAddU32, "add.u32");
RABIT_MM_SCALAR32(MmAddS32, "add.s32");
RABIT_MM_SCALAR32(MmMaxU32, "max.u32");
RABIT_MM_SCALAR32(MmMinU32, "min.u32");
RABIT_MM_SCALAR32(MmMaxS32, "max.s32");
RABIT_MM_SCALAR32(MmMinS32, "min.s32");
RABIT_MM_SCALAR32(MmOrB32, "or.b32");
#undef RABIT_MM_SCALAR64
#undef RABIT_MM_SCALAR32

__device__ __forceinline__ void MultimemStore(void *mc, const uint4 &v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
               ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <typename MM>
__device__ __forceinline__ void NvlsAllreduceBody(const Params &p) {
  __shared__ int s_verdict;
  const int me = p.rank;  // multicast needs real peers: no single-GPU emulation
  char *user = static_cast<char *>(p.user[0]);
  char *keep = static_cast<char *>(p.keep[0]);
  const bool aligned = (reinterpret_cast<uintptr_t>(user) % kVec) == 0;
  Stamp(p, 0);
  CopyTiles<true>(user, StageOf(p, me), p.bytes, aligned);
  Stamp(p, 1);
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[0], true, &s_verdict);
  if (verdict != kStOk) { CommitGate(p, me, verdict, &s_verdict); Report(p, me, verdict); return; }
  Stamp(p, 2);
  const size_t nvec_full = p.bytes / kVec;
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const int n = p.world;
  char *mc_stage = McStage(p);
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    if (TileOwner(k, n) != me) continue;
    const size_t v0 = tile * tile_vecs + threadIdx.x;
    uint4 acc[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
      if (v < nvec) acc[u] = MM::LdReduce(mc_stage + v * kVec);
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
      if (v < nvec) MultimemStore(mc_stage + v * kVec, acc[u]);
    }
  }
  Stamp(p, 3);
  verdict = FlagBarrier(p, me, MidRow{}, OpEpoch(p) + 1, 0u, false, &s_verdict);
  verdict = CommitGate(p, me, verdict, &s_verdict);  // fault-tolerant path: all blocks or none
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 4);
  // every tile of this channel now sits, fully reduced, in OUR stage: local copy-out
  const char *src = StageOf(p, me);
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
  Stamp(p, 5);
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
template <typename MM>
__global__ void __launch_bounds__(512) NvlsAllreduce(Params p) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  NvlsAllreduceBody<MM>(p);
  EndOp(p, me);
}

/*! \brief root: user -> multicast store into everybody's stage; all: barrier, local copy-out */
__device__ __forceinline__ void NvlsBroadcastBody(const Params &p) {
  __shared__ int s_verdict;
  const int me = p.rank;
  char *user = static_cast<char *>(p.user[0]);
  char *keep = static_cast<char *>(p.keep[0]);
  const bool aligned = (reinterpret_cast<uintptr_t>(user) % kVec) == 0;
  const size_t nvec_full = p.bytes / kVec;
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  if (me == p.root) {
    // stage locally first (handles unaligned / ragged user buffers), then fan out from the stage
    CopyTiles<true>(user, StageOf(p, me), p.bytes, aligned);
    __syncthreads();
    char *mc_stage = McStage(p);
    const char *src = StageOf(p, me);
    for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const size_t v0 = tile * tile_vecs + threadIdx.x;
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) MultimemStore(mc_stage + v * kVec, LoadVec(src + v * kVec));
      }
    }
  }
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[0], true, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  if (me == p.root && keep == nullptr) return;
  const char *src = StageOf(p, me);
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
__global__ void __launch_bounds__(512) NvlsBroadcast(Params p) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  NvlsBroadcastBody(p);
  EndOp(p, me);
}

}  // namespace device
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_COLL_NVLS_CUH_
