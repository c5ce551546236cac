/*!
 * rabit-b200: two-shot allreduce whose reduce-scatter phase is fed by the TMA engine.
 *
 * Same channel/ownership mapping and the same two flag barriers as TwoShotAllreduce, but the
 * owner does not issue per-thread peer loads: one elected thread posts `cp.async.bulk`
 * copies (SASS: UBLKCP) that pull the tile from EVERY rank's stage over NVLink into a
 * shared-memory ring guarded by mbarriers (expect_tx / complete_tx); all threads then
 * reduce out of shared memory and store the result into the local stage. With kDepth
 * stages in flight an SM keeps world * kSubTile * (kDepth-1) bytes of NVLink reads
 * outstanding without spending registers on them.
 * No counterpart in the reference (its ring reduce-scatter is allreduce_base.cc:857-946).
 */
#ifndef RABIT_SRC_DEVICE_COLL_TMA_CUH_
#define RABIT_SRC_DEVICE_COLL_TMA_CUH_

#include "coll_kernels.cuh"

namespace rabit {
namespace device {

constexpr int kTmaThreads = 512;
constexpr int kSubTile = kTmaThreads * kVec;  // 8 KiB per rank per pipeline stage

__device__ __forceinline__ uint32_t SmemAddr(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void MbarInit(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(SmemAddr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void MbarExpectTx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(SmemAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool MbarTryWait(uint64_t *bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(SmemAddr(bar)), "r"(parity) : "memory");
  return ok != 0;
}
/*! \brief global -> shared bulk copy, completion signalled on an mbarrier (bytes % 16 == 0) */
__device__ __forceinline__ void BulkLoad(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(SmemAddr(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(SmemAddr(bar)) : "memory");
}

/*! \brief dynamic shared memory: [kDepth mbarriers][kDepth][world][kSubTile] */
template <typename T, typename OP>
__device__ __forceinline__ void TwoShotAllreduceTmaBody(const Params &p, int depth) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ int s_verdict;
  uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw);
  unsigned char *ring = smem_raw + 128;
  const int me = p.emulate ? blockIdx.y : p.rank;
  char *user = static_cast<char *>(p.user[p.emulate ? me : 0]);
  char *keep = static_cast<char *>(p.keep[p.emulate ? me : 0]);
  const bool aligned = (reinterpret_cast<uintptr_t>(user) % kVec) == 0;
  const int n = p.world;
  if (threadIdx.x == 0) {
    for (int s = 0; s < depth; ++s) MbarInit(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  Stamp(p, 0);
  CopyTiles<true>(user, StageOf(p, me), p.bytes, aligned);
  Stamp(p, 1);
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[p.emulate ? me : 0], true, &s_verdict);
  if (verdict != kStOk) { CommitGate(p, me, verdict, &s_verdict); Report(p, me, verdict); return; }
  Stamp(p, 2);

  const size_t nvec_full = p.bytes / kVec;
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t padded = nvec * kVec;
  const size_t tile_bytes = static_cast<size_t>(blockDim.x) * kUnroll * kVec;
  const size_t ntiles = (padded + tile_bytes - 1) / tile_bytes;
  char *my_stage = StageOf(p, me);

  // enumerate the sub-tiles this rank owns in this channel: sub-tile q -> (byte offset, length)
  const size_t chan_tiles = ntiles > blockIdx.x ? (ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
  const size_t owned_tiles = chan_tiles > static_cast<size_t>(me) ? (chan_tiles - me + n - 1) / n : 0;
  const int subs_per_tile = static_cast<int>(tile_bytes / kSubTile);
  const size_t nsub = owned_tiles * subs_per_tile;
  auto sub_range = [&](size_t q, size_t *off, uint32_t *len) {
    const size_t k = me + (q / subs_per_tile) * n;              // k-th tile of the channel
    const size_t tile = blockIdx.x + k * gridDim.x;
    const size_t o = tile * tile_bytes + (q % subs_per_tile) * static_cast<size_t>(kSubTile);
    *off = o;
    *len = o >= padded ? 0u : static_cast<uint32_t>(padded - o < static_cast<size_t>(kSubTile) ? padded - o : kSubTile);
  };
  auto post = [&](size_t q) {  // thread 0 only
    size_t off; uint32_t len;
    sub_range(q, &off, &len);
    const int s = static_cast<int>(q % depth);
    if (len == 0) { MbarExpectTx(&full[s], 0); return; }
    MbarExpectTx(&full[s], len * n);
    for (int r = 0; r < n; ++r) {
      BulkLoad(ring + (static_cast<size_t>(s) * n + r) * kSubTile, StageOf(p, r) + off, len, &full[s]);
    }
  };
  if (threadIdx.x == 0) {
    for (size_t q = 0; q < nsub && q + 1 < static_cast<size_t>(depth); ++q) post(q);
  }
  for (size_t q = 0; q < nsub; ++q) {
    // the stage being refilled was consumed in iteration q-1 (block-wide sync at its end)
    if (threadIdx.x == 0 && q + depth - 1 < nsub) post(q + depth - 1);
    const int s = static_cast<int>(q % depth);
    const uint32_t parity = static_cast<uint32_t>((q / depth) & 1);
    while (!MbarTryWait(&full[s], parity)) {}
    size_t off; uint32_t len;
    sub_range(q, &off, &len);
    const uint32_t my_off = threadIdx.x * kVec;
    if (my_off < len) {
      const unsigned char *base = ring + static_cast<size_t>(s) * n * kSubTile + my_off;
      uint4 acc = *reinterpret_cast<const uint4 *>(base);
      for (int r = 1; r < n; ++r) {
        ReduceVec<T, OP>(&acc, *reinterpret_cast<const uint4 *>(base + static_cast<size_t>(r) * kSubTile));
      }
      StoreVec(my_stage + off + my_off, acc);
    }
    __syncthreads();
  }
  Stamp(p, 3);
  verdict = FlagBarrier(p, me, MidRow{}, OpEpoch(p) + 1, 0u, false, &s_verdict);
  verdict = CommitGate(p, me, verdict, &s_verdict);  // fault-tolerant path: all blocks or none
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 4);
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    const char *src = StageOf(p, static_cast<int>(k % n));
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
__global__ void __launch_bounds__(kTmaThreads) TwoShotAllreduceTma(Params p, int depth) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  TwoShotAllreduceTmaBody<T, OP>(p, depth);
  EndOp(p, me);
}

/*! \brief dynamic shared memory the TMA kernel needs for a given world and depth */
inline size_t TmaSmemBytes(int world, int depth) {
  return 128 + static_cast<size_t>(depth) * world * kSubTile;
}

}  // namespace device
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_COLL_TMA_CUH_
