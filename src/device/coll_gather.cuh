/*!
 * rabit-b200: allgather through the switch. Every rank multicast-stores its own byte range once
 * (multimem.st: one NVLink egress of S/N per GPU, the switch fans it out into every rank's stage), then
 * copies the other ranks' ranges out of its own stage at HBM speed. The pull kernel (coll_kernels.cuh)
 * moves (N-1)/N * S per GPU through peer loads instead; it stays the choice for small messages (one
 * barrier instead of two) and for boxes without a multicast mapping.
 * Byte ranges are arbitrary (reference ring allgather: allreduce_base.cc:779-843): whole 16-byte vectors go
 * through the multicast mapping, the ragged edges of a range are written byte-wise into every peer.
 * Stage layout: [kGatherHdr bytes of (begin, end) per rank | the conceptual buffer at its own offsets].
 */
#ifndef RABIT_SRC_DEVICE_COLL_GATHER_CUH_
#define RABIT_SRC_DEVICE_COLL_GATHER_CUH_

#include "coll_nvls.cuh"

namespace rabit {
namespace device {

constexpr size_t kGatherHdr = 256;   // 16 ranks x {u64 begin, u64 end}

/*! \brief 16 bytes from a possibly unaligned address */
__device__ __forceinline__ uint4 LoadVecAny(const char *ptr, bool aligned) {
  if (aligned) return LoadVec(ptr);
  uint4 v;
  char *d = reinterpret_cast<char *>(&v);
  for (int k = 0; k < kVec; ++k) d[k] = ptr[k];
  return v;
}
__device__ __forceinline__ void StoreVecAny(char *ptr, const uint4 &v, bool aligned) {
  if (aligned) { StoreVec(ptr, v); return; }
  const char *s = reinterpret_cast<const char *>(&v);
  for (int k = 0; k < kVec; ++k) ptr[k] = s[k];
}

/*!
 * \brief inplace != 0: the buffer is symmetric (same arena offset on every rank, p.sym_offset): every rank multicasts
 *  its range straight into everybody's BUFFER. No stage, no header, no copy-out: each byte crosses NVLink once and
 *  is written once. (Ragged range edges still go byte-wise into every peer.)
 */
__device__ __forceinline__ void AllgatherMcBody(const Params &p, int inplace) {
  __shared__ int s_verdict;
  const int me = p.rank, n = p.world;
  char *user = static_cast<char *>(p.user[0]);
  char *keep = static_cast<char *>(p.keep[0]);
  const size_t begin = p.slice_begin[0], end = p.slice_end[0];
  const bool aligned = (reinterpret_cast<uintptr_t>(user) % kVec) == 0;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  // nobody may still be reading the previous contents of the stages (or, in place, of its buffer)
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[0], true, &s_verdict, false);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  if (inplace) {
    char *mc_buf = p.mc_arena + p.sym_offset;
    if (begin < end) {
      const size_t vb = begin / kVec, ve = (end + kVec - 1) / kVec;
      const size_t ntiles = (ve - vb + tile_vecs - 1) / tile_vecs;
      for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
          const size_t v = vb + tile * tile_vecs + threadIdx.x + static_cast<size_t>(u) * blockDim.x;
          if (v >= ve) continue;
          const size_t b0 = v * kVec, b1 = b0 + kVec;
          if (b0 >= begin && b1 <= end) {
            MultimemStore(mc_buf + b0, LoadVec(user + b0));   // symmetric buffers are 256-byte aligned
          } else {
            const size_t e0 = b0 > begin ? b0 : begin, e1 = b1 < end ? b1 : end;
            for (size_t i = e0; i < e1; ++i) {
              const char c = user[i];
              for (int r = 0; r < n; ++r) if (r != me) (p.arena[r] + p.sym_offset)[i] = c;
            }
          }
        }
      }
    }
    verdict = FlagBarrier(p, me, MidRow{}, OpEpoch(p) + 1, 0u, false, &s_verdict);
    if (verdict != kStOk) { Report(p, me, verdict); return; }
    // (A result-cache copy is NOT taken here: a block only knows that the tiles pushed by the peers' same-numbered
    // blocks are complete, and the peers' ranges are not known in this variant. The host enqueues a device copy of
    // the whole buffer behind this kernel instead.)
    return;
  }
  char *mc = McStage(p);
  if (threadIdx.x == 0) {
    // every block announces the range (same words): a peer's block b learns it through ITS barrier with our block b
    uint4 h;
    h.x = static_cast<unsigned>(begin); h.y = static_cast<unsigned>(begin >> 32);
    h.z = static_cast<unsigned>(end); h.w = static_cast<unsigned>(end >> 32);
    MultimemStore(mc + static_cast<size_t>(me) * kVec, h);
  }
  if (begin < end) {
    const size_t vb = begin / kVec, ve = (end + kVec - 1) / kVec;   // vectors that intersect our range
    const size_t ntiles = (ve - vb + tile_vecs - 1) / tile_vecs;
    for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = vb + tile * tile_vecs + threadIdx.x + static_cast<size_t>(u) * blockDim.x;
        if (v >= ve) continue;
        const size_t b0 = v * kVec, b1 = b0 + kVec;
        if (b0 >= begin && b1 <= end) {
          MultimemStore(mc + kGatherHdr + b0, LoadVecAny(user + b0, aligned));
        } else {
          // ragged edge: only our bytes, straight into every rank's stage
          const size_t e0 = b0 > begin ? b0 : begin, e1 = b1 < end ? b1 : end;
          for (size_t i = e0; i < e1; ++i) {
            const char c = user[i];
            for (int r = 0; r < n; ++r) StageOf(p, r)[kGatherHdr + i] = c;
          }
        }
      }
    }
  }
  verdict = FlagBarrier(p, me, MidRow{}, OpEpoch(p) + 1, 0u, false, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  // copy-out: of every other rank's range, the tiles that ITS block b pushed (those are the ones our barrier covers)
  const char *stage = StageOf(p, me);
  for (int r = 0; r < n; ++r) {
    size_t rb, re;
    if (r == me) {
      if (keep == nullptr) continue;
      rb = begin; re = end;
    } else {
      rb = reinterpret_cast<const volatile unsigned long long *>(stage + static_cast<size_t>(r) * kVec)[0];
      re = reinterpret_cast<const volatile unsigned long long *>(stage + static_cast<size_t>(r) * kVec)[1];
    }
    if (rb >= re || re > p.bytes) continue;
    const char *src = r == me ? user : stage + kGatherHdr;
    const bool src_aligned = r == me ? aligned : true;
    const size_t vb = rb / kVec, ve = (re + kVec - 1) / kVec;
    const size_t ntiles = (ve - vb + tile_vecs - 1) / tile_vecs;
    for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = vb + tile * tile_vecs + threadIdx.x + static_cast<size_t>(u) * blockDim.x;
        if (v >= ve) continue;
        const size_t b0 = v * kVec, b1 = b0 + kVec;
        if (b0 >= rb && b1 <= re) {
          const uint4 x = LoadVecAny(src + b0, src_aligned);
          if (r != me) StoreVecAny(user + b0, x, aligned);
          if (keep != nullptr) StoreVec(keep + b0, x);
        } else {
          const size_t e0 = b0 > rb ? b0 : rb, e1 = b1 < re ? b1 : re;
          for (size_t i = e0; i < e1; ++i) {
            const char c = src[i];
            if (r != me) user[i] = c;
            if (keep != nullptr) keep[i] = c;
          }
        }
      }
    }
  }
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
__global__ void __launch_bounds__(512) AllgatherMc(Params p, int inplace) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  AllgatherMcBody(p, inplace);
  EndOp(p, me);
}

}  // namespace device
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_COLL_GATHER_CUH_
