/*!
 * rabit-b200: low-latency allreduce for small messages (<= kLLMaxBytes = 32 KiB per rank).
 *
 * The one-shot kernel needs a flag barrier (release fences + a round trip) after staging. For a
 * few hundred bytes that barrier IS the cost. Here data and flag travel in the same 8-byte word:
 * every rank pushes each 4-byte word of its contribution, tagged with the op's epoch, straight
 * into its slot of EVERY peer's LL area; a reader simply spins on a word until the tag matches.
 * 8-byte aligned stores are single transactions on NVLink, so a matching tag means the data half
 * is there too — no fences, no separate barrier, one NVLink one-way latency in total.
 * (Same idea as NCCL's LL protocol; 2x wire overhead, irrelevant at this size.)
 *
 * The consensus code rides in one extra tagged word per rank, so the fused consensus costs
 * nothing here either; a mismatch leaves the user buffer untouched -> kStMismatch.
 * Slots alternate with the op parity exactly like the stages do.
 * This is the device form of the reference's tiny tree allreduces (consensus words, scalars:
 * allreduce_robust.cc:1064,1178,1239 and allreduce_base.cc:476-668 for small counts).
 */
#ifndef RABIT_SRC_DEVICE_COLL_LL_CUH_
#define RABIT_SRC_DEVICE_COLL_LL_CUH_

#include "coll_kernels.cuh"

namespace rabit {
namespace device {

constexpr size_t kLLMaxBytes = 32768;                                  // payload per rank
constexpr size_t kLLSlotWords = kLLMaxBytes / 4 + 2;                   // + the consensus word (padded)
constexpr size_t kLLSlotBytes = kLLSlotWords * 8;                      // every word carries a tag
constexpr size_t kLLBytes = 2 * kMaxRanks * kLLSlotBytes;              // two parities x one slot per source rank

/*! \brief rank r's LL area: [parity][source rank][word] of {value, tag} */
__device__ __forceinline__ uint2 *LLSlot(const Params &p, int r, int src) {
  char *base = p.arena[r] + kCtrlBytes + 2 * p.stage_bytes;
  return reinterpret_cast<uint2 *>(base) + (static_cast<size_t>(OpEpoch(p) & 1u) * kMaxRanks + src) * kLLSlotWords;
}
__device__ __forceinline__ void StoreTagged(uint2 *addr, unsigned int value, unsigned int tag) {
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(addr), "r"(value), "r"(tag) : "memory");
}
__device__ __forceinline__ uint2 LoadTagged(const uint2 *addr) {
  uint2 v;
  asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(addr) : "memory");
  return v;
}

/*! \brief spin until the word carries `tag`; false on abort / timeout */
__device__ __forceinline__ bool WaitTagged(const Params &p, const uint2 *addr, unsigned int tag, unsigned int *value,
                                           int *verdict) {
  const unsigned long long deadline = GlobalTimer() + p.timeout_ns;
  unsigned int spins = 0;
  while (true) {
    const uint2 v = LoadTagged(addr);
    if (v.y == tag) { *value = v.x; return true; }
    if ((++spins & 0x3ff) == 0) {
      if (*p.abort_flag != 0) { *verdict = kStAborted; return false; }
      if (GlobalTimer() > deadline) { *verdict = kStTimeout; return false; }
    }
  }
}

template <typename T, typename OP>
__device__ __forceinline__ void OneShotLLBody(const Params &p) {
  __shared__ int s_verdict;
  const int me = p.emulate ? blockIdx.y : p.rank;
  const int ix = p.emulate ? me : 0;
  const int n = p.world;
  unsigned int *user = static_cast<unsigned int *>(p.user[ix]);
  unsigned int *keep = static_cast<unsigned int *>(p.keep[ix]);
  // never 0 (zeroed or never-written slot words must not look tagged), distinct for the two ops that share a parity
  const unsigned int tag = (OpEpoch(p) & 0x7fffffffu) + 1;
  const unsigned int words = static_cast<unsigned int>(p.bytes / 4);
  constexpr unsigned int kUnitWords = sizeof(T) > 4 ? 2 : 1;   // an element never straddles two threads
  const unsigned int units = words / kUnitWords;
  if (threadIdx.x == 0) s_verdict = kStOk;
  __syncthreads();
  // the grid splits the units evenly; block b of every rank works on the same range
  const unsigned int per_block = (units + gridDim.x - 1) / gridDim.x;
  const unsigned int u_begin = blockIdx.x * per_block;
  const unsigned int u_end = u_begin + per_block < units ? u_begin + per_block : units;
  // push: my words into slot `me` of every rank (own copy included, so that the reduce below is uniform)
  for (unsigned int w = u_begin * kUnitWords + threadIdx.x; w < u_end * kUnitWords; w += blockDim.x) {
    const unsigned int v = user[w];
    for (int q = 0; q < n; ++q) StoreTagged(LLSlot(p, q, me) + w, v, tag);
  }
  if (blockIdx.x == 0 && threadIdx.x < n) StoreTagged(LLSlot(p, threadIdx.x, me) + words, p.code[ix], tag);
  // consensus: everybody must be asking for the same thing
  int verdict = kStOk;
  if (threadIdx.x < n) {
    unsigned int code = 0;
    if (WaitTagged(p, LLSlot(p, me, threadIdx.x) + words, tag, &code, &verdict)) {
      if (code != p.code[ix]) verdict = kStMismatch;
    }
    if (verdict != kStOk) atomicMax(&s_verdict, verdict);
  }
  __syncthreads();
  if (s_verdict != kStOk || per_block > blockDim.x) {
    const int v = s_verdict != kStOk ? s_verdict : kStAborted;  // (a launch with too few threads is a host bug)
    __syncthreads();
    CommitGate(p, me, v, &s_verdict);
    Report(p, me, v);
    return;
  }
  // pull + reduce in rank order (identical on every rank => bit-identical results). The launch gives every
  // thread at most ONE unit, so the result waits in registers until the whole grid is known to be complete.
  const unsigned int u = u_begin + threadIdx.x;
  const bool mine = u < u_end;
  alignas(8) unsigned int acc[kUnitWords], in[kUnitWords];  // reinterpreted as (up to) 8-byte T below
  if (mine) {
    bool ok = true;
    for (int q = 0; q < n && ok; ++q) {
      for (unsigned int k = 0; k < kUnitWords && ok; ++k) {
        ok = WaitTagged(p, LLSlot(p, me, q) + u * kUnitWords + k, tag, &in[k], &verdict);
      }
      if (!ok) break;
      if (q == 0) {
        for (unsigned int k = 0; k < kUnitWords; ++k) acc[k] = in[k];
      } else {
        constexpr int kLanes = (kUnitWords * 4) / sizeof(T);
        T *a = reinterpret_cast<T *>(acc);
        const T *b = reinterpret_cast<const T *>(in);
#pragma unroll
        for (int l = 0; l < kLanes; ++l) a[l] = OP::Apply(a[l], b[l]);
      }
    }
    if (!ok) atomicMax(&s_verdict, verdict);
  }
  __syncthreads();
  verdict = s_verdict;
  __syncthreads();  // s_verdict is about to be reused by the gate
  verdict = CommitGate(p, me, verdict, &s_verdict);  // fault-tolerant path: all blocks or none
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  if (mine) {
    for (unsigned int k = 0; k < kUnitWords; ++k) {
      user[u * kUnitWords + k] = acc[k];
      if (keep != nullptr) keep[u * kUnitWords + k] = acc[k];
    }
  }
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
template <typename T, typename OP>
__global__ void __launch_bounds__(512) OneShotLL(Params p) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  OneShotLLBody<T, OP>(p);
  EndOp(p, me);
}

}  // namespace device
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_COLL_LL_CUH_
