/*!
 * rabit-b200: device-side building blocks of the peer-memory collectives (sm_100a).
 *
 * No counterpart in the reference (CPU/TCP only). What its AllreduceBase does with a
 * binary tree and a ring of sockets (src/allreduce_base.cc:476-977) is done here by
 * kernels that load/store the other GPUs' memory directly over NVLink 5 / NVSwitch:
 *
 *   symmetric arena   every rank maps every rank's arena; same layout everywhere:
 *                     [control block | stage 0 | stage 1]
 *   channels          block b of every rank works on the same tile set and only ever
 *                     synchronises with block b of the peers (per-block flag rows), so
 *                     there is no grid-wide barrier and blocks drift freely
 *   flag barrier      monotonically increasing 64-bit words written with st.release.sys
 *                     into the PEER's control block, polled locally with ld.acquire.sys
 *   fused consensus   the low 32 bits of the entry-barrier flag carry the rank's
 *                     protocol code (proto::Request::Code); a block that sees a peer with
 *                     a different code leaves without touching data -> Status::kMismatch
 *   bounded spin      %globaltimer deadline + host-mapped abort word: a dead peer turns
 *                     into an error code, never into a hung GPU
 */
#ifndef RABIT_SRC_DEVICE_COLL_COMMON_CUH_
#define RABIT_SRC_DEVICE_COLL_COMMON_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

namespace rabit {
namespace device {

constexpr int kMaxRanks = 16;
constexpr int kMaxBlocks = 296;          // 2 resident CTAs on each of 148 SMs
constexpr int kCtrlBytes = 256 << 10;    // control block at the start of every arena
constexpr int kVec = 16;                 // bytes moved per thread per access

// kernel exit codes written to the host-mapped status word (0 = fine)
constexpr int kStOk = 0;
constexpr int kStMismatch = 1;
constexpr int kStTimeout = 2;
constexpr int kStAborted = 3;

/*! \brief layout of the control block (per arena). Row = [block][source rank]. */
struct Control {
  unsigned long long entry[2][kMaxBlocks][kMaxRanks];  // entry barrier, double-buffered by epoch parity
  unsigned long long mid[kMaxBlocks][kMaxRanks];       // second barrier of two-phase kernels
  unsigned long long exit_[kMaxBlocks][kMaxRanks];     // exit barrier of the in-place kernels
  unsigned long long pipe_a[kMaxBlocks][kMaxRanks];    // pipelined kernel: rounds staged by rank r
  unsigned long long pipe_b[kMaxBlocks][kMaxRanks];    // pipelined kernel: rounds reduced by rank r
  // op counter kept ON THE DEVICE so that a captured CUDA graph can be replayed: every kernel reads it at its
  // start and the last block to finish bumps it (kept at the end: the offsets of the rows above never move)
  unsigned int epoch_dev;
  unsigned int done_count;
  // commit gate of the fault-tolerant path: every block posts (epoch + 1) << 32 | verdict in its own slot,
  // block 0 folds them and posts the grid's verdict in gate_go (no contended atomics)
  unsigned long long gate_slot[kMaxBlocks];
  unsigned long long gate_go;
  // multi-kernel ops: a kernel that fails leaves (1 << 32 | epoch of the op's FIRST kernel) here; the op's later
  // kernels look at it before they wait for anybody (device memory on purpose: a read of host memory in front of
  // the entry barrier makes its system-scope fence cost ~1 us per block, serialised over the grid - measured)
  unsigned long long op_failed;
  // peer-copy transport (coll_blob.cuh): piece counters written by the other end of a transfer.
  // deliver: one holder -> requesters; shift[d]: ring neighbours, d = 0 forward (from the predecessor), 1 backward
  unsigned long long blob_ready[kMaxRanks];   // [holder]: "piece k is in my window"
  unsigned long long blob_done[kMaxRanks];    // [requester]: "I have taken piece k"
  unsigned long long shift_ready[2];
  unsigned long long shift_done[2];
};
static_assert(sizeof(Control) <= kCtrlBytes, "control block too small");

/*! \brief everything a collective kernel needs; passed by value */
struct Params {
  char *arena[kMaxRanks];     // arena base of every rank as mapped HERE
  char *mc_arena;             // multicast alias of the arenas (NVLS), or NULL
  int world;
  int rank;                   // ignored in emulation (rank = blockIdx.y)
  int emulate;                // 1: all "ranks" are slices of this grid on one GPU (tests, smoke)
  unsigned int epoch;         // op counter, same on all ranks
  unsigned int code[kMaxRanks];   // consensus code (index 0 unless emulating)
  void *user[kMaxRanks];      // user buffer (index 0 unless emulating)
  void *keep[kMaxRanks];      // optional second destination = result-cache entry
  size_t bytes;               // payload bytes
  size_t stage_bytes;         // capacity of one stage (>= bytes rounded to 16)
  size_t sym_offset;          // in-place kernels: offset of the symmetric user buffer inside every arena
  int root;                   // broadcast
  size_t slice_begin[kMaxRanks], slice_end[kMaxRanks];  // allgather (index 0 unless emulating)
  volatile int *abort_flag;   // host-mapped; non-zero = give up
  int *status;                // host-mapped; first non-zero code wins (per emulated rank: status[r])
  unsigned long long timeout_ns;
  unsigned long long *trace;  // optional host-mapped phase timestamps of block 0 (NULL = off)
  int device_epoch;           // 1: take the op epoch from Control::epoch_dev instead of `epoch` (graph-replayable)
  int chained;                // index of this kernel inside a multi-kernel op (0 = first, or a single-kernel op):
                              // a later kernel leaves at once if an earlier one of the same op failed
  unsigned int block_base;    // first flag row of this kernel (chunk kernels of one op that run concurrently)
};

/*! \brief the op epoch of the running kernel. With device_epoch it was fetched from the control block by BeginOp
 *  into ONE shared word per block (Params stays untouched in the constant bank: writing into it would make the
 *  compiler spill the whole 800-byte struct to local memory in every thread). */
__device__ __forceinline__ unsigned int &SharedEpoch() {
  __shared__ unsigned int epoch_of_this_launch;
  return epoch_of_this_launch;
}
__device__ __forceinline__ unsigned int OpEpoch(const Params &p) {
  return p.device_epoch ? SharedEpoch() : p.epoch;
}

__device__ __forceinline__ Control *CtrlOf(const Params &p, int r) {
  return reinterpret_cast<Control *>(p.arena[r]);
}
__device__ __forceinline__ char *StageOf(const Params &p, int r) {
  return p.arena[r] + kCtrlBytes + static_cast<size_t>(OpEpoch(p) & 1u) * p.stage_bytes;
}
__device__ __forceinline__ char *McStage(const Params &p) {
  return p.mc_arena + kCtrlBytes + static_cast<size_t>(OpEpoch(p) & 1u) * p.stage_bytes;
}

__device__ __forceinline__ unsigned long long GlobalTimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void StoreReleaseSys(unsigned long long *addr, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ void StoreRelaxedSys(unsigned long long *addr, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(addr), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long LoadAcquireSys(const unsigned long long *addr) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(addr) : "memory");
  return v;
}

/*! \brief 16-byte streaming load that does not pollute L1 */
__device__ __forceinline__ uint4 LoadVec(const void *ptr) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(ptr));
  return v;
}
__device__ __forceinline__ void StoreVec(void *ptr, const uint4 &v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
               ::"l"(ptr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

/*!
 * \brief block-level flag barrier with the peers' same-numbered block.
 *  Every thread must call it. Thread i < world writes `value` into rank i's row (slot = my
 *  rank) and then waits until rank i has written >= value (compared on the high 32 bits)
 *  into my row. Returns kStOk, or an error every thread of the block agrees on.
 * \param row_of  functor: Control* -> pointer to this block's row of kMaxRanks words
 * \param low32   payload in the low half (consensus code on the entry barrier, else 0)
 * \param peer_low receives rank i's low half in thread i (only when want_low)
 */
template <typename RowOf>
__device__ __forceinline__ int FlagBarrier(const Params &p, int me, RowOf row_of, unsigned int seq,
                                           unsigned int low32, bool check_code, int *s_verdict,
                                           bool publishes = true) {
  __syncthreads();  // all of this block's prior global writes are ordered before the signal
  const int t = threadIdx.x;
  if (t == 0) *s_verdict = kStOk;
  __syncthreads();
  if (t < p.world) {
    const unsigned long long mine = (static_cast<unsigned long long>(seq) << 32) | low32;
    // publishes == false: this kernel has written nothing the peers will read before the barrier (its input was
    // put in place by a copy engine or an earlier kernel, complete at the kernel boundary), so the signal needs
    // no release fence. A system-scope fence queues behind PCIe traffic: ~45 us under H2D/D2H load (measured).
    if (publishes) StoreReleaseSys(row_of(CtrlOf(p, t)) + me, mine);
    else StoreRelaxedSys(row_of(CtrlOf(p, t)) + me, mine);
    const unsigned long long *slot = row_of(CtrlOf(p, me)) + t;
    const unsigned long long deadline = GlobalTimer() + p.timeout_ns;
    unsigned long long seen;
    int verdict = kStOk;
    unsigned int spins = 0;
    while (true) {
      seen = LoadAcquireSys(slot);
      // wrap-safe "seen >= seq": the 32-bit op counter may roll over in a long graph-replay loop
      if (static_cast<int>(static_cast<unsigned int>(seen >> 32) - seq) >= 0) break;
      if ((++spins & 0x3ff) == 0) {
        if (*p.abort_flag != 0) { verdict = kStAborted; break; }
        if (GlobalTimer() > deadline) { verdict = kStTimeout; break; }
      }
    }
    if (verdict == kStOk && check_code && static_cast<unsigned int>(seen >> 32) == seq &&
        static_cast<unsigned int>(seen) != low32) {
      verdict = kStMismatch;
    }
    if (verdict != kStOk) atomicMax(s_verdict, verdict);
  }
  __syncthreads();
  return *s_verdict;
}

/*! \brief what a failed kernel of the op that began at epoch `first_epoch` leaves in Control::op_failed */
__device__ __forceinline__ unsigned long long FailTag(unsigned int first_epoch) {
  return (1ull << 32) | first_epoch;
}
/*! \brief record the first failure of this (emulated) rank for the host, and for the op's later kernels */
__device__ __forceinline__ void Report(const Params &p, int me, int verdict) {
  if (threadIdx.x == 0 && verdict != kStOk) {
    atomicCAS(p.status + (p.emulate ? me : 0), kStOk, verdict);
    *reinterpret_cast<volatile unsigned long long *>(&CtrlOf(p, me)->op_failed) =
        FailTag(OpEpoch(p) - static_cast<unsigned int>(p.chained));
  }
}

/*!
 * \brief all-or-nothing commit point of ONE rank's grid, used on the fault-tolerant path only (a result-cache
 *  copy was requested). The flag barriers are per block channel, so when a peer dies in the middle of a kernel
 *  some blocks of a survivor may pass their last barrier while others time out. An allreduce must then not
 *  touch the caller's buffer at all: the engine retries the op from that buffer (the reference reduces inside
 *  its cache slot for the same reason, allreduce_robust.cc:279-283). Every block arrives here with its own
 *  verdict after the last barrier that guards the first write into the user buffer; blocks that passed wait
 *  for the whole grid and proceed only if nobody failed. Failed blocks arrive and leave at once.
 *  Local memory traffic only: every block posts its verdict in its own slot, block 0 folds them and posts the
 *  grid's verdict (two L2 round trips, no contended atomics). Needs the grid co-resident, which the launch
 *  configuration guarantees (<= one block per SM).
 */
__device__ __forceinline__ int CommitGate(const Params &p, int me, int verdict, int *s_verdict) {
  if (p.keep[p.emulate ? me : 0] == nullptr || gridDim.x == 1) return verdict;
  __syncthreads();
  Control *c = CtrlOf(p, me);
  // slot word = op epoch | "posted" marker | verdict: zeroed memory never looks posted, whatever the epoch
  const unsigned long long kMark = 0x80000000ull;
  const unsigned long long tag = (static_cast<unsigned long long>(OpEpoch(p)) << 32) | kMark;
  const int t = threadIdx.x;
  if (t == 0) {
    *s_verdict = verdict;
    asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(&c->gate_slot[blockIdx.x]),
                 "l"(tag | static_cast<unsigned long long>(verdict)) : "memory");
  }
  if (blockIdx.x != 0 && verdict != kStOk) return verdict;  // posted; nobody waits for a block that failed
  __syncthreads();
  const unsigned long long deadline = GlobalTimer() + p.timeout_ns;
  unsigned int spins = 0;
  if (blockIdx.x == 0) {
    // coordinator: thread i watches block i (its own failure is published without waiting for anybody)
    if (verdict == kStOk) {
      for (unsigned int b = t; b < gridDim.x; b += blockDim.x) {
        int seen = kStOk;
        while (true) {
          unsigned long long v;
          asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(&c->gate_slot[b]) : "memory");
          if ((v & ~(kMark - 1)) == tag) { if ((v & (kMark - 1)) != 0u) seen = kStAborted; break; }
          if ((++spins & 0x3ff) == 0) {
            if (*p.abort_flag != 0) { seen = kStAborted; break; }
            if (GlobalTimer() > deadline) { seen = kStTimeout; break; }
          }
        }
        if (seen != kStOk) atomicMax(s_verdict, seen);
      }
    }
    __syncthreads();
    if (t == 0) {
      asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(&c->gate_go),
                   "l"(tag | static_cast<unsigned long long>(*s_verdict)) : "memory");
    }
  } else if (t == 0) {
    int seen = kStOk;
    while (true) {
      unsigned long long v;
      asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(&c->gate_go) : "memory");
      if ((v & ~(kMark - 1)) == tag) { if ((v & (kMark - 1)) != 0u) seen = kStAborted; break; }
      if ((++spins & 0x3ff) == 0) {
        if (*p.abort_flag != 0) { seen = kStAborted; break; }
        if (GlobalTimer() > deadline) { seen = kStTimeout; break; }
      }
    }
    *s_verdict = seen;
  }
  __syncthreads();
  return *s_verdict;
}

/*! \brief first thing every kernel does: fetch the op epoch. All blocks read the same value because it is only
 *  advanced by EndOp of the LAST block of the launch. */
__device__ __forceinline__ void BeginOp(const Params &p, int me) {
  if (p.device_epoch) {
    if (threadIdx.x == 0) SharedEpoch() = *reinterpret_cast<volatile unsigned int *>(&CtrlOf(p, me)->epoch_dev);
    __syncthreads();
  }
}
/*! \brief last thing every kernel does (also on the error exits): count this block; the last one advances the epoch */
__device__ __forceinline__ void EndOp(const Params &p, int me) {
  if (!p.device_epoch) return;
  __syncthreads();
  if (threadIdx.x == 0) {
    Control *c = CtrlOf(p, me);
    const unsigned int blocks = gridDim.x;   // in emulation every virtual rank counts its own gridDim.x blocks
    if (atomicAdd(&c->done_count, 1u) == blocks - 1) {
      c->done_count = 0;
      __threadfence();
      *reinterpret_cast<volatile unsigned int *>(&c->epoch_dev) = OpEpoch(p) + 1;
    }
  }
}

/*! \brief true when an earlier kernel of the same multi-kernel op already reported a failure (block-uniform).
 *  After a failed chunk the peers may have left for the host protocol (consensus mismatch), so the remaining
 *  kernels of the op must not wait for them. */
__device__ __forceinline__ bool Poisoned(const Params &p, int me) {
  __shared__ int s_poison;
  if (p.chained == 0) return false;
  if (threadIdx.x == 0) {
    const unsigned long long seen = *reinterpret_cast<volatile unsigned long long *>(&CtrlOf(p, me)->op_failed);
    s_poison = seen == FailTag(OpEpoch(p) - static_cast<unsigned int>(p.chained)) ? 1 : 0;
  }
  __syncthreads();
  return s_poison != 0;
}

/*! \brief phase timestamp (block 0 / thread 0 only) for the tuning tools; no-op unless enabled */
__device__ __forceinline__ void Stamp(const Params &p, int idx) {
  if (p.trace != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) p.trace[idx] = GlobalTimer();
}

/*!
 * \brief which rank reduces the k-th tile of this block's channel. Offset by the channel number: a message of no
 *  more tiles than blocks gives every block exactly one tile (k = 0), and without the offset rank 0 would own,
 *  i.e. reduce and re-distribute, ALL of them while the other GPUs wait (256 KiB - 4 MiB messages).
 */
__device__ __forceinline__ int TileOwner(size_t k, int n) {
  return static_cast<int>((k + blockIdx.x) % static_cast<size_t>(n));
}

// ------------------------------------------------------------------ reduce ops
struct OpMax { template <typename T> __device__ __forceinline__ static T Apply(T a, T b) { return a < b ? b : a; } };
struct OpMin { template <typename T> __device__ __forceinline__ static T Apply(T a, T b) { return b < a ? b : a; } };
struct OpSum { template <typename T> __device__ __forceinline__ static T Apply(T a, T b) { return a + b; } };
struct OpOr {
  template <typename T> __device__ __forceinline__ static T Apply(T a, T b) { return a | b; }
};

/*! \brief acc = OP(acc, v) lane-wise over a 16-byte vector of T */
template <typename T, typename OP>
__device__ __forceinline__ void ReduceVec(uint4 *acc, const uint4 &v) {
  constexpr int kLanes = kVec / sizeof(T);
  T *a = reinterpret_cast<T *>(acc);
  const T *b = reinterpret_cast<const T *>(&v);
#pragma unroll
  for (int i = 0; i < kLanes; ++i) a[i] = OP::Apply(a[i], b[i]);
}

// ---------------------------------------------------------- user <-> stage copies
/*! \brief copy [0, bytes) between a user buffer and a 16-byte aligned stage, tiles of this block only.
 *  Tile t (kTileVecs vectors) belongs to block (t % gridDim.x). The ragged tail (< 16 B) is
 *  handled byte-wise by the block that owns the tile it falls into. */
constexpr int kUnroll = 4;

template <bool kToStage>
__device__ __forceinline__ void CopyTiles(char *user, char *stage, size_t bytes, bool user_aligned) {
  const size_t nvec = bytes / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const size_t v0 = tile * tile_vecs + threadIdx.x;
    if (user_aligned) {
      uint4 r[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) r[u] = LoadVec((kToStage ? user : stage) + v * kVec);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) StoreVec((kToStage ? stage : user) + v * kVec, r[u]);
      }
    } else {
      // unaligned user pointer: byte granularity on the user side
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) {
          for (int k = 0; k < kVec; ++k) {
            if (kToStage) stage[v * kVec + k] = user[v * kVec + k];
            else user[v * kVec + k] = stage[v * kVec + k];
          }
        }
      }
    }
  }
  // ragged tail
  const size_t tail0 = nvec * kVec;
  if (tail0 < bytes) {
    const size_t tail_tile = ntiles == 0 ? 0 : (nvec / tile_vecs);
    if (tail_tile % gridDim.x == blockIdx.x) {
      for (size_t i = tail0 + threadIdx.x; i < bytes; i += blockDim.x) {
        if (kToStage) stage[i] = user[i]; else user[i] = stage[i];
      }
    }
  }
}

}  // namespace device
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_COLL_COMMON_CUH_
