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
  unsigned int base = 0;
  __device__ __forceinline__ unsigned long long *operator()(Control *c) const {
    return &c->exit_[blockIdx.x + base][0];
  }
};

/*! \brief rank r's copy of the symmetric buffer */
__device__ __forceinline__ char *SymOf(const Params &p, int r) { return p.arena[r] + p.sym_offset; }

template <typename T, typename OP>
__device__ __forceinline__ void TwoShotInPlaceBody(const Params &p) {
  __shared__ int s_verdict;
  const int me = p.rank;
  char *mine = SymOf(p, me);
  char *keep = static_cast<char *>(p.keep[0]);
  Stamp(p, 0);
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[0], true, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 2);
  // symmetric buffers are allocated in 256-byte units, so the padded last vector is addressable
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const int n = p.world;
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    if (TileOwner(k, n) != me) continue;
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
  verdict = FlagBarrier(p, me, MidRow{}, OpEpoch(p) + 1, 0u, false, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 4);
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    const int owner = TileOwner(k, n);
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
  verdict = FlagBarrier(p, me, ExitRow{}, OpEpoch(p) + 1, 0u, false, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 5);
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
template <typename T, typename OP>
__global__ void __launch_bounds__(512) TwoShotInPlace(Params p) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  if (!Poisoned(p, me)) TwoShotInPlaceBody<T, OP>(p);
  EndOp(p, me);
}

/*! \brief switch-side reductions have a longer round trip than peer loads: keep 8 vectors per thread in flight */
constexpr int kNvlsUnroll = 8;

template <typename MM>
__device__ __forceinline__ void NvlsInPlaceBody(const Params &p) {
  __shared__ int s_verdict;
  const int me = p.rank;
  char *keep = static_cast<char *>(p.keep[0]);
  char *mc = p.mc_arena + p.sym_offset;
  Stamp(p, 0);
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[0], true, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  Stamp(p, 2);
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kNvlsUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const int n = p.world;
  for (size_t tile = blockIdx.x, k = 0; tile < ntiles; tile += gridDim.x, ++k) {
    if (TileOwner(k, n) != me) continue;
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
  verdict = FlagBarrier(p, me, MidRow{}, OpEpoch(p) + 1, 0u, false, &s_verdict);
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
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
template <typename MM>
__global__ void __launch_bounds__(512) NvlsInPlace(Params p) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  if (!Poisoned(p, me)) NvlsInPlaceBody<MM>(p);
  EndOp(p, me);
}

/*!
 * \brief zero-copy broadcast: the root stores its symmetric buffer through the multicast mapping, the
 *  switch writes it into every rank's copy of the buffer; one barrier tells everybody it is complete.
 */
__device__ __forceinline__ void NvlsBroadcastInPlaceBody(const Params &p) {
  __shared__ int s_verdict;
  const int me = p.rank;
  char *keep = static_cast<char *>(p.keep[0]);
  const size_t nvec = (p.bytes + kVec - 1) / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kNvlsUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  // nobody may still be using the previous contents of its buffer when the root starts overwriting it
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[0], true, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  if (me == p.root) {
    const char *src = SymOf(p, me);
    char *mc = p.mc_arena + p.sym_offset;
    for (size_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const size_t v0 = tile * tile_vecs + threadIdx.x;
      uint4 r[kNvlsUnroll];
#pragma unroll
      for (int u = 0; u < kNvlsUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) r[u] = LoadVec(src + v * kVec);
      }
#pragma unroll
      for (int u = 0; u < kNvlsUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) {
          MultimemStore(mc + v * kVec, r[u]);
          if (keep != nullptr) StoreVec(keep + v * kVec, r[u]);
        }
      }
    }
  }
  verdict = FlagBarrier(p, me, MidRow{}, OpEpoch(p) + 1, 0u, false, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  if (me != p.root && keep != nullptr) {
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
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
__global__ void __launch_bounds__(512) NvlsBroadcastInPlace(Params p) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  NvlsBroadcastInPlaceBody(p);
  EndOp(p, me);
}

/*!
 * \brief staged broadcast through the switch, chunk-pipelined: the root streams its (ordinary) user
 *  buffer into everybody's stage with multimem.st and publishes a per-block round counter every
 *  kBcastRoundTiles tiles; the other ranks copy finished rounds out of their stage while later rounds
 *  are still in flight. Needs 16-byte aligned buffers and a size that is a multiple of 16.
 */
constexpr int kBcastRoundTiles = 4;

__device__ __forceinline__ void NvlsBroadcastPipeBody(const Params &p, unsigned long long round_base) {
  __shared__ int s_verdict;
  __shared__ int s_abort;
  const int me = p.rank;
  char *user = static_cast<char *>(p.user[0]);
  char *keep = static_cast<char *>(p.keep[0]);
  const int b = blockIdx.x, G = gridDim.x, n = p.world;
  const size_t nvec = p.bytes / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const size_t chan_tiles = ntiles > static_cast<size_t>(b) ? (ntiles - b + G - 1) / G : 0;
  const size_t rounds = (chan_tiles + kBcastRoundTiles - 1) / kBcastRoundTiles;
  if (threadIdx.x == 0) s_abort = 0;
  // stage[parity] of every rank is free once everybody has entered this op (see coll_common.cuh)
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[0], true, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  if (me == p.root) {
    char *mc = McStage(p);
    for (size_t j = 0; j < rounds; ++j) {
      for (int i = 0; i < kBcastRoundTiles; ++i) {
        const size_t k = j * kBcastRoundTiles + i;
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
          if (v < nvec) {
            MultimemStore(mc + v * kVec, r[u]);
            if (keep != nullptr) StoreVec(keep + v * kVec, r[u]);
          }
        }
      }
      __syncthreads();
      if (threadIdx.x < n) StoreReleaseSys(&CtrlOf(p, threadIdx.x)->pipe_a[b][me], round_base + j + 1);
    }
  } else {
    const char *src = StageOf(p, me);
    const unsigned long long *slot = &CtrlOf(p, me)->pipe_a[b][p.root];
    for (size_t j = 0; j < rounds; ++j) {
      if (threadIdx.x == 0) {
        const unsigned long long deadline = GlobalTimer() + p.timeout_ns;
        unsigned int spins = 0;
        while (LoadAcquireSys(slot) < round_base + j + 1) {
          if ((++spins & 0x3ff) == 0) {
            if (*p.abort_flag != 0) { s_abort = kStAborted; break; }
            if (GlobalTimer() > deadline) { s_abort = kStTimeout; break; }
          }
        }
      }
      __syncthreads();
      if (s_abort != 0) break;
      for (int i = 0; i < kBcastRoundTiles; ++i) {
        const size_t k = j * kBcastRoundTiles + i;
        if (k >= chan_tiles) break;
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
    }
    if (s_abort != 0) Report(p, me, s_abort);
  }
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
__global__ void __launch_bounds__(512) NvlsBroadcastPipe(Params p, unsigned long long round_base) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  NvlsBroadcastPipeBody(p, round_base);
  EndOp(p, me);
}

/*!
 * \brief broadcast as scatter + multicast allgather. A single GPU streaming the whole buffer through
 *  multimem.st reaches only ~390-540 GB/s (measured, 8 and 2 B200); eight GPUs multicasting an eighth
 *  each saturate the fabric. So the root P2P-stores tile k of every channel into the buffer of rank
 *  k % world (its own share goes straight through the multicast mapping), and every rank re-emits the
 *  tiles it received with multimem.st. Rounds of `world` tiles are pipelined with per-block counters:
 *  A[root] "round scattered", B[i] "rank i has multicast its tile of the round".
 *  inplace != 0: source and destination are the symmetric user buffer (no copy-out at all);
 *  otherwise the data lands in the stage and each rank copies finished rounds out to its user buffer.
 *  direct != 0 (used for 2 ranks, where re-emitting would send every byte across twice): the root
 *  simply P2P-stores every tile into every peer and the peers only wait for A.
 */
__device__ __forceinline__ void BroadcastScatterMcBody(const Params &p, unsigned long long round_base, int inplace,
                                                          int direct) {
  __shared__ int s_verdict;
  __shared__ int s_abort;
  const int me = p.rank, n = p.world, root = p.root;
  const int b = blockIdx.x, G = gridDim.x;
  char *user = static_cast<char *>(p.user[0]);
  char *keep = static_cast<char *>(p.keep[0]);
  const size_t nvec = p.bytes / kVec;
  const size_t tile_vecs = static_cast<size_t>(blockDim.x) * kUnroll;
  const size_t ntiles = (nvec + tile_vecs - 1) / tile_vecs;
  const size_t chan_tiles = ntiles > static_cast<size_t>(b) ? (ntiles - b + G - 1) / G : 0;
  const size_t rounds = (chan_tiles + n - 1) / n;
  if (threadIdx.x == 0) s_abort = 0;
  int verdict = FlagBarrier(p, me, EntryRow{OpEpoch(p) & 1u}, OpEpoch(p) + 1, p.code[0], true, &s_verdict);
  if (verdict != kStOk) { Report(p, me, verdict); return; }
  char *my_dst = inplace ? SymOf(p, me) : StageOf(p, me);
  char *mc_dst = inplace ? p.mc_arena + p.sym_offset : McStage(p);
  Control *my_ctrl = CtrlOf(p, me);
  // one thread waits for a counter, the block follows; returns false when the op must be abandoned
  auto wait_for = [&](const unsigned long long *slot, unsigned long long target) -> bool {
    if (threadIdx.x == 0) {
      const unsigned long long deadline = GlobalTimer() + p.timeout_ns;
      unsigned int spins = 0;
      while (LoadAcquireSys(slot) < target) {
        if ((++spins & 0x3ff) == 0) {
          if (*p.abort_flag != 0) { s_abort = kStAborted; break; }
          if (GlobalTimer() > deadline) { s_abort = kStTimeout; break; }
        }
      }
    }
    __syncthreads();
    return s_abort == 0;
  };
  auto copy_out = [&](size_t j) {  // stage -> user (+keep) for every tile of round j
    for (int i = 0; i < n; ++i) {
      const size_t k = j * n + i;
      if (k >= chan_tiles) break;
      const size_t v0 = (b + k * G) * tile_vecs + threadIdx.x;
      uint4 r[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) r[u] = LoadVec(my_dst + v * kVec);
      }
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
        if (v < nvec) {
          if (!inplace) StoreVec(user + v * kVec, r[u]);
          if (keep != nullptr) StoreVec(keep + v * kVec, r[u]);
        }
      }
    }
  };
  if (me == root) {
    const char *src = inplace ? my_dst : user;
    for (size_t j = 0; j < rounds; ++j) {
      for (int i = 0; i < n; ++i) {
        const size_t k = j * n + i;
        if (k >= chan_tiles) break;
        const size_t v0 = (b + k * G) * tile_vecs + threadIdx.x;
        char *peer = inplace ? SymOf(p, i) : StageOf(p, i);
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
            if (direct) {
              for (int q = 0; q < n; ++q) {
                if (q != root) StoreVec((inplace ? SymOf(p, q) : StageOf(p, q)) + v * kVec, r[u]);
              }
            } else if (i == root) {
              MultimemStore(mc_dst + v * kVec, r[u]);
            } else {
              StoreVec(peer + v * kVec, r[u]);
            }
            if (keep != nullptr) StoreVec(keep + v * kVec, r[u]);
          }
        }
      }
      __syncthreads();
      if (threadIdx.x < n) StoreReleaseSys(&CtrlOf(p, threadIdx.x)->pipe_a[b][me], round_base + j + 1);
    }
    if (inplace && !direct) {
      // peers still multicast old tiles into OUR buffer: do not return (and let the caller refill it) before they are done
      for (int i = 0; i < n; ++i) {
        if (i != root && !wait_for(&my_ctrl->pipe_b[b][i], round_base + rounds)) break;
      }
    }
  } else {
    for (size_t j = 0; j <= rounds; ++j) {
      if (j < rounds) {
        if (!wait_for(&my_ctrl->pipe_a[b][root], round_base + j + 1)) break;
        if (direct) {
          if (!inplace || keep != nullptr) copy_out(j);
          continue;
        }
        const size_t k = j * n + me;
        if (k < chan_tiles) {
          const size_t v0 = (b + k * G) * tile_vecs + threadIdx.x;
          uint4 r[kUnroll];
#pragma unroll
          for (int u = 0; u < kUnroll; ++u) {
            const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
            if (v < nvec) r[u] = LoadVec(my_dst + v * kVec);
          }
#pragma unroll
          for (int u = 0; u < kUnroll; ++u) {
            const size_t v = v0 + static_cast<size_t>(u) * blockDim.x;
            if (v < nvec) MultimemStore(mc_dst + v * kVec, r[u]);
          }
        }
        __syncthreads();
        if (threadIdx.x < n) StoreReleaseSys(&CtrlOf(p, threadIdx.x)->pipe_b[b][me], round_base + j + 1);
      }
      if (direct) continue;
      if (j > 0) {
        // round j-1: every owner except the root announces its tile through B; the root's own tile and the
        // scatter itself are covered by A (already awaited above for round j-1)
        bool ok = true;
        for (int i = 0; i < n && ok; ++i) {
          if (i != root && i != me) ok = wait_for(&my_ctrl->pipe_b[b][i], round_base + j);
        }
        if (!ok) break;
        if (!inplace || keep != nullptr) copy_out(j - 1);
      }
    }
  }
  if (s_abort != 0) Report(p, me, s_abort);
}
/*! \brief launch wrapper: epoch from the control block (BeginOp), body, completion count (EndOp) */
__global__ void __launch_bounds__(512) BroadcastScatterMc(Params p, unsigned long long round_base, int inplace,
                                                          int direct) {
  const int me = p.emulate ? blockIdx.y : p.rank;
  BeginOp(p, me);
  BroadcastScatterMcBody(p, round_base, inplace, direct);
  EndOp(p, me);
}

/*! \brief rounds BroadcastScatterMc consumes */
inline unsigned long long ScatterRounds(size_t bytes, int grid, int world) {
  const size_t tile = 512 * kUnroll * kVec;
  const size_t ntiles = (bytes + tile - 1) / tile;
  const size_t per_chan = (ntiles + grid - 1) / grid;
  return (per_chan + world - 1) / world;
}

/*! \brief rounds NvlsBroadcastPipe consumes (same on every rank) */
inline unsigned long long BcastRounds(size_t bytes, int grid) {
  const size_t tile = 512 * kUnroll * kVec;
  const size_t ntiles = (bytes + tile - 1) / tile;
  const size_t per_chan = (ntiles + grid - 1) / grid;
  return (per_chan + kBcastRoundTiles - 1) / kBcastRoundTiles;
}

}  // namespace device
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_COLL_INPLACE_CUH_
