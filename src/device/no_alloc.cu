/*!
 * rabit-b200: an allreduce entry point that never allocates.
 *
 * RabitEmuCollective (device_capi.inc) gets the arenas its kernels synchronise and stage through
 * from cudaMalloc / cudaHostAlloc the first time it runs. Some hosts of this library cannot let a
 * plug-in do that: a process whose GPU memory has been handed over in full to another allocator
 * (a framework's caching pool, a fixed per-tenant budget) sees a late cudaMalloc fail or stall the
 * device, and embedders with a "no driver allocations after start-up" rule simply forbid it.
 * For them this file offers RabitEmuAllreduceNoAlloc: the same two-shot kernel, with every byte
 * of scratch — control flags, the two staging areas of every emulated rank, the status words —
 * living in ONE file-scope __device__ array that is part of the loaded module.
 *
 * Consequences the caller has to know (also in include/rabit_device.h):
 *   - the array is kNoAllocScratchBytes (about 6 MiB) and every GPU that loads the library holds
 *     it until the process exits, used or not;
 *   - payloads are limited to kNoAllocMaxBytes per rank and kNoAllocMaxRanks ranks; larger
 *     requests are refused with an error, never truncated;
 *   - all calls share the array and it keeps whatever the previous call left there. What the
 *     kernel needs zeroed (the flag rows and the status words) is cleared by this function on the
 *     caller's stream before every launch; the staging areas are fully overwritten before they
 *     are read. Two calls that overlap on the GPU (different streams, or different host threads)
 *     race on the array: serialise them;
 *   - the first call on each device looks the array's device address up (cudaGetSymbolAddress,
 *     with that device current). Make that first call outside any stream capture.
 * No counterpart in the reference.
 */
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>

// coll_kernels.cuh also DEFINES two non-template kernels that device_backend.cu already provides
// to the linker; give this translation unit's (unused) copies other names instead of touching
// the shared header.
#define BroadcastPull BroadcastPullNoAllocUnused
#define AllgatherPull AllgatherPullNoAllocUnused
#include "coll_kernels.cuh"
#undef BroadcastPull
#undef AllgatherPull

namespace rabit {
namespace device {
namespace {

constexpr int kNoAllocMaxRanks = 8;
constexpr size_t kNoAllocMaxBytes = 256 << 10;                       // payload per emulated rank
constexpr size_t kNoAllocStage = kNoAllocMaxBytes + 256;             // + room for the padded tail vector
constexpr size_t kNoAllocArena = kCtrlBytes + 2 * kNoAllocStage;     // control block + two stages
constexpr size_t kNoAllocWords = 256;                                // status[16] and the abort word
constexpr size_t kNoAllocScratchBytes = kNoAllocMaxRanks * kNoAllocArena + kNoAllocWords;
static_assert(kNoAllocArena % 256 == 0, "arenas must keep the 256-byte alignment of the array");

/*! \brief ALL scratch of the no-alloc entry point. In host code this name is only a handle for
 *  cudaGetSymbolAddress — it is not a device address and must never be passed to a kernel. */
static __device__ __align__(256) unsigned char g_no_alloc_scratch[kNoAllocScratchBytes];

constexpr int kMaxDevices = 64;
unsigned char *g_scratch_addr[kMaxDevices] = {nullptr};  // device address of the array, per device

typedef void (*Kern)(Params);
#define RABIT_NA_INT(T) {TwoShotAllreduce<T, OpMax>, TwoShotAllreduce<T, OpMin>, TwoShotAllreduce<T, OpSum>, TwoShotAllreduce<T, OpOr>}
#define RABIT_NA_FLT(T) {TwoShotAllreduce<T, OpMax>, TwoShotAllreduce<T, OpMin>, TwoShotAllreduce<T, OpSum>, nullptr}
const Kern kKernels[8][4] = {RABIT_NA_INT(signed char), RABIT_NA_INT(unsigned char), RABIT_NA_INT(int),
                             RABIT_NA_INT(unsigned int), RABIT_NA_INT(long long), RABIT_NA_INT(unsigned long long),
                             RABIT_NA_FLT(float), RABIT_NA_FLT(double)};
#undef RABIT_NA_INT
#undef RABIT_NA_FLT
const int kTypeSlot[10] = {0, 1, 2, 3, 4, 5, 6, 7, 4, 5};
const size_t kTypeBytes[10] = {1, 1, 4, 4, 8, 8, 4, 8, 8, 8};

int Fail(int code, const char *what, cudaError_t e) {
  if (e != cudaSuccess) {
    fprintf(stderr, "[rabit] RabitEmuAllreduceNoAlloc: %s: %s\n", what, cudaGetErrorString(e));
  } else {
    fprintf(stderr, "[rabit] RabitEmuAllreduceNoAlloc: %s\n", what);
  }
  return code;
}

}  // namespace
}  // namespace device
}  // namespace rabit

/*!
 * \brief nranks-wide allreduce of `count` elements, all ranks emulated on the current GPU (one
 *  cooperative grid, the two-shot kernel), using only the module's static scratch array.
 *
 *  Never calls cudaMalloc, cudaFree, cudaHostAlloc or any other allocating routine. The caller
 *  owns bufs[r] (device memory, reduced in place). Work is enqueued on `cuda_stream`; if
 *  status_out is not NULL the function additionally waits for the stream and stores one kernel
 *  status per rank there (0 ok, 1 consensus mismatch, 2 timeout), otherwise it returns right
 *  after the launch.
 *  The first call on a device resolves the scratch array's address and must not happen inside a
 *  stream capture. Calls share the scratch: do not let two of them overlap on the GPU.
 * \return 0 on success; -1 bad arguments; -2 dtype/op not supported; -3 request exceeds
 *  kNoAllocMaxRanks (8) ranks or kNoAllocMaxBytes (256 KiB) per rank; -4 CUDA error. A message
 *  goes to stderr for every non-zero return.
 */
extern "C" __attribute__((visibility("default")))
int RabitEmuAllreduceNoAlloc(int nranks, void **bufs, size_t count, int dtype, int op, void *cuda_stream,
                             int *status_out) {
  using namespace rabit::device;  // NOLINT(*)
  if (nranks < 1 || bufs == nullptr || count == 0) return Fail(-1, "bad arguments", cudaSuccess);
  if (dtype < 0 || dtype >= 10 || op < 0 || op > 3 || kKernels[kTypeSlot[dtype]][op] == nullptr) {
    return Fail(-2, "unsupported dtype/op combination", cudaSuccess);
  }
  const size_t bytes = count * kTypeBytes[dtype];
  if (nranks > kNoAllocMaxRanks || bytes > kNoAllocMaxBytes) {
    fprintf(stderr, "[rabit] RabitEmuAllreduceNoAlloc: %d ranks x %zu bytes exceeds the static scratch "
            "(at most %d ranks x %zu bytes); use RabitEmuCollective for larger requests\n",
            nranks, bytes, kNoAllocMaxRanks, kNoAllocMaxBytes);
    return -3;
  }
  for (int r = 0; r < nranks; ++r) if (bufs[r] == nullptr) return Fail(-1, "bufs[r] is NULL", cudaSuccess);
  cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return Fail(-4, "cudaGetDevice", e);
  if (dev < 0 || dev >= kMaxDevices) return Fail(-1, "device ordinal out of range", cudaSuccess);
  if (g_scratch_addr[dev] == nullptr) {
    // once per device, with that device current: the host-side symbol is only a handle
    unsigned char *addr = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void **>(&addr), g_no_alloc_scratch);
    if (e != cudaSuccess || addr == nullptr) return Fail(-4, "cudaGetSymbolAddress(scratch)", e);
    g_scratch_addr[dev] = addr;
  }
  unsigned char *base = g_scratch_addr[dev];
  unsigned char *words = base + static_cast<size_t>(kNoAllocMaxRanks) * kNoAllocArena;

  // the previous call's flags and status are still there: clear what must start at zero
  for (int r = 0; r < nranks; ++r) {
    e = cudaMemsetAsync(base + static_cast<size_t>(r) * kNoAllocArena, 0, sizeof(Control), stream);
    if (e != cudaSuccess) return Fail(-4, "cudaMemsetAsync(control)", e);
  }
  e = cudaMemsetAsync(words, 0, kNoAllocWords, stream);
  if (e != cudaSuccess) return Fail(-4, "cudaMemsetAsync(status)", e);

  Params p;
  std::memset(&p, 0, sizeof(p));
  for (int r = 0; r < nranks; ++r) {
    p.arena[r] = reinterpret_cast<char *>(base + static_cast<size_t>(r) * kNoAllocArena);
    p.user[r] = bufs[r];
  }
  p.world = nranks;
  p.emulate = 1;
  p.epoch = 0;  // control blocks were just zeroed, so every call starts a fresh epoch sequence
  p.bytes = bytes;
  p.stage_bytes = kNoAllocStage;
  p.status = reinterpret_cast<int *>(words);
  p.abort_flag = reinterpret_cast<volatile int *>(words + 128);
  p.timeout_ns = 2000ull * 1000000ull;

  const void *fn = reinterpret_cast<const void *>(kKernels[kTypeSlot[dtype]][op]);
  int occ = 0, sms = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, 512, 0);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) return Fail(-4, "occupancy query", e);
  const size_t tile = 512 * kUnroll * kVec;
  const int resident = (occ > 0 ? occ : 1) * sms;
  int g = static_cast<int>((bytes + tile - 1) / tile);
  if (g > resident / nranks) g = resident / nranks;
  if (g < 1) return Fail(-4, "GPU too small for a cooperative grid of this many ranks", cudaSuccess);
  void *args[1] = {&p};
  e = cudaLaunchCooperativeKernel(fn, dim3(g, nranks), dim3(512), args, 0, stream);
  if (e != cudaSuccess) return Fail(-4, "cudaLaunchCooperativeKernel", e);
  if (status_out != nullptr) {
    e = cudaMemcpyAsync(status_out, words, nranks * sizeof(int), cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) return Fail(-4, "status read-back", e);
  }
  return 0;
}
