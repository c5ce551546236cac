/*!
 * rabit-b200: C ABI of the device (sm_100a) side. Everything here is exported by lib/librabit*.so
 * when the library was built with nvcc (python build.py); none of it exists in the reference,
 * which has no GPU path at all.
 *
 * The fault-tolerant collectives need no extra API: RabitAllreduce / RabitBroadcast /
 * RabitAllgather (rabit/c_api.h) recognise CUDA device pointers and run the fused kernels.
 * This header adds (1) stream-ordered, non-fault-tolerant variants for training loops,
 * (2) symmetric (zero-copy) buffers, (3) tuning hooks and (4) the single-GPU rank emulation
 * used by tests, smoke() and the 1-GPU benchmark.
 *
 * dtype / op integers are the rabit::engine::mpi enums (see rabit/c_api.h).
 */
#ifndef RABIT_DEVICE_H_
#define RABIT_DEVICE_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- state ------------------------------------------------------------------------------ */
/* 1 when this thread's engine has an attached device backend */
int RabitDeviceActive(void);
/* 1 when the NVLS multicast mapping is up (multimem kernels selectable) */
int RabitDeviceHasMulticast(void);
/* human readable one-liner (arena kind and size, NVLS on/off ...); returns the full length */
int RabitDeviceDescribe(char *out, size_t max_len);
/* number of kernels of this library launched so far by this process */
unsigned long long RabitDeviceLaunchCount(void);

/* ---- stream-ordered collectives (enqueue only; no consensus, no result cache) ---------- */
/* CUDA stream used for all device work of this thread's engine (default: the legacy stream 0) */
void RabitDeviceSetStream(void *cuda_stream);
/* 0 on success, negative when there is no backend or the arguments are unusable */
int RabitDeviceAllreduceAsync(void *dptr, size_t count, int dtype, int op);
int RabitDeviceBroadcastAsync(void *dptr, size_t bytes, int root);
int RabitDeviceAllgatherAsync(void *dptr, size_t total_bytes, size_t slice_begin, size_t slice_end);
/* wait for everything enqueued: 0 ok, 1 consensus mismatch, 2 peer timeout, 3 peer lost */
int RabitDeviceSynchronize(void);

/* ---- symmetric buffers (zero-copy collectives) ------------------------------------------ */
/* Carve nbytes out of the symmetric heap (sized with rabit_device_heap=<bytes> at init). Collective:
 * same order and sizes on every rank. NULL when exhausted / no backend. Freed as a whole. */
void *RabitDeviceSymAlloc(size_t nbytes);
void RabitDeviceSymReset(void);

/* ---- tuning ------------------------------------------------------------------------------- */
/* algo: "auto", "ll", "oneshot", "twoshot", "tma", "nvls", "staged", "nostream" (NULL keeps it);
 * blocks: grid cap, 0 = default; trace != 0: block 0 records %globaltimer at its phase boundaries */
void RabitDeviceSetTuning(const char *algo, int blocks, int trace);
void RabitDeviceGetTrace(unsigned long long *out6);

/* ---- single-GPU emulation of an nranks-wide collective ----------------------------------- */
/* kind 0 allreduce (size = elements), 1 broadcast (size = bytes), 2 allgather (size = total bytes);
 * algo 1 one-shot, 2 two-shot, 3 two-shot/TMA, 9 tagged words (<= 32 KiB); bufs[r]/keeps[r] device pointers per
 * emulated rank; codes[r] consensus words (NULL = 0); status_out[r]: 0 ok, 1 mismatch, 2 timeout.
 * Allocates its arenas with cudaMalloc on first use. Returns 0 or a negative API error. */
int RabitEmuCollective(int kind, int algo, int nranks, void **bufs, void **keeps, size_t size, int dtype, int op,
                       int root, const size_t *slice_begin, const size_t *slice_end, const unsigned *codes,
                       int blocks, int timeout_ms, int *status_out);
void RabitEmuReset(void);
/* The chunked (streamed) allreduce on emulated ranks: chunks of chunk_bytes are copied into each rank's stage and
 * reduced by one StreamP2P launch each; fail_chunk >= 0 injects a consensus mismatch there, after which the call
 * restores bufs[r] from the stage copy like the multi-GPU path does. status_out[r]: first failure of rank r. */
int RabitEmuStreamAllreduce(int nranks, void **bufs, void **keeps, size_t count, int dtype, int op, size_t chunk_bytes,
                            int fail_chunk, int blocks, int *status_out);
/* Test hook: set the op counter of the emulated ranks (flag rows made consistent with it). */
int RabitEmuSetEpoch(int nranks, size_t payload_bytes, unsigned int epoch);

/* ---- allocation-free variant -------------------------------------------------------------- */
/* No counterpart in the reference. The emulated two-shot allreduce for callers that must not trigger
 * a GPU allocation: processes whose device memory is owned in full by another allocator (a
 * framework's caching pool, a fixed tenant budget) or that forbid driver allocations after start-up.
 * All scratch (flags, staging, status) is a static __device__ array inside the library: about 6 MiB
 * that every GPU loading the library holds until exit. Limits: at most 8 ranks and 256 KiB of
 * payload per rank; anything larger returns -3 with a message on stderr. The call clears the part
 * of the scratch that must start at zero, on `cuda_stream`, and then launches there; all calls share
 * the scratch, so calls that overlap on the GPU race. The first call on a device looks up the
 * array's address and must be made outside stream capture. bufs[r] are caller-owned device
 * buffers reduced in place. With status_out != NULL the call waits for the stream and reports one
 * status per rank (0 ok, 1 mismatch, 2 timeout). Returns 0, or -1 bad arguments, -2 unsupported
 * dtype/op, -3 too large, -4 CUDA error. */
int RabitEmuAllreduceNoAlloc(int nranks, void **bufs, size_t count, int dtype, int op, void *cuda_stream,
                             int *status_out);

#ifdef __cplusplus
}
#endif
#endif  /* RABIT_DEVICE_H_ */
