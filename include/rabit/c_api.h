/*!
 * rabit-b200: C ABI. The first 15 symbols are signature-compatible with reference
 * include/rabit/c_api.h:37-194 so python/rabit.py style ctypes callers and other language
 * bindings keep working. The RabitDevice* group is new: explicit CUDA-stream variants and
 * symmetric-memory management for the sm_100a kernels.
 *
 * dtype / op integers are engine::mpi::DataType / OpType values:
 *   dtype: 0 char, 1 uchar, 2 int, 3 uint, 4 long, 5 ulong, 6 float, 7 double,
 *          8 longlong, 9 ulonglong       op: 0 max, 1 min, 2 sum, 3 bitwise-or
 */
#ifndef RABIT_C_API_H_
#define RABIT_C_API_H_

#ifdef __cplusplus
#define RABIT_EXTERN_C extern "C"
#include <cstddef>
#else
#define RABIT_EXTERN_C
#include <stddef.h>
#include <stdbool.h>
#endif

#define RABIT_CAPI RABIT_EXTERN_C __attribute__((visibility("default")))

typedef unsigned long rbt_ulong;  /* NOLINT(*) */

/* ---- life cycle / identity ------------------------------------------------------ */
RABIT_CAPI bool RabitInit(int argc, char *argv[]);
RABIT_CAPI bool RabitFinalize(void);
RABIT_CAPI int RabitGetRingPrevRank(void);
RABIT_CAPI int RabitGetRank(void);
RABIT_CAPI int RabitGetWorldSize(void);
RABIT_CAPI int RabitIsDistributed(void);
RABIT_CAPI void RabitTrackerPrint(const char *msg);
/* copies at most max_len bytes of the host name into out_name, length into *out_len */
RABIT_CAPI void RabitGetProcessorName(char *out_name, rbt_ulong *out_len, rbt_ulong max_len);

/* ---- collectives (host or CUDA device pointers) --------------------------------- */
RABIT_CAPI void RabitBroadcast(void *sendrecv_data, rbt_ulong size, int root);
/* all counts in ELEMENTS of enum_dtype (the reference double-scales them) */
RABIT_CAPI void RabitAllgather(void *sendrecvbuf, size_t total_count, size_t begin_index,
                               size_t slice_count, size_t prev_slice_count, int enum_dtype);
RABIT_CAPI void RabitAllreduce(void *sendrecvbuf, size_t count, int enum_dtype, int enum_op,
                               void (*prepare_fun)(void *arg), void *prepare_arg);

/* ---- checkpoints: opaque blobs --------------------------------------------------- */
/* returned pointers stay valid until the next RabitLoadCheckPoint on this thread */
RABIT_CAPI int RabitLoadCheckPoint(char **out_global_model, rbt_ulong *out_global_len,
                                   char **out_local_model, rbt_ulong *out_local_len);
RABIT_CAPI void RabitCheckPoint(const char *global_model, rbt_ulong global_len,
                                const char *local_model, rbt_ulong local_len);
RABIT_CAPI int RabitVersionNumber(void);
RABIT_CAPI int RabitLinkTag(void);

/* ---- additions of this build ------------------------------------------------------ */
/* last error text of this thread ("" if none); C callers cannot catch C++ exceptions,
 * so every entry point above also records the message before propagating */
RABIT_CAPI const char *RabitGetLastError(void);
/* 1 when the sm_100a device backend is compiled in AND a usable GPU was found */
RABIT_CAPI int RabitDeviceAvailable(void);

#endif  /* RABIT_C_API_H_ */
