/*!
 * rabit-b200: the two one-thread kernels behind the peer-copy transport of recovery payloads and local
 * checkpoint replicas (device_backend_blob.inc). The bytes themselves are moved by copy engines
 * (cudaMemcpyAsync between mapped arenas) on a side stream; these kernels only order the two ends:
 *   SignalWord  st.release.sys of a counter into the OTHER rank's control block ("my window holds piece k",
 *               "I have taken piece k")
 *   WaitWord    bounded spin on a counter in THIS rank's control block; a dead peer becomes a status code
 * No counterpart in the reference, which streams these payloads over its TCP tree / ring
 * (allreduce_robust.cc:861-973, 1529-1587).
 */
#ifndef RABIT_SRC_DEVICE_COLL_BLOB_CUH_
#define RABIT_SRC_DEVICE_COLL_BLOB_CUH_

#include "coll_common.cuh"

namespace rabit {
namespace device {

__global__ void SignalWord(unsigned long long *remote, unsigned long long value) {
  StoreReleaseSys(remote, value);
}

__global__ void WaitWord(const unsigned long long *local, unsigned long long value, unsigned long long timeout_ns,
                         volatile int *abort_flag, int *status) {
  const unsigned long long deadline = GlobalTimer() + timeout_ns;
  unsigned int spins = 0;
  while (LoadAcquireSys(local) < value) {
    if ((++spins & 0xff) == 0) {
      if (*abort_flag != 0) { atomicCAS(status, kStOk, kStAborted); return; }
      if (GlobalTimer() > deadline) { atomicCAS(status, kStOk, kStTimeout); return; }
    }
  }
}

}  // namespace device
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_COLL_BLOB_CUH_
