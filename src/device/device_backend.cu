/*!
 * rabit-b200: sm_100a device backend — symmetric arena set-up and kernel dispatch.
 *
 * No counterpart in the reference. One process drives one GPU. At Attach every rank
 * allocates an arena with the CUDA VMM API (cuMemCreate, shareable as a POSIX fd), hands
 * the fd to the other ranks of the node over abstract unix sockets (SCM_RIGHTS) and maps
 * everybody's arena; the rendezvous itself (who is there, under which name) runs over
 * the host engine's ring, i.e. the tracker/TCP bootstrap is kept. If the VMM export is not
 * possible the arena falls back to cudaMalloc + cudaIpc handles. Where the driver supports
 * it, a multicast object (NVLS) is bound over all arenas so that kernels can use
 * multimem.ld_reduce / multimem.st (switch-side reduction and broadcast).
 *
 * The CUDA driver API is reached through cudaGetDriverEntryPoint, so the library links only
 * against the static runtime and still loads on machines without libcuda (CPU tests).
 */
#include <cuda.h>
#include <cuda_runtime.h>
#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstddef>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "coll_kernels.cuh"
#include "coll_tma.cuh"
#include "coll_nvls.cuh"
#include "coll_inplace.cuh"
#include "coll_stream.cuh"
#include "coll_blob.cuh"
#include "coll_gather.cuh"
#include "coll_ll.cuh"
#include "device_backend.h"
#include "../host/engine_core.h"
#include "rabit/internal/socket.h"
#include "rabit/internal/timer.h"

namespace rabit {
namespace engine {

using device::Params;
using device::kCtrlBytes;
using device::kMaxRanks;

#define CUDA_OK(expr)                                                                      \
  do {                                                                                     \
    cudaError_t e_ = (expr);                                                               \
    if (e_ != cudaSuccess) {                                                               \
      ::rabit::utils::Error("CUDA error %s at %s:%d: %s", cudaGetErrorName(e_), __FILE__, __LINE__, \
                   cudaGetErrorString(e_));                                                \
    }                                                                                      \
  } while (0)

// ------------------------------------------------------------ driver entry points
namespace {
struct Driver {
  bool ok = false;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
  CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*MemAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
  CUresult (*MemExportToShareableHandle)(void *, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long) = nullptr;
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle *, void *, CUmemAllocationHandleType) = nullptr;
  CUresult (*MemGetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle *, const CUmulticastObjectProp *) = nullptr;
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t,
                               unsigned long long) = nullptr;
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
  CUresult (*MulticastGetGranularity)(size_t *, const CUmulticastObjectProp *, CUmulticastGranularity_flags) = nullptr;
  CUresult (*DeviceGetAttribute)(int *, CUdevice_attribute, CUdevice) = nullptr;
  CUresult (*DeviceGet)(CUdevice *, int) = nullptr;
  CUresult (*GetErrorString)(CUresult, const char **) = nullptr;
};

template <typename F>
bool Resolve(const char *name, F *fn) {
  void *p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || p == nullptr ||
      q != cudaDriverEntryPointSuccess) {
    cudaGetLastError();
    return false;
  }
  *fn = reinterpret_cast<F>(p);
  return true;
}

Driver &Drv() {
  static Driver d;
  static bool tried = false;
  if (tried) return d;
  tried = true;
  bool ok = true;
  ok &= Resolve("cuMemCreate", &d.MemCreate);
  ok &= Resolve("cuMemRelease", &d.MemRelease);
  ok &= Resolve("cuMemAddressReserve", &d.MemAddressReserve);
  ok &= Resolve("cuMemAddressFree", &d.MemAddressFree);
  ok &= Resolve("cuMemMap", &d.MemMap);
  ok &= Resolve("cuMemUnmap", &d.MemUnmap);
  ok &= Resolve("cuMemSetAccess", &d.MemSetAccess);
  ok &= Resolve("cuMemExportToShareableHandle", &d.MemExportToShareableHandle);
  ok &= Resolve("cuMemImportFromShareableHandle", &d.MemImportFromShareableHandle);
  ok &= Resolve("cuMemGetAllocationGranularity", &d.MemGetAllocationGranularity);
  ok &= Resolve("cuDeviceGetAttribute", &d.DeviceGetAttribute);
  ok &= Resolve("cuDeviceGet", &d.DeviceGet);
  ok &= Resolve("cuGetErrorString", &d.GetErrorString);
  // multicast is optional
  Resolve("cuMulticastCreate", &d.MulticastCreate);
  Resolve("cuMulticastAddDevice", &d.MulticastAddDevice);
  Resolve("cuMulticastBindMem", &d.MulticastBindMem);
  Resolve("cuMulticastUnbind", &d.MulticastUnbind);
  Resolve("cuMulticastGetGranularity", &d.MulticastGetGranularity);
  d.ok = ok;
  return d;
}

const char *CuErr(CUresult r) {
  const char *s = "?";
  if (Drv().GetErrorString != nullptr) Drv().GetErrorString(r, &s);
  return s != nullptr ? s : "?";
}

inline size_t RoundUp(size_t v, size_t unit) { return (v + unit - 1) / unit * unit; }
}  // namespace

}  // namespace engine
}  // namespace rabit

#include "device_backend_impl.inc"
