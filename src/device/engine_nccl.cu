/*!
 * rabit-b200: NcclEngine — a library-backed, NON-fault-tolerant engine, the counterpart of the
 * reference's MPI engine (src/engine_mpi.cc:20-247: MPI::COMM_WORLD.Allreduce / Bcast behind
 * IEngine, no recovery, Allgather unsupported). It exists as the comparator ("what does the vendor
 * library do on the same buffers through the same API"), never as the product path.
 *
 *   rabit_engine=nccl
 *
 * Device pointers go to ncclAllReduce / ncclBroadcast / ncclAllGather-equivalent; host pointers and
 * opaque reducers use the TCP collectives of BaseEngine. Rendezvous: the ncclUniqueId of rank 0
 * travels over the host engine's broadcast, so the tracker bootstrap is shared with every other
 * engine. libnccl is dlopen'ed: the library has no link-time dependency on it.
 */
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cstring>

#include "../host/engine_core.h"
#include "device_backend.h"

namespace rabit {
namespace engine {

namespace {
// the handful of NCCL declarations needed (ABI-stable since NCCL 2.0), so that <nccl.h> is not required
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef enum { ncclSuccess = 0 } ncclResult_t;
typedef enum { ncclSum = 0, ncclProd = 1, ncclMax = 2, ncclMin = 3 } ncclRedOp_t;
typedef enum { ncclInt8 = 0, ncclUint8 = 1, ncclInt32 = 2, ncclUint32 = 3, ncclInt64 = 4, ncclUint64 = 5,
               ncclFloat32 = 7, ncclFloat64 = 8 } ncclDataType_t;

struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
  bool Load() {
    if (handle != nullptr) return true;
    for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
      handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
      if (handle != nullptr) break;
    }
    if (handle == nullptr) return false;
    GetUniqueId = reinterpret_cast<decltype(GetUniqueId)>(dlsym(handle, "ncclGetUniqueId"));
    CommInitRank = reinterpret_cast<decltype(CommInitRank)>(dlsym(handle, "ncclCommInitRank"));
    CommDestroy = reinterpret_cast<decltype(CommDestroy)>(dlsym(handle, "ncclCommDestroy"));
    AllReduce = reinterpret_cast<decltype(AllReduce)>(dlsym(handle, "ncclAllReduce"));
    Broadcast = reinterpret_cast<decltype(Broadcast)>(dlsym(handle, "ncclBroadcast"));
    GetErrorString = reinterpret_cast<decltype(GetErrorString)>(dlsym(handle, "ncclGetErrorString"));
    return GetUniqueId && CommInitRank && CommDestroy && AllReduce && Broadcast && GetErrorString;
  }
};

class NcclEngine : public BaseEngine {
 public:
  const char *Name() const override { return "nccl"; }
  bool Init(int argc, char *argv[]) override {
    comm_.cfg.device = "off";  // the symmetric arena of the product path is not needed here
    if (!BaseEngine::Init(argc, argv)) return false;
    if (comm_.world <= 1) return true;
    utils::Check(api_.Load(), "rabit_engine=nccl: cannot load libnccl.so.2 (%s)", dlerror());
    const char *ord = getenv("RABIT_DEVICE_ORDINAL");
    if (ord != nullptr) {
      int ndev = 0;
      cudaGetDeviceCount(&ndev);
      utils::Check(ndev > 0, "rabit_engine=nccl needs a GPU");
      cudaSetDevice(atoi(ord) % ndev);
    }
    ncclUniqueId id;
    std::memset(&id, 0, sizeof(id));
    if (comm_.rank == 0) Call(api_.GetUniqueId(&id), "ncclGetUniqueId");
    Status st = comm_.Broadcast(&id, sizeof(id), 0);
    utils::Check(Ok(st), "nccl bootstrap broadcast failed: %s", StatusName(st));
    Call(api_.CommInitRank(&nccl_, comm_.world, id, comm_.rank), "ncclCommInitRank");
    return true;
  }
  bool Shutdown(void) override {
    if (nccl_ != nullptr) { api_.CommDestroy(nccl_); nccl_ = nullptr; }
    return BaseEngine::Shutdown();
  }
  void AllreduceTyped(void *buf, size_t type_nbytes, size_t count, ReduceFunction reducer, int dtype, int op,
                      PreprocFunction prepare_fun, void *prepare_arg, const char *f, const int l,
                      const char *c) override {
    if (nccl_ == nullptr || !DeviceBackend::IsDevicePointer(buf) || dtype < 0 || op < 0) {
      BaseEngine::AllreduceTyped(buf, type_nbytes, count, reducer, dtype, op, prepare_fun, prepare_arg, f, l, c);
      return;
    }
    if (prepare_fun != nullptr) prepare_fun(prepare_arg);
    static const ncclDataType_t kType[10] = {ncclInt8, ncclUint8, ncclInt32, ncclUint32, ncclInt64, ncclUint64,
                                             ncclFloat32, ncclFloat64, ncclInt64, ncclUint64};
    utils::Check(op != mpi::kBitwiseOR, "rabit_engine=nccl: NCCL has no bitwise-or reduction");
    static const ncclRedOp_t kOp[3] = {ncclMax, ncclMin, ncclSum};
    const size_t n = count * type_nbytes / DTypeBytes(dtype);
    Call(api_.AllReduce(buf, buf, n, kType[dtype], kOp[op], nccl_, 0), "ncclAllReduce");
    Sync();
  }
  void Broadcast(void *buf, size_t size, int root, const char *f, const int l, const char *c) override {
    if (nccl_ == nullptr || !DeviceBackend::IsDevicePointer(buf)) {
      BaseEngine::Broadcast(buf, size, root, f, l, c);
      return;
    }
    Call(api_.Broadcast(buf, buf, size, ncclInt8, root, nccl_, 0), "ncclBroadcast");
    Sync();
  }
  void Allgather(void *buf, size_t total, size_t begin, size_t end, size_t prev, const char *f, const int l,
                 const char *c) override {
    // like the MPI engine of the reference, the library engine offers no allgather of its own
    utils::Check(!DeviceBackend::IsDevicePointer(buf), "rabit_engine=nccl: Allgather on device buffers is not "
                 "supported (the reference's MPI engine has none either, engine_mpi.cc:25-34)");
    BaseEngine::Allgather(buf, total, begin, end, prev, f, l, c);
  }

 private:
  void Call(ncclResult_t r, const char *what) {
    utils::Check(r == ncclSuccess, "%s failed: %s", what, api_.GetErrorString != nullptr ? api_.GetErrorString(r) : "?");
  }
  void Sync() {
    cudaError_t e = cudaStreamSynchronize(0);
    utils::Check(e == cudaSuccess, "CUDA error after NCCL call: %s", cudaGetErrorString(e));
  }
  NcclApi api_;
  ncclComm_t nccl_ = nullptr;
};
}  // namespace

/*! \brief factory used by engine.cc (a host-only build links a stub that returns NULL) */
BaseEngine *CreateNcclEngine() { return new NcclEngine(); }

}  // namespace engine
}  // namespace rabit
