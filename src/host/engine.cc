/*!
 * rabit-b200: engine selection and the free functions behind rabit.h.
 *
 * The reference picks the engine at LINK time by compiling a different engine*.cc into
 * each library (engine.cc:19-27, engine_base.cc, engine_mock.cc, engine_empty.cc). Here
 * the choice is made at Init from `rabit_engine=robust|mock|base|empty`; each shipped
 * shared object only differs in its compiled default (RABIT_DEFAULT_ENGINE), so
 * python's lib='standard'|'mock' switch keeps working.
 * Like the reference (engine.cc:30-40) the instance is thread-local, and GetEngine()
 * before Init yields a usable world-of-one engine (engine.cc:71-82).
 */
#include <memory>

#include "engine_core.h"
#include "rabit/internal/thread_local.h"

#ifndef RABIT_DEFAULT_ENGINE
#define RABIT_DEFAULT_ENGINE "robust"
#endif

namespace rabit {
namespace engine {

BaseEngine *CreateNcclEngine();  // src/device/engine_nccl.cu (stub: NULL)

namespace {
struct ThreadEngine {
  std::unique_ptr<IEngine> engine;   // what GetEngine returns after Init
  BaseEngine *networked = nullptr;   // same object when it is Base/Robust/Mock
  EmptyEngine fallback;              // world of one, before Init / after Finalize
};
typedef ThreadLocalStore<ThreadEngine> Store;

std::string PickEngine(int argc, char *argv[]) {
  std::string name = RABIT_DEFAULT_ENGINE;
  const char *env = getenv("rabit_engine");
  if (env != nullptr) name = env;
  for (int i = 0; i < argc; ++i) {
    if (argv[i] != nullptr && strncmp(argv[i], "rabit_engine=", 13) == 0) name = argv[i] + 13;
  }
  return name;
}
}  // namespace

bool Init(int argc, char *argv[]) {
  ThreadEngine *te = Store::Get();
  utils::Check(te->engine == nullptr, "rabit::Init is already called in this thread");
  const std::string name = PickEngine(argc, argv);
  if (name == "empty") {
    te->engine.reset(new EmptyEngine());
    return true;
  }
  BaseEngine *e = nullptr;
  if (name == "robust" || name == "standard") e = new RobustEngine();
  else if (name == "mock") e = new MockEngine();
  else if (name == "base") e = new BaseEngine();
  else if (name == "nccl") {
    e = CreateNcclEngine();
    utils::Check(e != nullptr, "rabit_engine=nccl needs a library built with CUDA (python build.py)");
  }
  else utils::Error("unknown rabit_engine \"%s\" (robust|mock|base|empty|nccl)", name.c_str());
  te->engine.reset(e);
  te->networked = e;
  if (!e->Init(argc, argv)) {
    te->engine.reset();
    te->networked = nullptr;
    return false;
  }
  return true;
}

bool Finalize(void) {
  ThreadEngine *te = Store::Get();
  utils::Check(te->engine != nullptr, "rabit::Finalize engine is not initialized or already been finalized.");
  bool ok = true;
  if (te->networked != nullptr) ok = te->networked->Shutdown();
  te->engine.reset();
  te->networked = nullptr;
  return ok;
}

IEngine *GetEngine(void) {
  ThreadEngine *te = Store::Get();
  return te->engine ? te->engine.get() : &te->fallback;
}

/*! \brief internal: the networked engine of this thread or NULL (tests, device API) */
BaseEngine *GetNetworkedEngine(void) { return Store::Get()->networked; }

void Allgather(void *sendrecvbuf, size_t total_size, size_t slice_begin, size_t slice_end,
               size_t size_prev_slice, const char *_file, const int _line, const char *_caller) {
  GetEngine()->Allgather(sendrecvbuf, total_size, slice_begin, slice_end, size_prev_slice,
                         _file, _line, _caller);
}

void Allreduce_(void *sendrecvbuf, size_t type_nbytes, size_t count, IEngine::ReduceFunction red,
                mpi::DataType dtype, mpi::OpType op, IEngine::PreprocFunction prepare_fun,
                void *prepare_arg, const char *_file, const int _line, const char *_caller) {
  GetEngine()->AllreduceTyped(sendrecvbuf, type_nbytes, count, red, static_cast<int>(dtype),
                              static_cast<int>(op), prepare_fun, prepare_arg, _file, _line, _caller);
}

// ------------------------------------------------------------- ReduceHandle
ReduceHandle::ReduceHandle(void)
    : handle_(nullptr), redfunc_(nullptr), htype_(nullptr), created_type_nbytes_(0),
      packed_dtype_(-1), packed_lanes_(0) {}
ReduceHandle::~ReduceHandle(void) {}

int ReduceHandle::TypeSize(const MPI::Datatype &dtype) { return static_cast<int>(dtype.type_size); }

void ReduceHandle::Init(IEngine::ReduceFunction redfunc, size_t) {
  utils::Assert(redfunc_ == nullptr, "cannot initialize reduce handle twice");
  redfunc_ = redfunc;
}

void ReduceHandle::SetPackedSum(mpi::DataType dtype, int lanes) {
  packed_dtype_ = static_cast<int>(dtype);
  packed_lanes_ = lanes;
}

void ReduceHandle::Allreduce(void *sendrecvbuf, size_t type_nbytes, size_t count,
                             IEngine::PreprocFunction prepare_fun, void *prepare_arg,
                             const char *_file, const int _line, const char *_caller) {
  utils::Assert(redfunc_ != nullptr, "must initialize handle to call AllReduce");
  // a declared packed sum is, for the kernels, a plain sum over count*lanes base values
  const int dtype = packed_lanes_ > 0 ? packed_dtype_ : -1;
  const int op = packed_lanes_ > 0 ? static_cast<int>(mpi::kSum) : -1;
  GetEngine()->AllreduceTyped(sendrecvbuf, type_nbytes, count, redfunc_, dtype, op, prepare_fun,
                              prepare_arg, _file, _line, _caller);
}

}  // namespace engine
}  // namespace rabit
