/*!
 * rabit-b200: the engine contract that the public templates in rabit.h forward to.
 *
 * Same surface as reference include/rabit/internal/engine.h (IEngine 32-209, Init/
 * Finalize/GetEngine 212-216, mpi enums 221-239, Allreduce_ 283-293, ReduceHandle
 * 298-343), so code written against rabit::engine (xgboost does this) ports unchanged.
 * Differences:
 *   - the engine is picked at RUN time (`rabit_engine=robust|mock|base|empty`, default set
 *     by which shared object was loaded) instead of by linking a different engine*.cc;
 *   - every collective also accepts CUDA device pointers; they are routed to the sm_100a
 *     peer-memory kernels (src/device) with the reduce operator fused into the transfer.
 *     `AllreduceTyped` carries the (dtype, op) pair the kernels need to do that.
 */
#ifndef RABIT_INTERNAL_ENGINE_H_
#define RABIT_INTERNAL_ENGINE_H_

#include <string>

#include "rabit/base.h"
#include "rabit/serializable.h"

// call-site capture used as the bootstrap-cache key (reference rabit.h:29-37)
#ifndef RABIT_API_CALLER_SIGNATURE
#define RABIT_API_CALLER_SIGNATURE
#if defined(__GNUC__) && !defined(__clang__)
#define _FILE __builtin_FILE()
#define _LINE __builtin_LINE()
#define _CALLER __builtin_FUNCTION()
#else
#define _FILE "N/A"
#define _LINE -1
#define _CALLER "N/A"
#endif
#endif  // RABIT_API_CALLER_SIGNATURE

/*! \brief stand-in for the MPI C++ type handle: reducers only ever need the item size */
namespace MPI {
class Datatype {
 public:
  size_t type_size;
  explicit Datatype(size_t nbytes) : type_size(nbytes) {}
};
}  // namespace MPI

namespace rabit {
namespace engine {

namespace mpi {
/*! \brief reduce operators; numeric values are part of the C/python ABI */
enum OpType { kMax = 0, kMin = 1, kSum = 2, kBitwiseOR = 3 };
/*! \brief element types; numeric values are part of the C/python ABI */
enum DataType {
  kChar = 0, kUChar = 1, kInt = 2, kUInt = 3, kLong = 4, kULong = 5,
  kFloat = 6, kDouble = 7, kLongLong = 8, kULongLong = 9
};
}  // namespace mpi

/*! \brief abstract engine; one instance per thread, NOT thread-safe */
class IEngine {
 public:
  /*! \brief lazy initialiser of the send buffer, skipped when the result is recovered */
  typedef void(PreprocFunction)(void *arg);
  /*! \brief dst[i] = reduce(dst[i], src[i]) for i in [0, count) */
  typedef void(ReduceFunction)(const void *src, void *dst, int count, const MPI::Datatype &dtype);

  virtual ~IEngine() {}
  /*!
   * \brief ring-ordered gather: rank k owns bytes [slice_begin, slice_end) of a buffer of
   *  total_size bytes; afterwards every rank holds all of it
   * \param size_prev_slice byte size of rank (k-1)'s slice
   */
  virtual void Allgather(void *sendrecvbuf, size_t total_size, size_t slice_begin,
                         size_t slice_end, size_t size_prev_slice,
                         const char *_file = _FILE, const int _line = _LINE,
                         const char *_caller = _CALLER) = 0;
  /*! \brief in-place allreduce of count items of type_nbytes each with a host reducer */
  virtual void Allreduce(void *sendrecvbuf_, size_t type_nbytes, size_t count,
                         ReduceFunction reducer, PreprocFunction prepare_fun = NULL,
                         void *prepare_arg = NULL, const char *_file = _FILE,
                         const int _line = _LINE, const char *_caller = _CALLER) = 0;
  /*! \brief copy size bytes from root to everyone */
  virtual void Broadcast(void *sendrecvbuf_, size_t size, int root,
                         const char *_file = _FILE, const int _line = _LINE,
                         const char *_caller = _CALLER) = 0;
  /*! \brief re-join after the caller caught an engine exception in-process */
  virtual void InitAfterException(void) = 0;
  /*! \return version of the checkpoint loaded, 0 = nothing stored yet (fresh start) */
  virtual int LoadCheckPoint(Serializable *global_model, Serializable *local_model = NULL) = 0;
  /*! \brief commit a new version of the models; clears the per-version result cache */
  virtual void CheckPoint(const Serializable *global_model,
                          const Serializable *local_model = NULL) = 0;
  /*! \brief like CheckPoint but only remembers the pointer; Save runs if a peer needs it */
  virtual void LazyCheckPoint(const Serializable *global_model) = 0;
  virtual int VersionNumber(void) const = 0;
  virtual int GetRingPrevRank(void) const = 0;
  virtual int GetRank(void) const = 0;
  virtual int GetWorldSize(void) const = 0;
  virtual bool IsDistributed(void) const = 0;
  virtual std::string GetHost(void) const = 0;
  virtual void TrackerPrint(const std::string &msg) = 0;

  // ---- additions of this build (have defaults, so foreign IEngine subclasses still work) ----
  /*!
   * \brief allreduce that also knows which built-in (dtype, op) the reducer implements,
   *  which lets a device-pointer call run the fused sm_100a kernel.
   *  dtype/op < 0 means "opaque reducer" and always takes the host path.
   */
  virtual void AllreduceTyped(void *sendrecvbuf_, size_t type_nbytes, size_t count,
                              ReduceFunction reducer, int dtype, int op,
                              PreprocFunction prepare_fun, void *prepare_arg,
                              const char *_file, const int _line, const char *_caller) {
    this->Allreduce(sendrecvbuf_, type_nbytes, count, reducer, prepare_fun, prepare_arg,
                    _file, _line, _caller);
  }
};

/*! \brief construct + connect this thread's engine from argv `name=value` pairs and env */
bool Init(int argc, char *argv[]);
/*! \brief orderly shutdown (waits until no peer can still need recovery data) */
bool Finalize(void);
/*! \brief this thread's engine; a world-size-1 default before Init */
IEngine *GetEngine(void);

void Allgather(void *sendrecvbuf, size_t total_size, size_t slice_begin, size_t slice_end,
               size_t size_prev_slice, const char *_file = _FILE, const int _line = _LINE,
               const char *_caller = _CALLER);

void Allreduce_(void *sendrecvbuf, size_t type_nbytes, size_t count,
                IEngine::ReduceFunction red, mpi::DataType dtype, mpi::OpType op,
                IEngine::PreprocFunction prepare_fun = NULL, void *prepare_arg = NULL,
                const char *_file = _FILE, const int _line = _LINE,
                const char *_caller = _CALLER);

/*!
 * \brief handle for user-defined reducers (Reducer<> / SerializeReducer<> sit on top).
 *  `SetPackedSum` is new: it declares that the opaque reducer is an element-wise sum over
 *  `lanes` values of `dtype` per item (e.g. XGBoost's {double grad, double hess} pair is
 *  kDouble x 2), which is what allows device buffers to use the fused pair-sum kernel.
 */
class ReduceHandle {
 public:
  ReduceHandle(void);
  ~ReduceHandle(void);
  void Init(IEngine::ReduceFunction redfunc, size_t type_nbytes);
  void SetPackedSum(mpi::DataType dtype, int lanes);
  void Allreduce(void *sendrecvbuf, size_t type_nbytes, size_t count,
                 IEngine::PreprocFunction prepare_fun = NULL, void *prepare_arg = NULL,
                 const char *_file = _FILE, const int _line = _LINE,
                 const char *_caller = _CALLER);
  static int TypeSize(const MPI::Datatype &dtype);

 protected:
  void *handle_;
  IEngine::ReduceFunction *redfunc_;
  void *htype_;
  size_t created_type_nbytes_;
  int packed_dtype_;
  int packed_lanes_;
};

}  // namespace engine
}  // namespace rabit
#endif  // RABIT_INTERNAL_ENGINE_H_
