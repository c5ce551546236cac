/*!
 * rabit-b200: definitions behind rabit.h.
 *
 * Parity map (reference include/rabit/internal/rabit-inl.h):
 *   TypeTag<T>           <- mpi::GetType<T> specialisations (21-62)
 *   op::*::Reduce        <- 67-94   (here __host__ __device__, so the SAME functor is the CPU
 *                                    oracle and the body of the fused CUDA kernels)
 *   ReduceLoop           <- op::Reducer (95-102)
 *   StructReduce*        <- ReducerSafe_/ReducerAlign_ (255-288)
 *   SerialMerge/Closure  <- SerializeReducerFunc_/SerializeReduceClosure (300-362)
 */
#ifndef RABIT_INTERNAL_RABIT_INL_H_
#define RABIT_INTERNAL_RABIT_INL_H_

#include <cstdarg>
#include <cstring>
#include <string>
#include <vector>

namespace rabit {
namespace engine {
namespace mpi {
/*! \brief compile-time C type -> DataType tag */
template <typename T> struct TypeTag;
#define RABIT_DECL_TYPETAG(T, TAG) \
  template <> struct TypeTag<T> { static const DataType value = TAG; }
RABIT_DECL_TYPETAG(char, kChar);
RABIT_DECL_TYPETAG(signed char, kChar);
RABIT_DECL_TYPETAG(unsigned char, kUChar);
RABIT_DECL_TYPETAG(int, kInt);
RABIT_DECL_TYPETAG(unsigned int, kUInt);
RABIT_DECL_TYPETAG(long, kLong);                 // NOLINT(*)
RABIT_DECL_TYPETAG(unsigned long, kULong);       // NOLINT(*)
RABIT_DECL_TYPETAG(float, kFloat);
RABIT_DECL_TYPETAG(double, kDouble);
RABIT_DECL_TYPETAG(long long, kLongLong);            // NOLINT(*)
RABIT_DECL_TYPETAG(unsigned long long, kULongLong);  // NOLINT(*)
#undef RABIT_DECL_TYPETAG
template <typename T>
inline DataType GetType(void) { return TypeTag<T>::value; }
}  // namespace mpi
}  // namespace engine

namespace op {
struct Max {
  static const engine::mpi::OpType kType = engine::mpi::kMax;
  template <typename T>
  RABIT_HOST_DEVICE static void Reduce(T &dst, const T &src) {  // NOLINT(*)
    dst = dst < src ? src : dst;  // unconditional store: the host loop below vectorises (maxps / pmax)
  }
};
struct Min {
  static const engine::mpi::OpType kType = engine::mpi::kMin;
  template <typename T>
  RABIT_HOST_DEVICE static void Reduce(T &dst, const T &src) {  // NOLINT(*)
    dst = src < dst ? src : dst;
  }
};
struct Sum {
  static const engine::mpi::OpType kType = engine::mpi::kSum;
  template <typename T>
  RABIT_HOST_DEVICE static void Reduce(T &dst, const T &src) {  // NOLINT(*)
    dst += src;
  }
};
struct BitOR {
  static const engine::mpi::OpType kType = engine::mpi::kBitwiseOR;
  template <typename T>
  RABIT_HOST_DEVICE static void Reduce(T &dst, const T &src) {  // NOLINT(*)
    dst |= src;
  }
};
/*! \brief host reducer with the engine's ReduceFunction signature */
template <typename OP, typename T>
inline void Reducer(const void *src_, void *dst_, int len, const MPI::Datatype &) {
  // the engine never folds a buffer into itself
  const T *__restrict__ src = static_cast<const T *>(src_);
  T *__restrict__ dst = static_cast<T *>(dst_);
  for (int i = 0; i < len; ++i) OP::Reduce(dst[i], src[i]);
}
}  // namespace op

// ------------------------------------------------------------------ forwards
inline bool Init(int argc, char *argv[]) { return engine::Init(argc, argv); }
inline bool Finalize(void) { return engine::Finalize(); }
inline int GetRank(void) { return engine::GetEngine()->GetRank(); }
inline int GetWorldSize(void) { return engine::GetEngine()->GetWorldSize(); }
inline bool IsDistributed(void) { return engine::GetEngine()->IsDistributed(); }
inline std::string GetProcessorName(void) { return engine::GetEngine()->GetHost(); }
inline int GetRingPrevRank(void) { return engine::GetEngine()->GetRingPrevRank(); }
inline void TrackerPrint(const std::string &msg) { engine::GetEngine()->TrackerPrint(msg); }
inline void TrackerPrintf(const char *fmt, ...) {
  char line[1 << 10];
  va_list ap; va_start(ap, fmt);
  vsnprintf(line, sizeof(line), fmt, ap);
  va_end(ap);
  TrackerPrint(std::string(line));
}

inline void Broadcast(void *sendrecv_data, size_t size, int root,
                      const char *_file, const int _line, const char *_caller) {
  engine::GetEngine()->Broadcast(sendrecv_data, size, root, _file, _line, _caller);
}
template <typename DType>
inline void Broadcast(std::vector<DType> *data, int root,
                      const char *_file, const int _line, const char *_caller) {
  // two engine ops (two sequence numbers): element count first, then the payload
  size_t n = data->size();
  Broadcast(&n, sizeof(n), root, _file, _line, _caller);
  if (data->size() != n) data->resize(n);
  if (n != 0) Broadcast(data->data(), n * sizeof(DType), root, _file, _line, _caller);
}
inline void Broadcast(std::string *data, int root,
                      const char *_file, const int _line, const char *_caller) {
  size_t n = data->length();
  Broadcast(&n, sizeof(n), root, _file, _line, _caller);
  if (data->length() != n) data->resize(n);
  if (n != 0) Broadcast(&(*data)[0], n, root, _file, _line, _caller);
}

template <typename OP, typename DType>
inline void Allreduce(DType *sendrecvbuf, size_t count, void (*prepare_fun)(void *),
                      void *prepare_arg, const char *_file, const int _line,
                      const char *_caller) {
  engine::Allreduce_(sendrecvbuf, sizeof(DType), count, op::Reducer<OP, DType>,
                     engine::mpi::GetType<DType>(), OP::kType, prepare_fun, prepare_arg,
                     _file, _line, _caller);
}
/*! \brief trampoline: void* -> std::function<void()> */
inline void InvokeLambda_(void *fun) { (*static_cast<std::function<void()> *>(fun))(); }

template <typename OP, typename DType>
inline void Allreduce(DType *sendrecvbuf, size_t count, std::function<void()> prepare_fun,
                      const char *_file, const int _line, const char *_caller) {
  engine::Allreduce_(sendrecvbuf, sizeof(DType), count, op::Reducer<OP, DType>,
                     engine::mpi::GetType<DType>(), OP::kType, InvokeLambda_, &prepare_fun,
                     _file, _line, _caller);
}
template <typename DType>
inline void Allgather(DType *sendrecvbuf, size_t total_count, size_t begin_index,
                      size_t slice_count, size_t prev_slice_count,
                      const char *_file, const int _line, const char *_caller) {
  const size_t w = sizeof(DType);
  engine::GetEngine()->Allgather(sendrecvbuf, total_count * w, begin_index * w,
                                 (begin_index + slice_count) * w, prev_slice_count * w,
                                 _file, _line, _caller);
}

inline int LoadCheckPoint(Serializable *global_model, Serializable *local_model) {
  return engine::GetEngine()->LoadCheckPoint(global_model, local_model);
}
inline void CheckPoint(const Serializable *global_model, const Serializable *local_model) {
  engine::GetEngine()->CheckPoint(global_model, local_model);
}
inline void LazyCheckPoint(const Serializable *global_model) {
  engine::GetEngine()->LazyCheckPoint(global_model);
}
inline int VersionNumber(void) { return engine::GetEngine()->VersionNumber(); }

// ------------------------------------------------------------ custom reducers
namespace detail {
/*! \brief struct items of any size/alignment: bounce through aligned temporaries */
template <typename DType, void (*freduce)(DType &, const DType &)>  // NOLINT(*)
inline void StructReduceUnaligned(const void *src_, void *dst_, int len, const MPI::Datatype &) {
  const char *s = static_cast<const char *>(src_);
  char *d = static_cast<char *>(dst_);
  for (int i = 0; i < len; ++i, s += sizeof(DType), d += sizeof(DType)) {
    DType a, b;
    std::memcpy(&a, d, sizeof(DType));
    std::memcpy(&b, s, sizeof(DType));
    freduce(a, b);
    std::memcpy(d, &a, sizeof(DType));
  }
}
/*! \brief item sizes 1/4/8 are always naturally aligned inside the engine's buffers */
template <typename DType, void (*freduce)(DType &, const DType &)>  // NOLINT(*)
inline void StructReduceAligned(const void *src_, void *dst_, int len, const MPI::Datatype &) {
  const DType *s = static_cast<const DType *>(src_);
  DType *d = static_cast<DType *>(dst_);
  for (int i = 0; i < len; ++i) freduce(d[i], s[i]);
}
/*! \brief merge serialised objects slot by slot */
template <typename DType>
inline void SerialMerge(const void *src_, void *dst_, int len, const MPI::Datatype &dtype) {
  const int slot = engine::ReduceHandle::TypeSize(dtype);
  char *s = const_cast<char *>(static_cast<const char *>(src_));
  char *d = static_cast<char *>(dst_);
  for (int i = 0; i < len; ++i, s += slot, d += slot) {
    DType from, into;
    utils::MemoryFixSizeBuffer fs(s, slot), fd(d, slot);
    from.Load(fs);
    into.Load(fd);
    into.Reduce(static_cast<const DType &>(from), slot);
    fd.Seek(0);
    into.Save(fd);
  }
}
/*! \brief lazy "serialise my objects into the slots" step, run as prepare_fun */
template <typename DType>
struct SerialPrepare {
  DType *objs;
  size_t slot, count;
  void (*user_fun)(void *);
  void *user_arg;
  std::string *store;
  static void Run(void *self_) {
    SerialPrepare *self = static_cast<SerialPrepare *>(self_);
    if (self->user_fun != NULL) self->user_fun(self->user_arg);
    for (size_t i = 0; i < self->count; ++i) {
      utils::MemoryFixSizeBuffer fs(&(*self->store)[0] + i * self->slot, self->slot);
      self->objs[i].Save(fs);
    }
  }
};
}  // namespace detail

template <typename DType, void (*freduce)(DType &, const DType &)>  // NOLINT(*)
inline Reducer<DType, freduce>::Reducer(void) {
  if (sizeof(DType) == 1 || sizeof(DType) == 4 || sizeof(DType) == 8) {
    handle_.Init(detail::StructReduceAligned<DType, freduce>, sizeof(DType));
  } else {
    handle_.Init(detail::StructReduceUnaligned<DType, freduce>, sizeof(DType));
  }
}
template <typename DType, void (*freduce)(DType &, const DType &)>  // NOLINT(*)
template <typename BaseT>
inline void Reducer<DType, freduce>::DeclarePackedSum(int lanes) {
  utils::Check(sizeof(BaseT) * lanes == sizeof(DType),
               "DeclarePackedSum: %d lanes of %lu bytes do not tile a %lu byte item", lanes,
               static_cast<unsigned long>(sizeof(BaseT)),     // NOLINT(*)
               static_cast<unsigned long>(sizeof(DType)));    // NOLINT(*)
  handle_.SetPackedSum(engine::mpi::GetType<BaseT>(), lanes);
}
template <typename DType, void (*freduce)(DType &, const DType &)>  // NOLINT(*)
inline void Reducer<DType, freduce>::Allreduce(DType *sendrecvbuf, size_t count,
                                               void (*prepare_fun)(void *), void *prepare_arg,
                                               const char *_file, const int _line,
                                               const char *_caller) {
  handle_.Allreduce(sendrecvbuf, sizeof(DType), count, prepare_fun, prepare_arg,
                    _file, _line, _caller);
}
template <typename DType, void (*freduce)(DType &, const DType &)>  // NOLINT(*)
inline void Reducer<DType, freduce>::Allreduce(DType *sendrecvbuf, size_t count,
                                               std::function<void()> prepare_fun,
                                               const char *_file, const int _line,
                                               const char *_caller) {
  this->Allreduce(sendrecvbuf, count, InvokeLambda_, &prepare_fun, _file, _line, _caller);
}

template <typename DType>
inline SerializeReducer<DType>::SerializeReducer(void) {
  handle_.Init(detail::SerialMerge<DType>, sizeof(DType));
}
template <typename DType>
inline void SerializeReducer<DType>::Allreduce(DType *sendrecvobj, size_t max_nbyte, size_t count,
                                               void (*prepare_fun)(void *), void *prepare_arg,
                                               const char *_file, const int _line,
                                               const char *_caller) {
  buffer_.resize(max_nbyte * count);
  detail::SerialPrepare<DType> prep;
  prep.objs = sendrecvobj; prep.slot = max_nbyte; prep.count = count;
  prep.user_fun = prepare_fun; prep.user_arg = prepare_arg; prep.store = &buffer_;
  handle_.Allreduce(BeginPtr(buffer_), max_nbyte, count, detail::SerialPrepare<DType>::Run, &prep,
                    _file, _line, _caller);
  for (size_t i = 0; i < count; ++i) {
    utils::MemoryFixSizeBuffer fs(BeginPtr(buffer_) + i * max_nbyte, max_nbyte);
    sendrecvobj[i].Load(fs);
  }
}
template <typename DType>
inline void SerializeReducer<DType>::Allreduce(DType *sendrecvobj, size_t max_nbyte, size_t count,
                                               std::function<void()> prepare_fun,
                                               const char *_file, const int _line,
                                               const char *_caller) {
  this->Allreduce(sendrecvobj, max_nbyte, count, InvokeLambda_, &prepare_fun,
                  _file, _line, _caller);
}
}  // namespace rabit
#endif  // RABIT_INTERNAL_RABIT_INL_H_
