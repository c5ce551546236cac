/*!
 * rabit-b200: public C++ API — reliable Allreduce / Broadcast / Allgather with in-memory,
 * versioned checkpoints.
 *
 * Call-compatible with reference include/rabit/rabit.h:66-456: a program written for
 * rabit compiles against this header unchanged (same namespaces, names, argument order,
 * defaulted call-site arguments). What is different is underneath: buffers may live in
 * GPU memory, in which case the collective runs as ONE sm_100a kernel that reads/writes
 * peer GPUs over NVLink with the reduce operator fused in.
 *
 * Threading: like the reference (rabit.h:178) nothing here is thread-safe; each thread
 * owns its own engine instance.
 */
#ifndef RABIT_RABIT_H_
#define RABIT_RABIT_H_

#include <functional>
#include <string>
#include <vector>

#include "rabit/base.h"
#include "rabit/internal/engine.h"
#include "rabit/internal/io.h"
#include "rabit/internal/timer.h"
#include "rabit/internal/utils.h"
#include "rabit/serializable.h"

namespace rabit {

/*! \brief built-in reduce operators; definitions in internal/rabit-inl.h */
namespace op {
struct Max;
struct Min;
struct Sum;
struct BitOR;
}  // namespace op

// ----------------------------------------------------------------- life cycle
/*!
 * \brief start the engine. Parameters are `name=value` strings (argv overrides env, env
 *  overrides defaults); see doc/parameters.md. Without a tracker the process is a world
 *  of one and every collective is a no-op that still runs prepare_fun.
 */
inline bool Init(int argc, char *argv[]);
/*! \brief leave the job; blocks until no peer can still need data from this rank */
inline bool Finalize();
inline int GetRank();
inline int GetWorldSize();
/*! \return whether a tracker is configured (false = single process) */
inline bool IsDistributed();
inline std::string GetProcessorName();
/*! \brief rank of the ring predecessor (the rank whose local checkpoint we replicate) */
inline int GetRingPrevRank();
/*! \brief print one line in the tracker's log (or locally when there is no tracker) */
inline void TrackerPrint(const std::string &msg);
inline void TrackerPrintf(const char *fmt, ...);

// ---------------------------------------------------------------- collectives
/*! \brief root's bytes end up in everybody's buffer */
inline void Broadcast(void *sendrecv_data, size_t size, int root,
                      const char *_file = _FILE, const int _line = _LINE,
                      const char *_caller = _CALLER);
/*! \brief vector flavour: receivers are resized (length travels first) */
template <typename DType>
inline void Broadcast(std::vector<DType> *sendrecv_data, int root,
                      const char *_file = _FILE, const int _line = _LINE,
                      const char *_caller = _CALLER);
inline void Broadcast(std::string *sendrecv_data, int root,
                      const char *_file = _FILE, const int _line = _LINE,
                      const char *_caller = _CALLER);
/*!
 * \brief in-place allreduce: sendrecvbuf[i] = OP over ranks of sendrecvbuf[i].
 * \param prepare_fun lazy initialiser of sendrecvbuf; NOT called when the result of this
 *   call is recovered from a peer's cache after a restart
 */
template <typename OP, typename DType>
inline void Allreduce(DType *sendrecvbuf, size_t count,
                      void (*prepare_fun)(void *) = NULL, void *prepare_arg = NULL,
                      const char *_file = _FILE, const int _line = _LINE,
                      const char *_caller = _CALLER);
template <typename OP, typename DType>
inline void Allreduce(DType *sendrecvbuf, size_t count, std::function<void()> prepare_fun,
                      const char *_file = _FILE, const int _line = _LINE,
                      const char *_caller = _CALLER);
/*!
 * \brief ring allgather in element units. The buffer holds total_count items; this rank
 *  owns [begin_index, begin_index + slice_count); prev_slice_count is the size of the
 *  ring predecessor's slice. (The reference header names the 4th argument slice_end but
 *  its definition, rabit-inl.h:201-213, and all callers pass a LENGTH — kept as length.)
 */
template <typename DType>
inline void Allgather(DType *sendrecvbuf, size_t total_count, size_t begin_index,
                      size_t slice_count, size_t prev_slice_count,
                      const char *_file = _FILE, const int _line = _LINE,
                      const char *_caller = _CALLER);

// ---------------------------------------------------------------- checkpoints
/*!
 * \brief fetch the newest checkpoint. Returns its version; 0 means nothing was ever
 *  checkpointed and the models are untouched (caller initialises them).
 *    int it = rabit::LoadCheckPoint(&model);
 *    if (it == 0) model.Init();
 *    for (; it < n; ++it) { ...collectives...; rabit::CheckPoint(&model); }
 */
inline int LoadCheckPoint(Serializable *global_model, Serializable *local_model = NULL);
/*! \brief commit: global_model must be identical on all ranks, local_model is per rank */
inline void CheckPoint(const Serializable *global_model, const Serializable *local_model = NULL);
/*! \brief commit without serialising now; the model must stay unchanged until the last
 *   collective of the next version has returned */
inline void LazyCheckPoint(const Serializable *global_model);
/*! \return number of checkpoints committed so far */
inline int VersionNumber();

// ------------------------------------------------------------ custom reducers
namespace engine { class ReduceHandle; }

/*! \brief allreduce over a POD struct with a user merge function */
template <typename DType, void (*freduce)(DType &dst, const DType &src)>  // NOLINT(*)
class Reducer {
 public:
  Reducer();
  inline void Allreduce(DType *sendrecvbuf, size_t count,
                        void (*prepare_fun)(void *) = NULL, void *prepare_arg = NULL,
                        const char *_file = _FILE, const int _line = _LINE,
                        const char *_caller = _CALLER);
  inline void Allreduce(DType *sendrecvbuf, size_t count, std::function<void()> prepare_fun,
                        const char *_file = _FILE, const int _line = _LINE,
                        const char *_caller = _CALLER);
  /*!
   * \brief (new) declare freduce to be a lane-wise sum of `lanes` values of BaseT, which
   *  lets device buffers take the fused pair-sum kernel, e.g. for {double g, h}:
   *  `red.DeclarePackedSum<double>(2)`.
   */
  template <typename BaseT>
  inline void DeclarePackedSum(int lanes);

 private:
  engine::ReduceHandle handle_;
};

/*!
 * \brief allreduce over objects that serialise into at most max_nbyte bytes each.
 *  DType provides Load(Stream&), Save(Stream&) and Reduce(const DType&, size_t max_nbyte).
 */
template <typename DType>
class SerializeReducer {
 public:
  SerializeReducer();
  inline void Allreduce(DType *sendrecvobj, size_t max_nbyte, size_t count,
                        void (*prepare_fun)(void *) = NULL, void *prepare_arg = NULL,
                        const char *_file = _FILE, const int _line = _LINE,
                        const char *_caller = _CALLER);
  inline void Allreduce(DType *sendrecvobj, size_t max_nbyte, size_t count,
                        std::function<void()> prepare_fun,
                        const char *_file = _FILE, const int _line = _LINE,
                        const char *_caller = _CALLER);

 private:
  engine::ReduceHandle handle_;
  std::string buffer_;
};

}  // namespace rabit

#include "rabit/internal/rabit-inl.h"
#endif  // RABIT_RABIT_H_
