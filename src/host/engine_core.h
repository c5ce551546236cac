/*!
 * rabit-b200: engine classes.
 *
 *   EmptyEngine   world of one, no transport      (reference src/engine_empty.cc:17-91)
 *   BaseEngine    collectives, no fault tolerance (reference AllreduceBase as an IEngine)
 *   RobustEngine  + result cache, consensus, checkpoints, recovery
 *                                                  (reference src/allreduce_robust.{h,cc})
 *   MockEngine    + scripted failures and timing stats (reference src/allreduce_mock.h)
 *
 * All of them accept CUDA device pointers; Base/Robust route those through DeviceBackend
 * (src/device) when it is available and otherwise stage through host memory.
 */
#ifndef RABIT_SRC_HOST_ENGINE_CORE_H_
#define RABIT_SRC_HOST_ENGINE_CORE_H_

#include <atomic>
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "comm.h"
#include "protocol.h"
#include "rabit/internal/engine.h"
#include "rabit/internal/io.h"
#include "rabit/internal/timer.h"

namespace rabit {
namespace engine {

class DeviceBackend;  // src/device/device_backend.h

/*! \brief where a pointer lives, as far as the engines care */
enum class MemKind { kHost, kDevice };

// ------------------------------------------------------------------ EmptyEngine
class EmptyEngine : public IEngine {
 public:
  void Allgather(void *, size_t, size_t, size_t, size_t, const char *, const int, const char *) override {}
  void Allreduce(void *, size_t, size_t, ReduceFunction, PreprocFunction prepare_fun, void *prepare_arg,
                 const char *, const int, const char *) override {
    if (prepare_fun != nullptr) prepare_fun(prepare_arg);
  }
  void Broadcast(void *, size_t, int, const char *, const int, const char *) override {}
  void InitAfterException(void) override {}
  int LoadCheckPoint(Serializable *, Serializable *) override { return 0; }
  void CheckPoint(const Serializable *, const Serializable *) override { ++version_; }
  void LazyCheckPoint(const Serializable *) override { ++version_; }
  int VersionNumber(void) const override { return version_; }
  int GetRingPrevRank(void) const override { return 0; }
  int GetRank(void) const override { return 0; }
  int GetWorldSize(void) const override { return 1; }
  bool IsDistributed(void) const override { return false; }
  std::string GetHost(void) const override { return std::string(); }
  void TrackerPrint(const std::string &msg) override { utils::Printf("%s", msg.c_str()); }

 private:
  int version_ = 0;
};

// ------------------------------------------------------------------- BaseEngine
class BaseEngine : public IEngine {
 public:
  BaseEngine();
  ~BaseEngine() override;
  /*! \brief parse parameters, meet the tracker, build links (and the device arena) */
  virtual bool Init(int argc, char *argv[]);
  virtual bool Shutdown(void);
  virtual void SetParam(const char *name, const char *val) { comm_.cfg.Set(name, val); }
  virtual const char *Name() const { return "base"; }

  void Allgather(void *sendrecvbuf, size_t total_size, size_t slice_begin, size_t slice_end,
                 size_t size_prev_slice, const char *_file, const int _line,
                 const char *_caller) override;
  void Allreduce(void *sendrecvbuf_, size_t type_nbytes, size_t count, ReduceFunction reducer,
                 PreprocFunction prepare_fun, void *prepare_arg, const char *_file,
                 const int _line, const char *_caller) override {
    this->AllreduceTyped(sendrecvbuf_, type_nbytes, count, reducer, -1, -1, prepare_fun,
                         prepare_arg, _file, _line, _caller);
  }
  void AllreduceTyped(void *sendrecvbuf_, size_t type_nbytes, size_t count, ReduceFunction reducer,
                      int dtype, int op, PreprocFunction prepare_fun, void *prepare_arg,
                      const char *_file, const int _line, const char *_caller) override;
  void Broadcast(void *sendrecvbuf_, size_t size, int root, const char *_file, const int _line,
                 const char *_caller) override;
  void InitAfterException(void) override {
    utils::Error("InitAfterException: the base engine is not fault tolerant");
  }
  int LoadCheckPoint(Serializable *, Serializable *) override { return 0; }
  void CheckPoint(const Serializable *, const Serializable *) override { ++version_number_; }
  void LazyCheckPoint(const Serializable *) override { ++version_number_; }
  int VersionNumber(void) const override { return version_number_; }
  int GetRingPrevRank(void) const override { return comm_.ring_prev_rank; }
  int GetRank(void) const override { return comm_.rank; }
  int GetWorldSize(void) const override { return comm_.world < 1 ? 1 : comm_.world; }
  bool IsDistributed(void) const override { return comm_.distributed(); }
  std::string GetHost(void) const override { return comm_.host; }
  void TrackerPrint(const std::string &msg) override { comm_.TrackerPrint(msg); }

  /*! \brief transport, exposed for the device backend's bootstrap and for tests */
  HostComm &comm() { return comm_; }
  DeviceBackend *device() { return device_.get(); }
  /*! \brief classify a user pointer (host unless CUDA says otherwise) */
  MemKind Classify(const void *ptr) const;
  /*! \brief outcome of the collective device bootstrap that follows every rendezvous */
  enum class AttachResult { kAttached, kNoDevice, kLinkError };
  /*! \brief agree with all ranks on using the device backend, then build the arena (or drop it everywhere) */
  AttachResult AttachDevice();
  /*! \brief the device bootstrap that follows the FIRST rendezvous of this process (Init) */
  virtual void AttachAfterStart();

 protected:
  /*! \brief Hadoop streaming status line (reference allreduce_base.h:248-253) */
  void ReportStatus() const {
    if (comm_.cfg.hadoop_mode != 0) {
      fprintf(stderr, "reporter:status:Rabit Phase[%03d] Operation %03d\n", version_number_, seq_);
    }
  }
  HostComm comm_;
  std::unique_ptr<DeviceBackend> device_;
  int version_number_ = 0;
  int seq_ = 0;  // collectives since the last checkpoint
  bool initialized_ = false;
};

// ----------------------------------------------------------------- RobustEngine
/*! \brief bytes per element of an mpi::DataType value (0 if unknown) */
inline size_t DTypeBytes(int dtype) {
  static const size_t w[] = {1, 1, 4, 4, sizeof(long), sizeof(unsigned long), 4, 8, 8, 8};  // NOLINT(*)
  return (dtype >= 0 && dtype < 10) ? w[dtype] : 0;
}

/*!
 * \brief seq-indexed byte blobs: the per-version result cache. An entry lives either in
 *  host memory or in GPU memory (written there by the collective kernel's epilogue); a host
 *  view of a device entry is materialised only if a peer actually asks for it.
 */
/*!
 * \brief host bytes of a cache entry: uninitialised and recycled. A fresh std::string per collective costs a
 *  zero fill plus one page fault per 4 KiB of a large message every time; blocks of 64 KiB and more therefore
 *  go back to a small per-thread pool and are handed out again, still mapped.
 */
class HostBytes {
 public:
  HostBytes() {}
  HostBytes(HostBytes &&o) noexcept : p_(o.p_), size_(o.size_), cap_(o.cap_) { o.p_ = nullptr; o.size_ = o.cap_ = 0; }
  HostBytes &operator=(HostBytes &&o) noexcept {
    if (this != &o) { clear(); p_ = o.p_; size_ = o.size_; cap_ = o.cap_; o.p_ = nullptr; o.size_ = o.cap_ = 0; }
    return *this;
  }
  HostBytes(const HostBytes &) = delete;
  HostBytes &operator=(const HostBytes &) = delete;
  ~HostBytes() { clear(); }
  /*! \brief make room for n bytes; the content is NOT preserved and NOT initialised */
  void resize(size_t n);
  /*! \brief give the block back */
  void clear();
  char *data() { return p_; }
  const char *data() const { return p_; }
  size_t size() const { return size_; }
  char &operator[](size_t i) { return p_[i]; }

 private:
  char *p_ = nullptr;
  size_t size_ = 0, cap_ = 0;
};

class ResultStore {
 public:
  struct Entry {
    int seq = -1;
    size_t bytes = 0;
    HostBytes host;
    void *dev = nullptr;
  };
  ~ResultStore() { Clear(); }
  void SetDevice(DeviceBackend *d) { device_ = d; }
  /*! \brief host scratch for the op in flight; becomes an entry with Commit */
  char *AllocTemp(size_t nbytes);
  /*! \brief device scratch for the op in flight */
  void *AllocTempDevice(size_t nbytes);
  void DiscardTemp();
  void Commit(int seq);
  void Clear();
  int LastSeq() const { return entries_.empty() ? -1 : entries_.back().seq; }
  void DropLast();
  Entry *Find(int seq);
  /*! \brief host-readable bytes of an entry (copies device entries down on first use) */
  char *HostView(Entry *e);
  size_t Count() const { return entries_.size(); }
  size_t DeviceBytes() const { return device_bytes_; }
  /*! \brief move the oldest device entries to host memory until at most `budget` bytes stay on the GPU */
  void SpillTo(size_t budget);

 private:
  void Release(Entry *e);
  std::vector<Entry> entries_;
  Entry temp_;
  DeviceBackend *device_ = nullptr;
  size_t device_bytes_ = 0;
};

class RobustEngine : public BaseEngine {
 public:
  RobustEngine();
  ~RobustEngine() override;
  bool Init(int argc, char *argv[]) override;
  bool Shutdown(void) override;
  const char *Name() const override { return "robust"; }

  void Allgather(void *sendrecvbuf, size_t total_size, size_t slice_begin, size_t slice_end,
                 size_t size_prev_slice, const char *_file, const int _line,
                 const char *_caller) override;
  void AllreduceTyped(void *sendrecvbuf_, size_t type_nbytes, size_t count, ReduceFunction reducer,
                      int dtype, int op, PreprocFunction prepare_fun, void *prepare_arg,
                      const char *_file, const int _line, const char *_caller) override;
  void Broadcast(void *sendrecvbuf_, size_t size, int root, const char *_file, const int _line,
                 const char *_caller) override;
  void InitAfterException(void) override;
  int LoadCheckPoint(Serializable *global_model, Serializable *local_model) override;
  void CheckPoint(const Serializable *global_model, const Serializable *local_model) override {
    this->Commit(global_model, local_model, false);
  }
  void LazyCheckPoint(const Serializable *global_model) override {
    this->Commit(global_model, nullptr, true);
  }

  // ---- hooks for tests (reference allreduce_robust.h:663-666) ----
  /*! \brief called by the watchdog when recovery does not finish in time; default aborts */
  void (*on_timeout)(int rank, int seconds) = nullptr;
  /*! \brief number of link failures handled so far */
  int recover_count() const { return recover_counter_; }
  uint64_t nvlink_payload_bytes() const { return nvlink_payload_bytes_; }
  uint64_t tcp_payload_bytes() const { return tcp_payload_bytes_; }
  /*! \brief feed a primitive's status through the failure handler (tests) */
  bool HandleStatus(Status s) { return this->Healthy(s); }
  /*! \brief true while a watchdog thread is counting down */
  bool watchdog_running() const { return watchdog_.joinable() && !watchdog_done_.load(); }
  bool watchdog_fired() const { return watchdog_fired_.load(); }

 protected:
  /*! \brief one collective's worth of work on host memory, parameterised by the op kind */
  enum class OpKind { kAllreduce, kBroadcast, kAllgather };
  struct OpArgs {
    OpKind kind;
    void *buf; size_t bytes;
    size_t elem_bytes, count; ReduceFunction *reducer; int dtype, op;  // allreduce
    int root;                                                          // broadcast
    size_t slice_begin, slice_end, size_prev_slice;                    // allgather
    PreprocFunction *prepare_fun; void *prepare_arg;
    std::string key;
  };
  void RunOp(OpArgs *a);
  /*! \brief does this allreduce on a HOST buffer go through the device kernels (H2D | kernel | D2H)? */
  bool HostOverDevice(const OpArgs &a) const;
  Status TryOp(const OpArgs &a, char *slot, const char *input);

  /*!
   * \brief the consensus loop. Returns true when the request was satisfied by recovery
   *  (barrier passed, or `buf` filled with the cached result), false when the caller must
   *  now execute the step itself.
   */
  bool Consensus(void *buf, size_t size, uint32_t flags, uint32_t seq,
                 const std::function<void *()> *lazy_buf = nullptr, proto::InlineOp *inline_op = nullptr);
  /*! \brief failure handler: kOk -> true; otherwise tear down, re-rendezvous, return false */
  bool Healthy(Status s);
  void StartWatchdog();
  void StopWatchdog();
  void Rejoin();
  /*! \brief a peer may die while this process is still in Init (it is the replacement of another one, or the job is
   *  just starting): that is one more recovery, not the end of this worker */
  void AttachAfterStart() override;

  // ---- data movers run by every rank once Decide() says so ----
  Status ServeResult(void *buf, size_t size, uint32_t seq, bool requester);
  Status ServeLoad(bool requester);
  Status ServeCache(bool requester, uint32_t max_count);
  /*! \brief make every rank hold its own + its `k` ring predecessors' local blobs */
  Status CompleteLocal(int which);
  /*! \brief gather `nwords` 64-bit words from every rank into a world x nwords table */
  Status GatherTable(const uint64_t *mine, int nwords, std::vector<uint64_t> *table);
  /*!
   * \brief recovery payloads go over NVLink instead of the TCP ring when the device arena is up: the lowest-ranked
   *  holder puts the blob into its arena window, the requesters' copy engines pull it from there on a side
   *  stream; nobody else takes part. Returns false if that path is not usable (then the ring delivers).
   */
  bool DeviceDeliver(const std::vector<char> &has, const std::vector<char> &need, const void *src_host,
                     const void *src_dev, void *dst_host, size_t size, Status *st);
  /*! \brief move one blob from its holders to the ranks that need it, along the ring */
  Status RingDeliver(const std::vector<char> &has, const std::vector<char> &need, void *buf, size_t size);

  void Commit(const Serializable *global_model, const Serializable *local_model, bool lazy);
  void LocalModelCheck(bool with_local);
  void MaterializeLazy();
  void DropStaleResult();
  int ReplicaRound() const;

  // bootstrap cache (ops issued before LoadCheckPoint when rabit_bootstrap_cache=1)
  bool BootLookup(const std::string &key, void *buf, size_t size) const;
  std::string BootSerialize() const;
  void BootDeserialize(const std::string &blob);

  ResultStore results_;
  std::vector<std::pair<std::string, std::string> > boot_cache_;  // key -> result, insertion order
  bool checkpoint_loaded_ = false;
  /*! \brief all ranks are known to be at the same op: device collectives may use the fused kernel path */
  bool in_sync_ = false;
  std::string global_blob_;                  // [int32 version][Save bytes]
  const Serializable *lazy_model_ = nullptr;
  /*! \brief inside Commit/Shutdown, before barrier 1 has passed: the caller has already moved its model on to the
   *  NEXT version, so a lazy checkpoint pointer must not be serialised for a loader of the current one */
  bool at_barrier1_ = false;
  // local models: slot j holds the blob of rank (me - j); two generations
  std::vector<std::string> local_[2];
  std::vector<char> local_has_[2];
  int local_gen_ = 0;
  int num_local_replica_ = 0;
  int use_local_model_ = -1;
  int recover_counter_ = 0;
  // bytes of recovery payloads / checkpoint replicas this rank sent or received, by transport (rabit_debug log)
  uint64_t nvlink_payload_bytes_ = 0, tcp_payload_bytes_ = 0;
  std::thread watchdog_;
  std::atomic<bool> watchdog_cancel_{false};
  std::atomic<bool> watchdog_done_{true};
  std::atomic<bool> watchdog_fired_{false};
};

// ------------------------------------------------------------------- MockEngine
class MockEngine : public RobustEngine {
 public:
  const char *Name() const override { return "mock"; }
  void Allgather(void *sendrecvbuf, size_t total_size, size_t slice_begin, size_t slice_end,
                 size_t size_prev_slice, const char *_file, const int _line,
                 const char *_caller) override;
  void AllreduceTyped(void *sendrecvbuf_, size_t type_nbytes, size_t count, ReduceFunction reducer,
                      int dtype, int op, PreprocFunction prepare_fun, void *prepare_arg,
                      const char *_file, const int _line, const char *_caller) override;
  void Broadcast(void *sendrecvbuf_, size_t size, int root, const char *_file, const int _line,
                 const char *_caller) override;
  int LoadCheckPoint(Serializable *global_model, Serializable *local_model) override;
  void CheckPoint(const Serializable *global_model, const Serializable *local_model) override;
  void LazyCheckPoint(const Serializable *global_model) override;

 private:
  /*! \brief die here if (rank, version, seq, trial) was scripted */
  void Verify(const char *what);
  double tsum_allreduce_ = 0.0, tsum_allgather_ = 0.0, time_checkpoint_ = 0.0;
};

}  // namespace engine
}  // namespace rabit
#endif  // RABIT_SRC_HOST_ENGINE_CORE_H_
