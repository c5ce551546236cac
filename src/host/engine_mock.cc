/*!
 * rabit-b200: MockEngine — scripted failures for tests and timing statistics.
 * `mock=rank,version,seqno,ntrial` kills the matching rank at the entry of the matching
 * collective or checkpoint, exactly where the reference does (allreduce_mock.h:47-137,
 * 196-202). "Kill" = throw rabit::Error; an uncaught one aborts the process and the
 * launcher restarts it with DMLC_NUM_ATTEMPT + 1, so the same key does not fire twice.
 */
#include <sstream>

#include "engine_core.h"

namespace rabit {
namespace engine {

namespace {
/*! \brief routes a global model through the local-model path (force_local=1) */
struct NullModel : public Serializable {
  void Load(Stream *) override {}
  void Save(Stream *) const override {}
};
struct PairModel : public Serializable {
  Serializable *a, *b;
  const Serializable *ca, *cb;
  PairModel(Serializable *a_, Serializable *b_) : a(a_), b(b_), ca(a_), cb(b_) {}
  PairModel(const Serializable *a_, const Serializable *b_) : a(nullptr), b(nullptr), ca(a_), cb(b_) {}
  void Load(Stream *fi) override {
    if (a != nullptr) a->Load(fi);
    if (b != nullptr) b->Load(fi);
  }
  void Save(Stream *fo) const override {
    if (ca != nullptr) ca->Save(fo);
    if (cb != nullptr) cb->Save(fo);
  }
};
}  // namespace

void MockEngine::Verify(const char *what) {
  Config &cfg = comm_.cfg;
  if (cfg.mock_keys.count(MockKey(comm_.rank, version_number_, seq_, cfg.num_trial)) == 0) return;
  cfg.num_trial += 1;
  fflush(stdout);
  utils::Error("[%d]@@@Hit Mock Error:%s ", comm_.rank, what);
}

void MockEngine::AllreduceTyped(void *buf, size_t type_nbytes, size_t count, ReduceFunction reducer,
                                int dtype, int op, PreprocFunction prepare_fun, void *prepare_arg,
                                const char *_file, const int _line, const char *_caller) {
  Verify("AllReduce");
  const double t0 = utils::GetTime();
  RobustEngine::AllreduceTyped(buf, type_nbytes, count, reducer, dtype, op, prepare_fun, prepare_arg,
                               _file, _line, _caller);
  tsum_allreduce_ += utils::GetTime() - t0;
}

void MockEngine::Allgather(void *buf, size_t total_size, size_t slice_begin, size_t slice_end,
                           size_t size_prev_slice, const char *_file, const int _line,
                           const char *_caller) {
  Verify("Allgather");
  const double t0 = utils::GetTime();
  RobustEngine::Allgather(buf, total_size, slice_begin, slice_end, size_prev_slice, _file, _line, _caller);
  tsum_allgather_ += utils::GetTime() - t0;
}

void MockEngine::Broadcast(void *buf, size_t size, int root, const char *_file, const int _line,
                           const char *_caller) {
  Verify("Broadcast");
  RobustEngine::Broadcast(buf, size, root, _file, _line, _caller);
}

int MockEngine::LoadCheckPoint(Serializable *global_model, Serializable *local_model) {
  tsum_allreduce_ = tsum_allgather_ = 0.0;
  time_checkpoint_ = utils::GetTime();
  if (comm_.cfg.force_local == 0) return RobustEngine::LoadCheckPoint(global_model, local_model);
  NullModel none;
  PairModel both(global_model, local_model);
  return RobustEngine::LoadCheckPoint(&none, &both);
}

void MockEngine::CheckPoint(const Serializable *global_model, const Serializable *local_model) {
  Verify("CheckPoint");
  const double t0 = utils::GetTime();
  const double since_last = t0 - time_checkpoint_;
  if (comm_.cfg.force_local == 0) {
    RobustEngine::CheckPoint(global_model, local_model);
  } else {
    NullModel none;
    PairModel both(global_model, local_model);
    RobustEngine::CheckPoint(&none, &both);
  }
  time_checkpoint_ = utils::GetTime();
  if (comm_.cfg.report_stats != 0 && comm_.rank == 0) {
    size_t local_bytes = 0;
    for (int g = 0; g < 2; ++g) for (const std::string &b : local_[g]) local_bytes += b.size();
    std::ostringstream os;
    os << "[v" << version_number_ << "] global_size=" << global_blob_.size()
       << ",local_size=" << local_bytes
       << ",check_tcost=" << (time_checkpoint_ - t0) << " sec"
       << ",allreduce_tcost=" << tsum_allreduce_ << " sec"
       << ",allgather_tcost=" << tsum_allgather_ << " sec"
       << ",between_chpt=" << since_last << "sec\n";
    this->TrackerPrint(os.str());
  }
  tsum_allreduce_ = tsum_allgather_ = 0.0;
}

void MockEngine::LazyCheckPoint(const Serializable *global_model) {
  Verify("LazyCheckPoint");
  RobustEngine::LazyCheckPoint(global_model);
}

}  // namespace engine
}  // namespace rabit
