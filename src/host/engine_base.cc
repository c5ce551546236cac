/*!
 * rabit-b200: BaseEngine — collectives without fault tolerance. Any transport failure is
 * fatal here (reference AllreduceBase used as an engine: allreduce_base.h:104-160).
 */
#include <unistd.h>

#include "../device/device_backend.h"
#include "engine_core.h"

#include <dirent.h>

namespace rabit {
namespace engine {

BaseEngine::BaseEngine() {}
BaseEngine::~BaseEngine() {}

MemKind BaseEngine::Classify(const void *ptr) const {
  if (ptr == nullptr) return MemKind::kHost;
  return DeviceBackend::IsDevicePointer(ptr) ? MemKind::kDevice : MemKind::kHost;
}

bool BaseEngine::Init(int argc, char *argv[]) {
  comm_.cfg.Load(argc, argv);
  // Hadoop streaming sniffing (reference allreduce_base.cc:66-100): task ids and world size
  // come from the job environment when rabit runs inside a mapper.
  if (comm_.cfg.task_id == "NULL") {
    const char *tip = getenv("mapred_tip_id");
    if (tip == nullptr) tip = getenv("mapreduce_task_id");
    if (tip != nullptr) { comm_.cfg.task_id = tip; comm_.cfg.hadoop_mode = 1; }
  }
  if (comm_.cfg.hadoop_mode != 0) {
    const char *att = getenv("mapred_task_id");
    if (att == nullptr) att = getenv("mapreduce_task_attempt_id");
    if (att != nullptr) {
      const char *us = strrchr(att, '_');
      int n;
      if (us != nullptr && sscanf(us + 1, "%d", &n) == 1) comm_.cfg.num_trial = n;
    }
    const char *nmap = getenv("mapred_map_tasks");
    if (nmap == nullptr) nmap = getenv("mapreduce_job_maps");
    if (nmap != nullptr) comm_.cfg.world_size_hint = atoi(nmap);
  }
  if (comm_.cfg.role != "worker") {
    // servers / schedulers of a dmlc job have nothing to do in an allreduce program
    fprintf(stderr, "Rabit Module currently only work with dmlc worker, quit this program by exit 0\n");
    exit(0);
  }
  comm_.host = utils::SockAddr::GetHostName();
  // Standby worker (started by the launcher next to the live one, RABIT_STANDBY_FIFO): everything up to here - the
  // program's own start-up, library loading, and now the CUDA context - is done ahead of time; the process then
  // sleeps until the launcher hands it the life of a worker that died (one line: the attempt number) or ends the
  // job (end of file). Cuts the dominant part of recover-after-kill: starting a process and its CUDA context.
  if (const char *fifo = getenv("RABIT_STANDBY_FIFO")) {
    if (comm_.cfg.device != "off") device_.reset(DeviceBackend::Create(comm_.cfg));
    FILE *f = fopen(fifo, "r");
    int attempt = 0;
    if (f == nullptr || fscanf(f, "%d", &attempt) != 1) exit(0);   // the job ended without needing us
    fclose(f);
    comm_.cfg.num_trial = attempt;
    unsetenv("RABIT_STANDBY_FIFO");
  }
  if (!comm_.Connect("start")) {
    if (comm_.job_over) {
      // A replacement process that arrives after its peers have left: they could only leave because every rank,
      // the one we replace included, had reached the final barrier. Nothing is left to compute or to re-join.
      fprintf(stderr, "[rabit] the job finished while this worker was being restarted; nothing to re-join, exit 0\n");
      exit(0);
    }
    // Carrying on would mean running as a silent world of one next to a job that is waiting for this rank: stop
    // here, the launcher starts the worker again (the reference throws from its tracker connect as well).
    utils::Error("cannot join the job: the rendezvous through tracker %s:%d failed", comm_.cfg.tracker_uri.c_str(),
                 comm_.cfg.tracker_port);
  }
  if (comm_.world > 1 || comm_.cfg.device_force) {
    const double t_dev = utils::GetTime();
    if (!device_) device_.reset(DeviceBackend::Create(comm_.cfg));
    utils::Check(device_ != nullptr || comm_.cfg.device != "on",
                 "rabit_device=on but this rank has no usable GPU (or the library was built without CUDA)");
    if (getenv("RABIT_RECOVER_PROFILE") != nullptr) {
      utils::HandleLogInfo("[%d] init: CUDA context + backend set-up %.1f ms (attempt %d)\n", comm_.rank,
                           (utils::GetTime() - t_dev) * 1e3, comm_.cfg.num_trial);
    }
    this->AttachAfterStart();
  } else {
    device_.reset();
  }
  initialized_ = true;
  return true;
}

void BaseEngine::AttachAfterStart() {
  // (a peer that dies this early is fatal for an engine without recovery; RobustEngine re-joins instead)
  utils::Check(this->AttachDevice() != AttachResult::kLinkError, "lost a peer while setting up the device arena");
}

BaseEngine::AttachResult BaseEngine::AttachDevice() {
  // Whether the job uses the NVLink arena is ONE decision of all ranks. A rank without a usable GPU (none
  // visible, wrong architecture, rabit_device=off only there) must not skip the bootstrap while the others run
  // its host collectives: the byte streams of different collectives would be paired. So everybody takes part in
  // this agreement, with or without a backend, after every rendezvous.
  int agreed[2] = {device_ ? 1 : 0, comm_.AllLinksSharedMemory() ? 1 : 0};
  if (comm_.world > 1) {
    // (the second value rides along: do all links of the job carry their payload through shared memory? The choice
    // between tree and ring depends on it and has to be the same everywhere, HostComm::UseRing)
    comm_.shm_everywhere = false;
    const Status st = comm_.Allreduce(agreed, sizeof(int), 2, [](const void *s, void *d, int n, const MPI::Datatype &) {
      const int *src = static_cast<const int *>(s);
      int *dst = static_cast<int *>(d);
      for (int i = 0; i < n; ++i) dst[i] = std::min(dst[i], src[i]);
    });
    if (!Ok(st)) return AttachResult::kLinkError;
    comm_.shm_everywhere = agreed[1] != 0;
  }
  const int have = agreed[0];
  if (have == 0) {
    if (device_) {
      utils::Check(comm_.cfg.device != "on", "rabit_device=on but another rank has no usable GPU");
      if (comm_.rank == 0) fprintf(stderr, "[rabit] a rank has no usable GPU: device buffers are staged through the TCP collectives\n");
      device_.reset();
    }
    return AttachResult::kNoDevice;
  }
  if (device_->Attach(&comm_)) return AttachResult::kAttached;
  if (comm_.LinksBroken()) return AttachResult::kLinkError;
  // declined by all ranks alike (several hosts, no peer access ...): the same answer everywhere
  utils::Check(comm_.cfg.device != "on", "rabit_device=on but the device arena could not be set up");
  device_.reset();
  return AttachResult::kNoDevice;
}

bool BaseEngine::Shutdown(void) {
  if (device_) { device_->Detach(); device_.reset(); }
  comm_.CloseLinks(true);
  comm_.SendShutdown();
  if (comm_.cfg.debug) {
    // (every re-join opens sockets and shared-memory rings: this line is how the kill campaigns watch for leaks)
    int nfd = 0;
    if (DIR *d = opendir("/proc/self/fd")) {
      while (readdir(d) != nullptr) ++nfd;
      closedir(d);
    }
    utils::HandleLogInfo("[%d] open file descriptors after shutdown: %d\n", comm_.rank, nfd - 3);  // ".", "..", the DIR itself
  }
  initialized_ = false;
  return true;
}

namespace {
void Fatal(Status s, const char *what) {
  utils::Error("%s failed: %s (the base engine cannot recover; use the robust engine)", what, StatusName(s));
}
}  // namespace

void BaseEngine::AllreduceTyped(void *buf, size_t type_nbytes, size_t count, ReduceFunction reducer,
                                int dtype, int op, PreprocFunction prepare_fun, void *prepare_arg,
                                const char *, const int, const char *) {
  if (prepare_fun != nullptr) prepare_fun(prepare_arg);
  if (comm_.world <= 1 && !(device_ && comm_.cfg.device_force)) return;
  if (Classify(buf) == MemKind::kDevice) {
    if (device_ != nullptr && dtype >= 0 && op >= 0) {
      Status s = device_->Allreduce(buf, count, dtype, op, 1, 0, nullptr);
      if (!Ok(s)) Fatal(s, "device allreduce");
      return;
    }
    // opaque reducer, or no arena attached (several hosts, rabit_device=off): stage through the host
    std::string tmp(type_nbytes * count, '\0');
    DeviceBackend::RawCopy(&tmp[0], buf, tmp.size(), false);
    Status s = comm_.Allreduce(&tmp[0], type_nbytes, count, reducer);
    if (!Ok(s)) Fatal(s, "allreduce");
    DeviceBackend::RawCopy(buf, tmp.data(), tmp.size(), true);
    return;
  }
  Status s = comm_.Allreduce(buf, type_nbytes, count, reducer);
  if (!Ok(s)) Fatal(s, "allreduce");
  ++seq_;
}

void BaseEngine::Broadcast(void *buf, size_t size, int root, const char *, const int, const char *) {
  if (comm_.world <= 1) return;
  if (Classify(buf) == MemKind::kDevice) {
    if (device_ != nullptr) {
      Status s = device_->Broadcast(buf, size, root, 0, nullptr);
      if (!Ok(s)) Fatal(s, "device broadcast");
      return;
    }
    std::string tmp(size, '\0');
    if (comm_.rank == root) DeviceBackend::RawCopy(&tmp[0], buf, size, false);
    Status s = comm_.Broadcast(&tmp[0], size, root);
    if (!Ok(s)) Fatal(s, "broadcast");
    if (comm_.rank != root) DeviceBackend::RawCopy(buf, tmp.data(), size, true);
    return;
  }
  Status s = comm_.Broadcast(buf, size, root);
  if (!Ok(s)) Fatal(s, "broadcast");
  ++seq_;
}

void BaseEngine::Allgather(void *buf, size_t total_size, size_t slice_begin, size_t slice_end,
                           size_t size_prev_slice, const char *, const int, const char *) {
  if (comm_.world <= 1) return;
  if (Classify(buf) == MemKind::kDevice) {
    if (device_ != nullptr) {
      Status s = device_->Allgather(buf, total_size, slice_begin, slice_end, 0, nullptr);
      if (!Ok(s)) Fatal(s, "device allgather");
      return;
    }
    std::string tmp(total_size, '\0');
    DeviceBackend::RawCopy(&tmp[slice_begin], static_cast<char *>(buf) + slice_begin, slice_end - slice_begin, false);
    Status s = comm_.AllgatherRing(&tmp[0], total_size, slice_begin, slice_end, size_prev_slice);
    if (!Ok(s)) Fatal(s, "allgather");
    DeviceBackend::RawCopy(buf, tmp.data(), total_size, true);
    return;
  }
  Status s = comm_.AllgatherRing(buf, total_size, slice_begin, slice_end, size_prev_slice);
  if (!Ok(s)) Fatal(s, "allgather");
  ++seq_;
}

}  // namespace engine
}  // namespace rabit
