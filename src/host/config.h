/*!
 * rabit-b200: flat `name=value` configuration.
 *
 * Sources in rising priority: compiled defaults -> environment (only the names in
 * kEnvNames) -> argv entries of the form name=value. Same names, defaults and parsing
 * quirks as the reference (allreduce_base.cc:18-46,150-216; allreduce_robust.cc:88-98;
 * allreduce_mock.h:32-45), tabulated in SURVEY.md section 5.6, plus the rabit_device_*
 * family that configures the sm_100a backend.
 */
#ifndef RABIT_SRC_HOST_CONFIG_H_
#define RABIT_SRC_HOST_CONFIG_H_

#include <cstdlib>
#include <cstring>
#include <set>
#include <string>
#include <tuple>
#include <vector>

#include "rabit/internal/utils.h"

namespace rabit {
namespace engine {

/*! \brief scripted failure point: (rank, version, seqno, restart count) */
typedef std::tuple<int, int, int, int> MockKey;

struct Config {
  // ---- bootstrap ----
  std::string tracker_uri = "NULL";
  int tracker_port = 9000;
  std::string task_id = "NULL";
  std::string role = "worker";
  int world_size_hint = -1;
  int connect_retry = 5;
  int slave_port = 0;          // 0 = let the OS pick; >0 = scan [port, port + nport_trial)
  int nport_trial = 1000;
  // ---- host collectives ----
  size_t tree_reduce_minsize = 64 << 10;     // bytes a tree node gathers before it folds and forwards (reference: 1 MiB;
                                             // smaller here so that the levels of the tree still overlap, doc/parameters.md)
  size_t reduce_ring_mincount = 256 << 10;   // elements; above this the ring is used (reference: 32 Ki; see doc/parameters.md)
  bool reduce_ring_mincount_set = false;     // given by the user: then it also holds when every link is shared memory
  size_t reduce_buffer = 256UL << 20;        // cap of a per-link staging ring
  size_t shm_ring = 512 << 10;               // bytes per direction of a same-host link's shared-memory ring (0 = TCP only);
                                             // measured: a ring that stays in the L2 beats a large one at every message size
  int shm_spin_us = 50;                      // look at the rings this long before sleeping on the doorbell
  bool shm_spin_set = false;                 // given by the user: then it also holds when the workers outnumber the cores
  int shm_test_split = 0;                    // test hook: ranks of different (rank % split) pretend to be on different hosts
  bool tcp_no_delay = true;                  // reference default is off (Nagle stalls)
  int hadoop_mode = 0;
  // ---- robustness ----
  bool bootstrap_cache = false;
  bool debug = false;
  bool timeout_enabled = false;
  int timeout_sec = 1800;
  int global_replica = 5;
  int local_replica = 0;
  // ---- mock ----
  int num_trial = 0;
  int report_stats = 0;
  int force_local = 0;
  std::set<MockKey> mock_keys;
  // ---- engine / device (new) ----
  std::string engine = "";                   // robust|mock|base|empty ; "" = library default
  std::string device = "auto";               // auto|on|off
  size_t device_arena = 0;                   // bytes of symmetric staging; 0 = grow on demand
  size_t device_heap = 0;                    // bytes of symmetric heap for zero-copy buffers (RabitDeviceSymAlloc)
  size_t device_oneshot_max = 0;             // bytes: <= -> one-shot kernel; 0 = table by world size
  size_t device_nvls_min = 0;                // bytes: >= -> multimem kernel when available; 0 = table by world size
  size_t device_ll_max = 0;                  // bytes: <= -> tagged-word kernel; 0 = table by world size
  size_t device_stream_min = 0;              // bytes: >= -> chunked, copy-engine fed allreduce of DEVICE buffers; 0 = table
                                             // (128 MiB with 2 ranks, off otherwise), 1 = off
  size_t device_stream_chunk = 0;            // bytes per chunk of that path; 0 = table (512 MiB)
  size_t device_stream_first = 0;            // bytes of its first chunk; 0 = table (64 MiB)
  std::string device_tune = "";              // file with measured thresholds (benchmarks/algo_sweep.py --emit-table)
  int device_blocks = 0;                     // 0 = table default
  int device_timeout_ms = 20000;             // bounded spin before a peer is declared dead
  std::string device_algo = "auto";          // auto|oneshot|twoshot|tma|nvls
  bool device_force = false;                 // run kernels even at world size 1
  size_t device_cache_bytes = 16UL << 30;    // device-resident result cache budget
  size_t device_blob_min = 32UL << 10;       // bytes: recovery payloads / checkpoint replicas >= this move by NVLink peer copies
  size_t device_host_min = 4UL << 20;        // bytes: HOST buffers >= this ride PCIe + NVLink instead of TCP; 0 = off
  size_t device_host_chunk = 64UL << 20;     // pipeline granule of that path (H2D | kernel | D2H overlap)

  /*! \brief names looked up in the environment at Init */
  static const std::vector<std::string> &EnvNames() {
    static const std::vector<std::string> names = {
      "DMLC_TASK_ID", "DMLC_ROLE", "DMLC_NUM_ATTEMPT", "DMLC_TRACKER_URI", "DMLC_TRACKER_PORT",
      "DMLC_WORKER_CONNECT_RETRY", "rabit_global_replica", "rabit_local_replica",
      "rabit_tracker_uri", "rabit_tracker_port", "rabit_task_id", "rabit_bootstrap_cache",
      "rabit_debug", "rabit_timeout", "rabit_timeout_sec", "rabit_enable_tcp_no_delay",
      "rabit_reduce_ring_mincount", "rabit_tree_reduce_minsize", "rabit_reduce_buffer",
      "rabit_engine", "rabit_device", "rabit_device_arena", "rabit_device_algo",
      "rabit_device_blocks", "rabit_device_timeout_ms", "rabit_device_force",
      "rabit_device_oneshot_max", "rabit_device_nvls_min", "rabit_device_ll_max", "rabit_device_heap", "rabit_device_cache",
      "rabit_device_stream_min", "rabit_device_stream_chunk", "rabit_device_stream_first", "rabit_device_tune",
      "rabit_device_host_min", "rabit_device_host_chunk", "rabit_device_blob_min",
      "rabit_world_size", "rabit_slave_port", "rabit_shm_ring", "rabit_shm_spin_us", "rabit_shm_test_split"};
    return names;
  }

  /*! \brief "64", "64B", "16K", "256MB", "1G" -> bytes */
  static size_t ParseBytes(const char *name, const char *val) {
    char *end = nullptr;
    unsigned long long n = strtoull(val, &end, 10);  // NOLINT(*)
    utils::Check(end != val, "invalid size for %s: \"%s\"", name, val);
    switch (*end) {
      case '\0': case 'B': case 'b': return n;
      case 'K': case 'k': return n << 10;
      case 'M': case 'm': return n << 20;
      case 'G': case 'g': return n << 30;
      default:
        utils::Error("invalid size for %s: \"%s\" (use B, K, M or G)", name, val);
    }
  }

  /*! \brief apply one pair; unknown names are ignored (programs mix their own args in) */
  void Set(const char *name, const char *val) {
    auto is = [name](const char *k) { return strcmp(name, k) == 0; };
    if (is("rabit_tracker_uri") || is("DMLC_TRACKER_URI")) tracker_uri = val;
    else if (is("rabit_tracker_port") || is("DMLC_TRACKER_PORT")) tracker_port = atoi(val);
    else if (is("rabit_task_id") || is("DMLC_TASK_ID")) task_id = val;
    else if (is("DMLC_ROLE")) role = val;
    else if (is("rabit_world_size")) world_size_hint = atoi(val);
    else if (is("rabit_hadoop_mode")) hadoop_mode = atoi(val);
    else if (is("rabit_slave_port")) slave_port = atoi(val);
    else if (is("rabit_tree_reduce_minsize")) tree_reduce_minsize = static_cast<size_t>(atol(val));
    else if (is("rabit_reduce_ring_mincount")) {
      long v = atol(val);  // NOLINT(*)
      utils::Check(v >= 0, "rabit_reduce_ring_mincount should be non-negative");
      reduce_ring_mincount = static_cast<size_t>(v);
      reduce_ring_mincount_set = true;
    }
    else if (is("rabit_reduce_buffer")) reduce_buffer = ParseBytes(name, val);
    else if (is("rabit_shm_ring")) shm_ring = ParseBytes(name, val);
    else if (is("rabit_shm_spin_us")) { shm_spin_us = atoi(val); shm_spin_set = true; }
    else if (is("rabit_shm_test_split")) shm_test_split = atoi(val);
    else if (is("DMLC_WORKER_CONNECT_RETRY")) connect_retry = atoi(val);
    else if (is("rabit_bootstrap_cache")) bootstrap_cache = utils::StringToBool(val);
    else if (is("rabit_debug")) debug = utils::StringToBool(val);
    else if (is("rabit_timeout")) timeout_enabled = utils::StringToBool(val);
    else if (is("rabit_timeout_sec")) {
      timeout_sec = atoi(val);
      utils::Check(timeout_sec >= 0, "rabit_timeout_sec should be non negative second");
    }
    else if (is("rabit_enable_tcp_no_delay")) tcp_no_delay = utils::StringToBool(val);
    else if (is("rabit_global_replica")) global_replica = atoi(val);
    else if (is("rabit_local_replica")) local_replica = atoi(val);
    else if (is("rabit_num_trial") || is("DMLC_NUM_ATTEMPT")) num_trial = atoi(val);
    else if (is("report_stats")) report_stats = atoi(val);
    else if (is("force_local")) force_local = atoi(val);
    else if (is("mock")) {
      int r, v, s, n;
      utils::Check(sscanf(val, "%d,%d,%d,%d", &r, &v, &s, &n) == 4,
                   "invalid mock parameter \"%s\", expect rank,version,seqno,ntrial", val);
      mock_keys.insert(MockKey(r, v, s, n));
    }
    else if (is("rabit_engine")) engine = val;
    else if (is("rabit_device")) device = val;
    else if (is("rabit_device_arena")) device_arena = ParseBytes(name, val);
    else if (is("rabit_device_heap")) device_heap = ParseBytes(name, val);
    else if (is("rabit_device_oneshot_max")) device_oneshot_max = ParseBytes(name, val);
    else if (is("rabit_device_ll_max")) device_ll_max = ParseBytes(name, val);
    else if (is("rabit_device_stream_min")) device_stream_min = ParseBytes(name, val);
    else if (is("rabit_device_stream_chunk")) device_stream_chunk = ParseBytes(name, val);
    else if (is("rabit_device_stream_first")) device_stream_first = ParseBytes(name, val);
    else if (is("rabit_device_tune")) device_tune = val;
    else if (is("rabit_device_nvls_min")) device_nvls_min = ParseBytes(name, val);
    else if (is("rabit_device_cache")) device_cache_bytes = ParseBytes(name, val);
    else if (is("rabit_device_host_min")) device_host_min = ParseBytes(name, val);
    else if (is("rabit_device_blob_min")) device_blob_min = ParseBytes(name, val);
    else if (is("rabit_device_host_chunk")) device_host_chunk = std::max<size_t>(ParseBytes(name, val), 64UL << 10);
    else if (is("rabit_device_blocks")) device_blocks = atoi(val);
    else if (is("rabit_device_timeout_ms")) device_timeout_ms = atoi(val);
    else if (is("rabit_device_algo")) device_algo = val;
    else if (is("rabit_device_force")) device_force = utils::StringToBool(val);
  }

  /*! \brief env first, then argv pairs */
  void Load(int argc, char *argv[]) {
    for (const std::string &n : EnvNames()) {
      const char *v = getenv(n.c_str());
      if (v != nullptr) this->Set(n.c_str(), v);
    }
    for (int i = 0; i < argc; ++i) {
      if (argv[i] == nullptr) continue;
      const char *eq = strchr(argv[i], '=');
      if (eq == nullptr || eq == argv[i]) continue;
      std::string name(argv[i], eq - argv[i]);
      this->Set(name.c_str(), eq + 1);
    }
  }
};

}  // namespace engine
}  // namespace rabit
#endif  // RABIT_SRC_HOST_CONFIG_H_
