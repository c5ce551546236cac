/*!
 * rabit-b200: single-process unit tests (no tracker, world of one), own 40-line harness.
 * Covers what the reference's gtest suite covers (test/cpp/allreduce_base_test.cc: parameter
 * parsing; allreduce_mock_test.cc: scripted failures throw; allreduce_robust_test.cc: the
 * recovery watchdog; test_io.cc, test_utils.cc) plus the consensus decision table, the
 * result store, streams and the tracker-free code paths of this build.
 */
#include <rabit/rabit.h>

#include <chrono>
#include <cstdio>
#include <functional>
#include <string>
#include <thread>
#include <vector>

#include "../src/host/engine_core.h"
#include "../src/host/shm_ring.h"

using namespace rabit;          // NOLINT(*)
using namespace rabit::engine;  // NOLINT(*)

namespace {
struct Case { const char *name; std::function<void()> fn; };
std::vector<Case> &Cases() { static std::vector<Case> c; return c; }
struct Reg { Reg(const char *n, std::function<void()> f) { Cases().push_back(Case{n, f}); } };
#define TEST(name) void name(); Reg reg_##name(#name, name); void name()
#define EXPECT(...) do { if (!(__VA_ARGS__)) { throw std::runtime_error(std::string("EXPECT failed: ") + #__VA_ARGS__ + \
  " at line " + std::to_string(__LINE__)); } } while (0)
template <typename F> bool Throws(F f) { try { f(); } catch (const rabit::Error &) { return true; } return false; }

std::vector<char *> Argv(std::vector<std::string> &v) {  // NOLINT(*)
  std::vector<char *> out;
  for (auto &s : v) out.push_back(&s[0]);
  return out;
}
}  // namespace

// ---- parameters (parity: allreduce_base_test.cc) ----
TEST(config_defaults_and_argv) {
  Config c;
  EXPECT(c.tracker_uri == "NULL" && c.tracker_port == 9000 && c.reduce_ring_mincount == (256u << 10));
  std::vector<std::string> a = {"prog", "rabit_task_id=7", "rabit_bootstrap_cache=1", "rabit_debug=true",
                                "rabit_reduce_ring_mincount=10", "rabit_reduce_buffer=16K", "mock=1,2,3,4",
                                "mock=0,0,1,0", "rabit_timeout=1", "rabit_timeout_sec=7", "notaparam", "=x"};
  auto argv = Argv(a);
  c.Load(static_cast<int>(argv.size()), argv.data());
  EXPECT(c.task_id == "7" && c.bootstrap_cache && c.debug && c.reduce_ring_mincount == 10);
  EXPECT(c.reduce_buffer == (16u << 10) && c.mock_keys.size() == 2 && c.timeout_enabled && c.timeout_sec == 7);
  EXPECT(c.mock_keys.count(MockKey(1, 2, 3, 4)) == 1);
}
TEST(config_bad_values_throw) {
  Config c;
  EXPECT(Throws([&] { c.Set("mock", "1,2,3"); }));
  EXPECT(Throws([&] { c.Set("rabit_reduce_buffer", "12Q"); }));
  EXPECT(Throws([&] { c.Set("rabit_timeout_sec", "-1"); }));
  EXPECT(Config::ParseBytes("x", "2G") == (2ull << 30) && Config::ParseBytes("x", "3MB") == (3u << 20));
}
TEST(config_env_is_overridden_by_argv) {
  setenv("rabit_debug", "1", 1);
  setenv("DMLC_TASK_ID", "envtask", 1);
  Config c;
  std::vector<std::string> a = {"prog", "rabit_task_id=argtask"};
  auto argv = Argv(a);
  c.Load(static_cast<int>(argv.size()), argv.data());
  EXPECT(c.debug && c.task_id == "argtask");
  unsetenv("rabit_debug"); unsetenv("DMLC_TASK_ID");
}

// ---- utils / io (parity: test_utils.cc, test_io.cc) ----
TEST(assert_check_error_throw) {
  EXPECT(Throws([] { utils::Assert(false, "boom %d", 1); }));
  EXPECT(Throws([] { utils::Check(false, "boom"); }));
  EXPECT(Throws([] { utils::Error("boom"); }));
  EXPECT(!Throws([] { utils::Assert(true, "fine"); }));
  EXPECT(utils::StringToBool("true") && utils::StringToBool("TRUE") && utils::StringToBool("3") &&
         !utils::StringToBool("false") && !utils::StringToBool("0"));
}
TEST(fix_size_buffer_seek_end) {
  char mem[16] = {0};
  utils::MemoryFixSizeBuffer fs(mem, sizeof(mem));
  fs.Seek(utils::kSeekEnd);
  EXPECT(fs.Tell() == 16 && fs.AtEnd());
  fs.Seek(4);
  int v = 0x11223344;
  fs.Write(&v, sizeof(v));
  fs.Seek(4);
  int w = 0;
  EXPECT(fs.Read(&w, sizeof(w)) == sizeof(w) && w == v);
  fs.Seek(14);
  EXPECT(Throws([&] { fs.Write(&v, sizeof(v)); }));
  EXPECT(fs.Read(&w, 8) == 2);
}
TEST(buffer_stream_roundtrip_pod_encoding) {
  std::string store;
  utils::MemoryBufferStream fo(&store);
  std::vector<float> v = {1.5f, 2.5f, -3.0f};
  std::string s = "hello";
  fo.Write(v); fo.Write(s); fo.Write(7);
  // dmlc-compatible layout: u64 count + raw payload
  EXPECT(store.size() == 8 + 12 + 8 + 5 + 4);
  uint64_t n; std::memcpy(&n, store.data(), 8);
  EXPECT(n == 3);
  utils::MemoryBufferStream fi(&store);
  std::vector<float> v2; std::string s2; int k = 0;
  EXPECT(fi.Read(&v2) && fi.Read(&s2) && fi.Read(&k));
  EXPECT(v2 == v && s2 == s && k == 7 && fi.AtEnd());
  EXPECT(BeginPtr(std::vector<int>()) == nullptr);
}

// ---- world-of-one behaviour (no Init / tracker "NULL") ----
static int g_prepared = 0;
TEST(before_init_is_world_of_one) {
  EXPECT(rabit::GetRank() == 0 && rabit::GetWorldSize() == 1 && !rabit::IsDistributed());
  float x = 3.0f;
  g_prepared = 0;
  rabit::Allreduce<op::Sum>(&x, 1, [](void *) { ++g_prepared; }, nullptr);
  EXPECT(g_prepared == 1 && x == 3.0f);
  std::string s = "abc";
  rabit::Broadcast(&s, 0);
  EXPECT(s == "abc");
}
struct IntModel : public Serializable {
  int v = 0;
  void Load(Stream *fi) override { fi->Read(&v, sizeof(v)); }
  void Save(Stream *fo) const override { fo->Write(&v, sizeof(v)); }
};
static void RunSingle(const char *engine) {
  std::vector<std::string> a = {"prog", std::string("rabit_engine=") + engine};
  auto argv = Argv(a);
  EXPECT(rabit::Init(static_cast<int>(argv.size()), argv.data()));
  IntModel m;
  EXPECT(rabit::LoadCheckPoint(&m) == 0);
  int buf[3] = {1, 2, 3};
  rabit::Allreduce<op::Max>(buf, 3);
  rabit::Allreduce<op::BitOR>(buf, 3);
  EXPECT(buf[0] == 1 && buf[2] == 3);
  std::vector<int> all = {5, 6};
  rabit::Allgather(all.data(), 2, 0, 2, 2);
  rabit::CheckPoint(&m);
  EXPECT(rabit::VersionNumber() == 1);
  rabit::LazyCheckPoint(&m);
  EXPECT(rabit::VersionNumber() == 2);
  EXPECT(rabit::Finalize());
}
TEST(single_process_all_engines) {
  RunSingle("robust"); RunSingle("mock"); RunSingle("base"); RunSingle("empty");
  std::vector<std::string> a = {"prog", "rabit_engine=nonsense"};
  auto argv = Argv(a);
  EXPECT(Throws([&] { rabit::Init(2, argv.data()); }));
}

// ---- custom reducers ----
struct Pair { double g, h; };
static void PairSum(Pair &d, const Pair &s) { d.g += s.g; d.h += s.h; }  // NOLINT(*)
struct Blob {
  std::vector<int> items;
  void Load(Stream &fi) { fi.Read(&items); }    // NOLINT(*)
  void Save(Stream &fo) const { fo.Write(items); }  // NOLINT(*)
  void Reduce(const Blob &o, size_t) { items.insert(items.end(), o.items.begin(), o.items.end()); }
};
TEST(reducers_compile_and_run_single) {
  Reducer<Pair, PairSum> red;
  red.DeclarePackedSum<double>(2);
  Pair p[2] = {{1, 2}, {3, 4}};
  red.Allreduce(p, 2);
  EXPECT(p[1].h == 4);
  SerializeReducer<Blob> sred;
  Blob b; b.items = {1, 2, 3};
  sred.Allreduce(&b, 64, 1);
  EXPECT(b.items.size() == 3 && b.items[2] == 3);
  // the host reduce kernels themselves
  Pair a[2] = {{1, 1}, {2, 2}}, c[2] = {{10, 20}, {30, 40}};
  detail::StructReduceUnaligned<Pair, PairSum>(c, a, 2, MPI::Datatype(sizeof(Pair)));
  EXPECT(a[0].g == 11 && a[1].h == 42);
  float x[3] = {1, 5, 3}, y[3] = {4, 2, 3};
  op::Reducer<op::Max, float>(y, x, 3, MPI::Datatype(4));
  EXPECT(x[0] == 4 && x[1] == 5 && x[2] == 3);
  unsigned u[2] = {1, 2}, w[2] = {4, 2};
  op::Reducer<op::BitOR, unsigned>(w, u, 2, MPI::Datatype(4));
  EXPECT(u[0] == 5 && u[1] == 2);
}

// ---- mock (parity: allreduce_mock_test.cc) ----
TEST(mock_key_throws_at_the_scripted_op) {
  std::vector<std::string> a = {"prog", "rabit_engine=mock", "mock=0,0,0,0"};
  auto argv = Argv(a);
  EXPECT(rabit::Init(static_cast<int>(argv.size()), argv.data()));
  int v = 1;
  EXPECT(Throws([&] { rabit::Allreduce<op::Sum>(&v, 1); }));
  // ntrial was bumped by the hit, so the same op goes through now
  EXPECT(!Throws([&] { rabit::Allreduce<op::Sum>(&v, 1); }));
  rabit::Finalize();
  std::vector<std::string> b = {"prog", "rabit_engine=mock", "mock=0,0,0,0"};
  auto argv2 = Argv(b);
  EXPECT(rabit::Init(static_cast<int>(argv2.size()), argv2.data()));
  std::string s = "x";
  EXPECT(Throws([&] { rabit::Broadcast(&s, 0); }));
  rabit::Finalize();
}

// ---- result store ----
TEST(result_store_semantics) {
  ResultStore rs;
  char *t = rs.AllocTemp(4); std::memcpy(t, "abcd", 4); rs.Commit(0);
  t = rs.AllocTemp(2); std::memcpy(t, "xy", 2); rs.Commit(1);
  EXPECT(rs.Count() == 2 && rs.LastSeq() == 1);
  EXPECT(rs.Find(0) != nullptr && rs.Find(0)->bytes == 4 && std::string(rs.HostView(rs.Find(0)), 4) == "abcd");
  EXPECT(rs.Find(5) == nullptr);
  rs.DropLast();
  EXPECT(rs.LastSeq() == 0 && rs.Find(1) == nullptr);
  rs.AllocTemp(8); rs.DiscardTemp();
  rs.Clear();
  EXPECT(rs.Count() == 0 && rs.LastSeq() == -1);
}

TEST(host_bytes_blocks_are_recycled) {
  // medium (slab) and large blocks come back from the pool still mapped; small ones are plain malloc
  for (size_t n : {size_t(100) << 10, size_t(5) << 20}) {
    HostBytes a; a.resize(n);
    EXPECT(a.size() == n && a.data() != nullptr);
    std::memset(a.data(), 7, n);
    char *first = a.data();
    a.clear();
    EXPECT(a.size() == 0 && a.data() == nullptr);
    HostBytes b; b.resize(n - 100);   // close enough in size: same block again
    EXPECT(b.data() == first && b.size() == n - 100);
    HostBytes c(std::move(b));
    EXPECT(c.data() == first && b.data() == nullptr);
    HostBytes d; d.resize(n);         // the first block is taken: a different one
    EXPECT(d.data() != first);
  }
  HostBytes s; s.resize(10); s.resize(4); EXPECT(s.size() == 4);
  s.resize(0); EXPECT(s.size() == 0);
}

// ---- consensus decision table ----
using namespace rabit::engine::proto;  // NOLINT(*)
static Summary Fold(const std::vector<Request> &rs) {
  Summary s = Summary::Of(rs[0]);
  for (size_t i = 1; i < rs.size(); ++i) { Summary o = Summary::Of(rs[i]); SummaryReduce(&o, &s, 1, MPI::Datatype(sizeof(s))); }
  return s;
}
TEST(decide_plain_ops) {
  std::vector<Request> same = {{3, 0, 0}, {3, 0, 0}, {3, 0, 0}};
  EXPECT(Decide(Fold(same), same[0]).action == kExecute);
  std::vector<Request> lag = {{5, 0, 0}, {2, 0, 0}, {5, 0, 0}};
  Summary s = Fold(lag);
  EXPECT(Decide(s, lag[1]).action == kServeResult && Decide(s, lag[1]).requester && s.min_seq == 2);
  EXPECT(Decide(s, lag[0]).action == kServeResult && !Decide(s, lag[0]).requester);
}
TEST(decide_commit_barriers) {
  Request ck{kSeqSpecial, kCheckPoint, 0}, ack{kSeqSpecial, kCheckAck, 0}, op3{3, 0, 0}, op0{0, 0, 0};
  EXPECT(Decide(Fold({ck, ck}), ck).action == kDone);
  // a laggard keeps the barrier closed and gets served
  Summary s = Fold({ck, op3});
  EXPECT(Decide(s, ck).action == kServeResult && !Decide(s, ck).requester && Decide(s, op3).requester);
  // half the job passed barrier 1: stragglers at barrier 1 go first, ackers wait
  s = Fold({ck, ack});
  EXPECT(Decide(s, ck).action == kDone && Decide(s, ack).action == kRetry);
  // someone already in the next version: ackers are released, the runner waits
  s = Fold({ack, op0});
  EXPECT(Decide(s, ack).action == kDone && Decide(s, op0).action == kRetry);
}
TEST(decide_load_checkpoint) {
  Request ld{kSeqSpecial, kLoadCheck, 0}, op3{3, 0, 0}, ck{kSeqSpecial, kCheckPoint, 0}, ack{kSeqSpecial, kCheckAck, 0};
  EXPECT(Decide(Fold({ld, ld, ld}), ld).action == kExecute);  // fresh start
  Summary s = Fold({ld, op3, op3});
  EXPECT(Decide(s, ld).action == kServeLoad && Decide(s, ld).requester && !Decide(s, op3).requester);
  // all survivors at barrier 1: commit first, loader waits and will see the new version
  s = Fold({ld, ck, ck});
  EXPECT(Decide(s, ck).action == kDone && Decide(s, ld).action == kRetry);
  // mixed survivors: serve the loader now
  s = Fold({ld, ck, op3});
  EXPECT(Decide(s, ld).action == kServeLoad);
  // ack phase: load is served with the new version before ackers leave
  s = Fold({ld, ack});
  EXPECT(Decide(s, ack).action == kServeLoad && !Decide(s, ack).requester && Decide(s, ld).requester);
  // the loader's own acknowledgement releases the loader only: a rank at commit barrier 1 must NOT slip through
  // on it while a third rank is still inside the version's last collective
  Request lack{kSeqSpecial, kLoadAck, 0};
  s = Fold({lack, ck, op3});
  EXPECT(Decide(s, lack).action == kDone && Decide(s, ck).action == kRetry && Decide(s, op3).action == kRetry);
  s = Fold({lack, ck, ck});
  EXPECT(Decide(s, lack).action == kDone && Decide(s, ck).action == kRetry);
  // loaded the new version from ranks parked at barrier 2: everybody who acknowledged leaves in the same round
  // (a loader released alone could reach the NEXT commit while the others still wait at barrier 2 of this one)
  s = Fold({lack, ack, ack});
  EXPECT(Decide(s, lack).action == kDone && Decide(s, ack).action == kDone);
  s = Fold({lack, ack, ck});    // a knocker of the same commit goes first
  EXPECT(Decide(s, ck).action == kDone && Decide(s, lack).action == kRetry && Decide(s, ack).action == kRetry);
  s = Fold({lack, ld, op3});    // a second rank is still loading: it is served before anybody leaves
  EXPECT(Decide(s, ld).action == kServeLoad && Decide(s, ld).requester && !Decide(s, lack).requester);
}
TEST(decide_bootstrap_cache) {
  Request b0{0, kLoadCache, 0}, b1{0, kLoadCache, 1}, op2{2, 0, 1}, ld{kSeqSpecial, kLoadCheck, 1};
  EXPECT(Decide(Fold({b0, b0}), b0).action == kExecute);
  Summary s = Fold({b0, op2});
  EXPECT(Decide(s, b0).action == kServeCache && Decide(s, b0).requester && !Decide(s, op2).requester);
  s = Fold({b0, b1});  // one rank failed the tail of the bootstrap op
  EXPECT(Decide(s, b0).requester && !Decide(s, b1).requester && Decide(s, b1).action == kServeCache);
  s = Fold({b0, ld});  // the cache is synced before any checkpoint is loaded
  EXPECT(Decide(s, ld).action == kServeCache && !Decide(s, ld).requester);
}
TEST(decide_local_checkpoint_order) {
  Request lc{kSeqLocalCheckPoint, 0, 0}, la{kSeqLocalAck, 0, 0}, ck{kSeqSpecial, kCheckPoint, 0}, op1{1, 0, 0};
  EXPECT(Decide(Fold({lc, lc}), lc).action == kExecute);
  EXPECT(Decide(Fold({la, la}), la).action == kDone);
  Summary s = Fold({lc, la});
  EXPECT(Decide(s, lc).action == kServeResult && Decide(s, lc).requester && s.min_seq == kSeqLocalCheckPoint);
  s = Fold({la, ck});
  EXPECT(Decide(s, la).requester && s.min_seq == kSeqLocalAck);
  s = Fold({op1, lc});
  EXPECT(Decide(s, op1).requester && s.min_seq == 1);
  Request rq{3, kCheckAck, 0};
  EXPECT(rq.Code() == ((3u << 5) | 4u));
}

// ---- watchdog (parity: allreduce_robust_test.cc) ----
static std::atomic<int> g_timeouts{0};
static void OnTimeout(int, int) { ++g_timeouts; }
TEST(watchdog_idle_when_disabled) {
  RobustEngine e;
  e.on_timeout = OnTimeout;
  EXPECT(e.HandleStatus(Status::kOk));
  EXPECT(!e.watchdog_running() && !e.watchdog_fired());
}
TEST(watchdog_fires_then_reset_by_success) {
  // world of one: Connect() is a no-op, so HandleStatus(error) returns immediately after arming
  g_timeouts = 0;
  RobustEngine e;
  e.on_timeout = OnTimeout;
  e.SetParam("rabit_timeout", "true");
  e.SetParam("rabit_timeout_sec", "1");
  EXPECT(!e.HandleStatus(Status::kPeerClosed));
  EXPECT(e.watchdog_running() && e.recover_count() == 1);
  EXPECT(!e.HandleStatus(Status::kSockError));  // a second failure does not start a second watchdog
  EXPECT(e.recover_count() == 2);
  std::this_thread::sleep_for(std::chrono::milliseconds(1400));
  EXPECT(e.watchdog_fired() && g_timeouts == 1);
  // success cancels/joins it; a later failure may arm a new one
  EXPECT(e.HandleStatus(Status::kOk));
  EXPECT(!e.watchdog_running());
  EXPECT(!e.HandleStatus(Status::kPeerClosed));
  EXPECT(e.watchdog_running());
  EXPECT(e.HandleStatus(Status::kOk));
  EXPECT(!e.watchdog_running() && g_timeouts == 1);
}
TEST(watchdog_cancelled_by_quick_recovery) {
  g_timeouts = 0;
  RobustEngine e;
  e.on_timeout = OnTimeout;
  e.SetParam("rabit_timeout", "1");
  e.SetParam("rabit_timeout_sec", "5");
  EXPECT(!e.HandleStatus(Status::kPeerClosed));
  EXPECT(e.HandleStatus(Status::kOk));
  std::this_thread::sleep_for(std::chrono::milliseconds(300));
  EXPECT(!e.watchdog_fired() && g_timeouts == 0);
}

// ---- shared-memory ring of same-host links (src/host/shm_ring.h) ----
TEST(shm_ring_streams_bytes_across_the_wrap) {
  ShmRing tx, rx;
  EXPECT(tx.Create(64 << 10));
  EXPECT(rx.Open(static_cast<int>(getpid()), tx.fd(), 64 << 10));   // the consumer's way in: /proc/<pid>/fd/<n>
  EXPECT(!rx.Open(static_cast<int>(getpid()), tx.fd(), 128 << 10)); // a size that does not match is refused
  EXPECT(rx.Open(static_cast<int>(getpid()), tx.fd(), 64 << 10));
  EXPECT(tx.Space() == (64 << 10) && rx.Avail() == 0);
  const size_t total = 3000000;    // ~46 laps
  std::thread producer([&]() {
    std::vector<char> piece(40000);
    size_t sent = 0;
    unsigned seed = 12345;
    while (sent < total) {
      seed = seed * 1103515245u + 12345u;
      size_t n = std::min<size_t>(1 + (seed >> 8) % piece.size(), total - sent);
      for (size_t i = 0; i < n; ++i) piece[i] = static_cast<char>((sent + i) * 31 % 251);
      size_t off = 0;
      while (off < n) off += tx.Write(piece.data() + off, n - off);
      sent += n;
    }
  });
  std::vector<char> out(7777);
  size_t got = 0;
  bool good = true;
  unsigned seed = 999;
  while (got < total) {
    seed = seed * 1103515245u + 12345u;
    if ((seed >> 10) % 3 == 0) {
      // zero-copy path: what lies contiguously in the mapping
      const char *span;
      size_t n = std::min(rx.Peek(&span), total - got);
      for (size_t i = 0; i < n; ++i) good = good && span[i] == static_cast<char>((got + i) * 31 % 251);
      rx.Consume(n);
      got += n;
    } else {
      size_t n = rx.Read(out.data(), std::min(out.size(), total - got));
      for (size_t i = 0; i < n; ++i) good = good && out[i] == static_cast<char>((got + i) * 31 % 251);
      got += n;
    }
  }
  producer.join();
  EXPECT(good);
  EXPECT(rx.Avail() == 0 && tx.Space() == (64 << 10));
  // the sleep / wake notes
  rx.SetConsumerWaiting(1);
  EXPECT(tx.TakeConsumerWaiting());
  EXPECT(!tx.TakeConsumerWaiting());
  tx.SetProducerWaiting(1);
  EXPECT(rx.TakeProducerWaiting() && !rx.TakeProducerWaiting());
  uint64_t a[2], b[2];
  HostToken(a); HostToken(b);
  EXPECT(a[0] == b[0] && a[1] == b[1] && (a[0] != 0 || a[1] != 0));
  tx.Close(); rx.Close();
  EXPECT(!tx.Active() && !rx.Active());
}

int main(int argc, char *argv[]) {
  int failed = 0;
  for (const Case &c : Cases()) {
    if (argc > 1 && std::string(c.name).find(argv[1]) == std::string::npos) continue;
    try {
      c.fn();
      std::printf("[ OK ] %s\n", c.name);
    } catch (const std::exception &e) {
      ++failed;
      std::printf("[FAIL] %s: %s\n", c.name, e.what());
    }
    std::fflush(stdout);
  }
  std::printf("%d failed\n", failed);
  return failed == 0 ? 0 : 1;
}
