/*!
 * rabit-b200: C ABI over the C++ API (parity: reference src/c_api.cc:221-333).
 * Deliberate differences, see SURVEY.md 2.7: checkpoint blobs carry their full 64-bit
 * length (the reference truncates it to 16 bits, #1); RabitAllgather takes ELEMENT counts
 * and scales once (#2); all ten dtypes are dispatched (the reference stops at double).
 */
#include "rabit/c_api.h"

#include <string>

#include "../device/device_backend.h"
#include "engine_core.h"
#include "rabit/rabit.h"

namespace {
using rabit::engine::mpi::DataType;
using rabit::engine::mpi::OpType;

thread_local std::string g_last_error;
thread_local std::string g_ckpt_global, g_ckpt_local;

/*! \brief run fn, remember the message of anything it throws, and rethrow */
template <typename Fn>
auto Guard(Fn fn) -> decltype(fn()) {
  try {
    return fn();
  } catch (const std::exception &e) {
    g_last_error = e.what();
    throw;
  }
}

template <typename OP, typename T>
void ReduceAs(void *buf, size_t count, void (*prep)(void *), void *arg) {
  rabit::Allreduce<OP>(static_cast<T *>(buf), count, prep, arg);
}
template <typename OP>
void ReduceArith(void *buf, size_t count, int dtype, void (*prep)(void *), void *arg) {
  switch (dtype) {
    case rabit::engine::mpi::kChar: return ReduceAs<OP, char>(buf, count, prep, arg);
    case rabit::engine::mpi::kUChar: return ReduceAs<OP, unsigned char>(buf, count, prep, arg);
    case rabit::engine::mpi::kInt: return ReduceAs<OP, int>(buf, count, prep, arg);
    case rabit::engine::mpi::kUInt: return ReduceAs<OP, unsigned>(buf, count, prep, arg);
    case rabit::engine::mpi::kLong: return ReduceAs<OP, long>(buf, count, prep, arg);                     // NOLINT(*)
    case rabit::engine::mpi::kULong: return ReduceAs<OP, unsigned long>(buf, count, prep, arg);           // NOLINT(*)
    case rabit::engine::mpi::kFloat: return ReduceAs<OP, float>(buf, count, prep, arg);
    case rabit::engine::mpi::kDouble: return ReduceAs<OP, double>(buf, count, prep, arg);
    case rabit::engine::mpi::kLongLong: return ReduceAs<OP, long long>(buf, count, prep, arg);            // NOLINT(*)
    case rabit::engine::mpi::kULongLong: return ReduceAs<OP, unsigned long long>(buf, count, prep, arg);  // NOLINT(*)
    default: rabit::utils::Error("unknown data_type %d", dtype);
  }
}
void ReduceBits(void *buf, size_t count, int dtype, void (*prep)(void *), void *arg) {
  typedef rabit::op::BitOR OP;
  switch (dtype) {
    case rabit::engine::mpi::kChar: return ReduceAs<OP, char>(buf, count, prep, arg);
    case rabit::engine::mpi::kUChar: return ReduceAs<OP, unsigned char>(buf, count, prep, arg);
    case rabit::engine::mpi::kInt: return ReduceAs<OP, int>(buf, count, prep, arg);
    case rabit::engine::mpi::kUInt: return ReduceAs<OP, unsigned>(buf, count, prep, arg);
    case rabit::engine::mpi::kLong: return ReduceAs<OP, long>(buf, count, prep, arg);                     // NOLINT(*)
    case rabit::engine::mpi::kULong: return ReduceAs<OP, unsigned long>(buf, count, prep, arg);           // NOLINT(*)
    case rabit::engine::mpi::kLongLong: return ReduceAs<OP, long long>(buf, count, prep, arg);            // NOLINT(*)
    case rabit::engine::mpi::kULongLong: return ReduceAs<OP, unsigned long long>(buf, count, prep, arg);  // NOLINT(*)
    case rabit::engine::mpi::kFloat: case rabit::engine::mpi::kDouble:
      rabit::utils::Error("DataType does not support bitwise or operation");
    default: rabit::utils::Error("unknown data_type %d", dtype);
  }
}

/*! \brief opaque blob as a Serializable: u64 length + bytes */
struct BlobOut : public rabit::Serializable {
  const char *data; size_t len;
  BlobOut(const char *d, size_t n) : data(d), len(n) {}
  void Load(rabit::Stream *) override { rabit::utils::Error("BlobOut is write-only"); }
  void Save(rabit::Stream *fo) const override {
    uint64_t n = len;
    fo->Write(&n, sizeof(n));
    if (n != 0) fo->Write(data, len);
  }
};
struct BlobIn : public rabit::Serializable {
  std::string *dst;
  explicit BlobIn(std::string *d) : dst(d) {}
  void Load(rabit::Stream *fi) override {
    uint64_t n = 0;
    rabit::utils::Assert(fi->Read(&n, sizeof(n)) == sizeof(n), "checkpoint blob: missing length");
    dst->resize(n);
    if (n != 0) rabit::utils::Assert(fi->Read(&(*dst)[0], n) == n, "checkpoint blob: truncated");
  }
  void Save(rabit::Stream *) const override { rabit::utils::Error("BlobIn is read-only"); }
};
}  // namespace

bool RabitInit(int argc, char *argv[]) { return Guard([&] { return rabit::Init(argc, argv); }); }
bool RabitFinalize(void) { return Guard([&] { return rabit::Finalize(); }); }
int RabitGetRingPrevRank(void) { return rabit::GetRingPrevRank(); }
int RabitGetRank(void) { return rabit::GetRank(); }
int RabitGetWorldSize(void) { return rabit::GetWorldSize(); }
int RabitIsDistributed(void) { return rabit::IsDistributed() ? 1 : 0; }
void RabitTrackerPrint(const char *msg) { Guard([&] { rabit::TrackerPrint(std::string(msg)); return 0; }); }

void RabitGetProcessorName(char *out_name, rbt_ulong *out_len, rbt_ulong max_len) {
  std::string s = rabit::GetProcessorName();
  if (s.length() > max_len) s.resize(max_len > 0 ? max_len - 1 : 0);
  std::memcpy(out_name, s.data(), s.length());
  if (s.length() < max_len) out_name[s.length()] = '\0';
  *out_len = static_cast<rbt_ulong>(s.length());
}

void RabitBroadcast(void *sendrecv_data, rbt_ulong size, int root) {
  Guard([&] { rabit::Broadcast(sendrecv_data, size, root); return 0; });
}

void RabitAllgather(void *sendrecvbuf, size_t total_count, size_t begin_index, size_t slice_count,
                    size_t prev_slice_count, int enum_dtype) {
  Guard([&] {
    const size_t w = rabit::engine::DTypeBytes(enum_dtype);
    rabit::utils::Check(w != 0, "unknown data_type %d", enum_dtype);
    rabit::engine::GetEngine()->Allgather(sendrecvbuf, total_count * w, begin_index * w,
                                          (begin_index + slice_count) * w, prev_slice_count * w);
    return 0;
  });
}

void RabitAllreduce(void *sendrecvbuf, size_t count, int enum_dtype, int enum_op,
                    void (*prepare_fun)(void *arg), void *prepare_arg) {
  Guard([&] {
    switch (enum_op) {
      case rabit::engine::mpi::kMax: ReduceArith<rabit::op::Max>(sendrecvbuf, count, enum_dtype, prepare_fun, prepare_arg); break;
      case rabit::engine::mpi::kMin: ReduceArith<rabit::op::Min>(sendrecvbuf, count, enum_dtype, prepare_fun, prepare_arg); break;
      case rabit::engine::mpi::kSum: ReduceArith<rabit::op::Sum>(sendrecvbuf, count, enum_dtype, prepare_fun, prepare_arg); break;
      case rabit::engine::mpi::kBitwiseOR: ReduceBits(sendrecvbuf, count, enum_dtype, prepare_fun, prepare_arg); break;
      default: rabit::utils::Error("unknown enum_op %d", enum_op);
    }
    return 0;
  });
}

int RabitLoadCheckPoint(char **out_global_model, rbt_ulong *out_global_len,
                        char **out_local_model, rbt_ulong *out_local_len) {
  return Guard([&] {
    BlobIn g(&g_ckpt_global), l(&g_ckpt_local);
    int version;
    if (out_local_model == nullptr) {
      version = rabit::LoadCheckPoint(&g, nullptr);
    } else {
      version = rabit::LoadCheckPoint(&g, &l);
      *out_local_model = rabit::BeginPtr(g_ckpt_local);
      *out_local_len = static_cast<rbt_ulong>(g_ckpt_local.length());
    }
    *out_global_model = rabit::BeginPtr(g_ckpt_global);
    *out_global_len = static_cast<rbt_ulong>(g_ckpt_global.length());
    return version;
  });
}

void RabitCheckPoint(const char *global_model, rbt_ulong global_len, const char *local_model,
                     rbt_ulong local_len) {
  Guard([&] {
    BlobOut g(global_model, global_len), l(local_model, local_len);
    rabit::CheckPoint(&g, local_model == nullptr ? nullptr : &l);
    return 0;
  });
}

int RabitVersionNumber(void) { return rabit::VersionNumber(); }
int RabitLinkTag(void) { return 0; }

const char *RabitGetLastError(void) { return g_last_error.c_str(); }
int RabitDeviceAvailable(void) {
  return rabit::engine::DeviceBackend::Compiled() ? 1 : 0;
}
