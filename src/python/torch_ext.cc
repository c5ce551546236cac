/*!
 * rabit-b200: PyTorch extension module (pybind11 / libtorch), the native form of rabit_b200/torch_binding.py.
 *
 * The ctypes binding pays ~4-5 us of interpreter work per call (argument marshalling, stream lookup through
 * Python), which is half the cost of a small stream-ordered collective. Here a call is: read the tensor's
 * pointer / size / dtype, take torch's current CUDA stream, call the C entry point. The engine library is NOT
 * linked: rabit_b200.rabit loads one of lib/librabit*.so (which one decides the engine) and torch_binding hands
 * this module the addresses of that library's C entry points, so both bindings drive the same engine instance.
 * No counterpart in the reference (its python module is ctypes only, python/rabit.py).
 */
#include <ATen/cuda/CUDAContext.h>
#include <torch/extension.h>

#include <cstdint>
#include <stdexcept>
#include <string>

namespace {

struct Api {
  int (*allreduce_async)(void *, size_t, int, int) = nullptr;
  int (*broadcast_async)(void *, size_t, int) = nullptr;
  int (*allgather_async)(void *, size_t, size_t, size_t) = nullptr;
  int (*synchronize)(void) = nullptr;
  void (*set_stream)(void *) = nullptr;
  void (*allreduce)(void *, size_t, int, int, void (*)(void *), void *) = nullptr;
  void (*broadcast)(void *, unsigned long, int) = nullptr;  // NOLINT(*)
  unsigned long long (*launch_count)(void) = nullptr;       // NOLINT(*)
  bool ok = false;
};

Api g_api;

template <typename F>
void Take(const pybind11::dict &addr, const char *name, F *fn) {
  if (!addr.contains(name)) throw std::runtime_error(std::string("rabit torch extension: no address for ") + name);
  *fn = reinterpret_cast<F>(static_cast<uintptr_t>(addr[name].cast<unsigned long long>()));  // NOLINT(*)
}

/*!
 * \brief called by torch_binding with the addresses of the C entry points of the library that rabit.py loaded
 *  (taken through ctypes from that very handle): a process may hold several lib/librabit*.so, and a lookup in
 *  the global namespace would silently pick the first one.
 */
void BindLibrary(const pybind11::dict &addr) {
  Api api;
  Take(addr, "RabitDeviceAllreduceAsync", &api.allreduce_async);
  Take(addr, "RabitDeviceBroadcastAsync", &api.broadcast_async);
  Take(addr, "RabitDeviceAllgatherAsync", &api.allgather_async);
  Take(addr, "RabitDeviceSynchronize", &api.synchronize);
  Take(addr, "RabitDeviceSetStream", &api.set_stream);
  Take(addr, "RabitAllreduce", &api.allreduce);
  Take(addr, "RabitBroadcast", &api.broadcast);
  Take(addr, "RabitDeviceLaunchCount", &api.launch_count);
  api.ok = true;
  g_api = api;
}

Api &GetApi() {
  if (!g_api.ok) throw std::runtime_error("rabit engine library is not bound: call rabit_b200.rabit.init() first");
  return g_api;
}

/*! \brief torch dtype -> rabit::engine::mpi::DataType */
int DTypeEnum(const at::Tensor &t) {
  switch (t.scalar_type()) {
    case at::kChar: return 0;
    case at::kByte: return 1;
    case at::kInt: return 2;
    case at::kLong: return 4;
    case at::kFloat: return 6;
    case at::kDouble: return 7;
    default: throw std::invalid_argument("dtype is not supported by rabit allreduce");
  }
}

void *g_last_stream = reinterpret_cast<void *>(-1);

/*! \brief make the engine launch on torch's current stream of the tensor's device */
void BindStream(Api &api, const at::Tensor &t) {
  if (!t.is_cuda()) return;
  void *s = at::cuda::getCurrentCUDAStream(t.get_device()).stream();
  if (s != g_last_stream) {
    api.set_stream(s);
    g_last_stream = s;
  }
}

void Check(const at::Tensor &t) {
  if (!t.is_contiguous()) throw std::invalid_argument("rabit collectives need contiguous tensors");
}

at::Tensor AllreduceAsync(at::Tensor t, int op) {
  Api &api = GetApi();
  Check(t);
  if (!t.is_cuda()) throw std::invalid_argument("allreduce_async_ needs a CUDA tensor");
  BindStream(api, t);
  const int rc = api.allreduce_async(t.data_ptr(), static_cast<size_t>(t.numel()), DTypeEnum(t), op);
  if (rc != 0) throw std::runtime_error("RabitDeviceAllreduceAsync failed (" + std::to_string(rc) + ")");
  return t;
}

at::Tensor BroadcastAsync(at::Tensor t, int root) {
  Api &api = GetApi();
  Check(t);
  BindStream(api, t);
  const int rc = api.broadcast_async(t.data_ptr(), static_cast<size_t>(t.numel()) * t.element_size(), root);
  if (rc != 0) throw std::runtime_error("RabitDeviceBroadcastAsync failed (" + std::to_string(rc) + ")");
  return t;
}

at::Tensor AllgatherAsync(at::Tensor out, int64_t begin, int64_t count) {
  Api &api = GetApi();
  Check(out);
  if (!out.is_cuda()) throw std::invalid_argument("allgather_async_ needs a CUDA tensor");
  BindStream(api, out);
  const size_t es = out.element_size();
  const int rc = api.allgather_async(out.data_ptr(), static_cast<size_t>(out.numel()) * es, static_cast<size_t>(begin) * es,
                                     static_cast<size_t>(begin + count) * es);
  if (rc != 0) throw std::runtime_error("RabitDeviceAllgatherAsync failed (" + std::to_string(rc) + ")");
  return out;
}

/*! \brief the blocking, fault-tolerant call (no lazy prepare here: torch_binding.allreduce_ handles that in Python) */
at::Tensor Allreduce(at::Tensor t, int op) {
  Api &api = GetApi();
  Check(t);
  BindStream(api, t);
  const int dt = DTypeEnum(t);
  void *ptr = t.data_ptr();
  const size_t n = static_cast<size_t>(t.numel());
  {
    pybind11::gil_scoped_release nogil;   // the call blocks on the other ranks
    api.allreduce(ptr, n, dt, op, nullptr, nullptr);
  }
  return t;
}

at::Tensor Broadcast(at::Tensor t, int root) {
  Api &api = GetApi();
  Check(t);
  BindStream(api, t);
  void *ptr = t.data_ptr();
  const unsigned long nbytes = static_cast<unsigned long>(t.numel()) * t.element_size();  // NOLINT(*)
  {
    pybind11::gil_scoped_release nogil;
    api.broadcast(ptr, nbytes, root);
  }
  return t;
}

void Synchronize() {
  const int rc = GetApi().synchronize();
  if (rc != 0) {
    static const char *what[] = {"", "consensus mismatch", "peer timeout", "peer lost"};
    throw std::runtime_error(std::string("device collective failed: ") + (rc > 0 && rc < 4 ? what[rc] : "no backend"));
  }
}

void ResetStreamCache() { g_last_stream = reinterpret_cast<void *>(-1); }

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  m.doc() = "rabit-b200 collectives on torch tensors (native binding)";
  m.def("allreduce_async_", &AllreduceAsync, "enqueue an in-place allreduce on the current CUDA stream");
  m.def("broadcast_async_", &BroadcastAsync, "enqueue an in-place broadcast on the current CUDA stream");
  m.def("allgather_async_", &AllgatherAsync, "enqueue a ring-ordered allgather on the current CUDA stream");
  m.def("allreduce_", &Allreduce, "blocking, fault-tolerant in-place allreduce");
  m.def("broadcast_", &Broadcast, "blocking, fault-tolerant in-place broadcast");
  m.def("synchronize", &Synchronize, "wait for the enqueued collectives");
  m.def("launch_count", []() { return GetApi().launch_count(); });
  m.def("reset_stream_cache", &ResetStreamCache);
  m.def("bind_library", &BindLibrary, "addresses of the engine's C entry points (from the ctypes handle in use)");
  m.def("unbind_library", []() { g_api = Api(); });
}
