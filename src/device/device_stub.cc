/*!
 * rabit-b200: link-time stand-in for the sm_100a backend when the library is built
 * without nvcc (RABIT_CUDA=0). Every pointer is a host pointer, there is no backend.
 */
#include "device_backend.h"

namespace rabit {
namespace engine {
DeviceBackend *CreateFakeBackend(const Config &cfg);  // src/host/fake_device.cc
DeviceBackend *DeviceBackend::Create(const Config &cfg) { return cfg.device == "fake" ? CreateFakeBackend(cfg) : nullptr; }
bool DeviceBackend::IsDevicePointer(const void *) { return false; }
bool DeviceBackend::Compiled() { return false; }
void DeviceBackend::RawCopy(void *, const void *, size_t, bool) {}
class BaseEngine;
BaseEngine *CreateNcclEngine() { return nullptr; }
}  // namespace engine
}  // namespace rabit
