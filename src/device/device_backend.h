/*!
 * rabit-b200: what the engines need from the GPU side, as an interface so that the host
 * layer builds and runs (CPU-only) without CUDA. device_backend.cu implements it on
 * sm_100a; device_stub.cc is linked instead when the library is built without nvcc.
 *
 * No counterpart in the reference (it has no GPU path). The operations mirror the host
 * primitives of HostComm — allreduce / broadcast / allgather that may FAIL with a Status —
 * so RobustEngine drives both through the same recovery logic.
 */
#ifndef RABIT_SRC_DEVICE_DEVICE_BACKEND_H_
#define RABIT_SRC_DEVICE_DEVICE_BACKEND_H_

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "../host/comm.h"

namespace rabit {
namespace engine {

/*! \brief counters the benchmark and tests read back */
struct DeviceStats {
  uint64_t kernel_launches = 0;   // kernels of this library launched so far
  uint64_t fused_consensus_ok = 0;
  uint64_t fused_consensus_mismatch = 0;
  uint64_t timeouts = 0;
  uint64_t streamed_ops = 0;      // allreduces that ran as chunked, copy-engine fed ops
  uint64_t blob_bytes = 0;        // recovery / replica payload bytes this rank offered over NVLink
  int nvls = 0;                   // multicast path usable
  int p2p = 0;                    // peer mappings established
};

class DeviceBackend {
 public:
  virtual ~DeviceBackend() {}
  /*! \brief NULL when not compiled in, no GPU, or cfg.device == "off" */
  static DeviceBackend *Create(const Config &cfg);
  /*! \brief does ptr point into CUDA device memory? (false without CUDA) */
  static bool IsDevicePointer(const void *ptr);
  static bool Compiled();
  /*!
   * \brief plain cudaMemcpy between a device pointer and host memory, usable WITHOUT a backend instance:
   *  the engines stage device buffers through the host TCP collectives when no arena could be attached
   *  (ranks on different hosts, more than one NVLink domain, rabit_device=off). to_device: host -> device.
   */
  static void RawCopy(void *dst, const void *src, size_t nbytes, bool to_device);

  /*!
   * \brief (re)build the symmetric arena: allocate if needed, exchange handles with all
   *  peers over comm's ring, map everybody. Collective; called after every rendezvous.
   */
  virtual bool Attach(HostComm *comm) = 0;
  /*! \brief drop peer mappings (before a re-rendezvous) */
  virtual void Detach() = 0;
  virtual bool attached() const = 0;

  /*!
   * \brief in-place allreduce on device memory. `code` is this rank's consensus word
   *  (proto::Request::Code()); the kernel compares all ranks' codes in its entry barrier
   *  and returns kMismatch without touching data if they differ.
   * \param lanes values of dtype per item (1 for built-in types, 2 for a {g,h} pair ...)
   * \param keep if not NULL, the result is additionally written here (device memory):
   *   that copy is the rabit result cache entry, fused into the kernel's epilogue
   */
  virtual Status Allreduce(void *dptr, size_t count, int dtype, int op, int lanes,
                           uint32_t code, void *keep, int flags = 0) = 0;
  /*! \brief flags of Allreduce / Broadcast */
  enum {
    /*! \brief pick the kernel from rank-uniform values only (no zero-copy, no alignment-dependent variant):
     *  set by the engine after a kernel reported a consensus mismatch although every rank is at this very op,
     *  i.e. the ranks passed differently placed or aligned buffers */
    kOpConservative = 1
  };
  /*! \brief largest allreduce (bytes) that can run as ONE recoverable op right now, after trying to make room
   *  for `want` bytes (collective: may re-allocate and re-exchange the arena). HOST buffers count too. */
  virtual size_t MaxOpBytes(size_t want) = 0;
  /*! \brief op counters of the flag protocols; the engines keep them equal on all ranks (consensus envelope) */
  virtual uint32_t OpEpoch() = 0;
  virtual uint64_t OpRounds() const = 0;
  virtual void AdoptEpoch(uint32_t epoch, uint64_t rounds) = 0;
  virtual Status Broadcast(void *dptr, size_t size, int root, uint32_t code, void *keep, int flags = 0) = 0;
  virtual Status Allgather(void *dptr, size_t total, size_t slice_begin, size_t slice_end,
                           uint32_t code, void *keep) = 0;

  // ---- recovery payloads over NVLink: copy engines on a side stream, only the two ends take part ----
  /*!
   * \brief move `size` bytes from rank `holder` (src_dev if not NULL, else src_host) to every rank r with need[r]
   *  (landing in dst_host there). ALL ranks call it with the same holder / need / size; the others return at once.
   */
  virtual Status BlobDeliver(int holder, const std::vector<char> &need, const void *src_host, const void *src_dev,
                             void *dst_host, size_t size) = 0;
  /*!
   * \brief ring neighbour exchange of host blobs: offer `out` to the successor (forward) or predecessor and take
   *  `in_len` bytes from the other neighbour. Both ends of each transfer must call it (a length of 0 = nothing).
   */
  virtual Status BlobShift(const void *out, size_t out_len, void *in, size_t in_len, bool forward) = 0;

  // ---- plain memory services for the engines (staging, result cache) ----
  virtual void *Alloc(size_t nbytes) = 0;
  virtual void Free(void *dptr) = 0;
  virtual void CopyD2H(void *dst, const void *src, size_t nbytes) = 0;
  virtual void CopyH2D(void *dst, const void *src, size_t nbytes) = 0;
  virtual void CopyD2D(void *dst, const void *src, size_t nbytes) = 0;

  // ---- copy engines next to the collective stream: host buffers that ride PCIe + NVLink in chunks ----
  static const int kCopySlots = 4;
  /*! \brief enqueue host -> device on the copy-in stream and remember its completion in `slot` */
  virtual void PrefetchH2D(void *dst, const void *src, size_t nbytes, int slot) = 0;
  /*! \brief the collective stream waits (on the device) for the copy remembered in `slot` */
  virtual void AwaitPrefetch(int slot) = 0;
  /*! \brief enqueue device -> host on the copy-out stream; the source must already be complete */
  virtual void DrainD2H(void *dst, const void *src, size_t nbytes) = 0;
  /*! \brief wait for everything enqueued by DrainD2H (and PrefetchH2D) */
  virtual void DrainSync() = 0;

  virtual const DeviceStats &stats() const = 0;
  virtual std::string Describe() const = 0;
};

}  // namespace engine
}  // namespace rabit
#endif  // RABIT_SRC_DEVICE_DEVICE_BACKEND_H_
