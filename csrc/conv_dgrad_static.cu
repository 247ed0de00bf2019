// drn_conv_dgrad_static - the data gradient of a strided convolution for callers that must not
// allocate device memory.
//
// Who needs this: the ordinary entry point, drn_conv_dgrad, asks the caller for a scratch buffer
// (the zero-inserted copy of dy).  Some callers have nowhere to get one from at the moment of the
// call - for example a host process that handed the whole GPU to its own memory pool at start-up,
// so that a fresh cudaMalloc would fail or fragment the pool, or a latency-critical loop in which
// cudaMalloc/cudaFree are forbidden because they synchronise the device.  For them this variant
// keeps the scratch inside the library: a fixed array in device global memory that exists for as
// long as the library is loaded.
//
// What to know before using it:
//   * The array is DRN_DGRAD_STATIC_SCRATCH_BYTES (16 MiB) and is reserved on every GPU whose
//     context loads this library's device code, whether or not the function is ever called.  With
//     CUDA's default lazy module loading that reservation happens when the first kernel of the
//     library runs on the device (or at this function's first call) - not at dlopen.  A host pool
//     that grabs "all remaining memory" must therefore be created AFTER that first call, otherwise
//     the reservation itself can fail with an out-of-memory error.
//   * Problems whose zero-inserted dy ( 2 * N * ((P-1)*stride+1) * ((Q-1)*stride+1) * K bytes,
//     see drn_conv_dgrad_scratch_bytes ) does not fit are refused with an error; use
//     drn_conv_dgrad with your own buffer for those.
//   * The first call on each device looks up the device address of the array
//     (cudaGetSymbolAddress).  Make that first call during start-up, before any stream capture is
//     in progress on that device; later calls only enqueue kernels.
//   * All calls on a device share the one array.  Each call rewrites every byte of the region it
//     uses before reading it (the zero-insertion kernel writes the zeros as well as the samples),
//     so nothing needs clearing between calls - but two calls that may run on the GPU at the same
//     time (different streams, or graphs launched concurrently) would overwrite each other's
//     scratch.  Serialise them.
//   * It never calls cudaMalloc, cudaFree or any allocating library routine.
// This entry point has no counterpart in the reference.
#include <cuda_runtime.h>

#include <atomic>

#include "host.cuh"
#include "../include/drn_b200.h"

namespace {

static __device__ __align__(256) unsigned char g_dgrad_scratch[DRN_DGRAD_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
// device address of g_dgrad_scratch, per device; host threads may race to fill an entry (they all
// store the same value), hence atomics
std::atomic<void*> g_scratch_addr[kMaxDevices];

// In host code `g_dgrad_scratch` is only a handle for the runtime, not a device address: the real
// address differs per device and has to be asked for with that device current.
int scratch_for_current_device(void** out) {
  int dev = -1;
  DRN_CHECK_CUDA(cudaGetDevice(&dev));
  DRN_REQUIRE(dev >= 0 && dev < kMaxDevices, "conv_dgrad_static: device ordinal %d out of range", dev);
  void* p = g_scratch_addr[dev].load(std::memory_order_acquire);
  if (p == nullptr) {
    DRN_CHECK_CUDA(cudaGetSymbolAddress(&p, g_dgrad_scratch));
    DRN_REQUIRE(p != nullptr, "conv_dgrad_static: cudaGetSymbolAddress returned a null address");
    g_scratch_addr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return 0;
}

}  // namespace

extern "C" int drn_conv_dgrad_static(const drn_conv_desc* desc, const void* dy, const void* w_dgrad,
                                     void* dx, const void* side, float* stats, int stats_ld,
                                     const void* bn_scale_shift, int mode, void* stream) {
  using namespace drn;
  const long long need = drn_conv_dgrad_scratch_bytes(desc);
  DRN_REQUIRE(need <= (long long)DRN_DGRAD_STATIC_SCRATCH_BYTES,
              "conv_dgrad_static: this problem needs %lld bytes of scratch but the built-in buffer "
              "holds %lld; call drn_conv_dgrad with a caller-provided buffer instead",
              need, (long long)DRN_DGRAD_STATIC_SCRATCH_BYTES);
  void* scratch = nullptr;
  if (int e = scratch_for_current_device(&scratch)) return e;
  return drn_conv_dgrad(desc, dy, w_dgrad, dx, side, stats, stats_ld, bn_scale_shift, mode, scratch,
                        (long long)DRN_DGRAD_STATIC_SCRATCH_BYTES, stream);
}
