// Host-side plumbing shared by the C-ABI launchers: error reporting, the driver
// entry points for tensor-map encoding (resolved at run time so the library links
// without libcuda).  Tensor maps are encoded per launch (a few hundred nanoseconds each on the
// host; inside a captured CUDA graph they are encoded once, at capture).
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <string>
#include <unordered_map>

namespace drn {

// Last error text, per thread; returned by drn_last_error().
std::string& last_error();
int fail(const char* fmt, ...);

#define DRN_CHECK_CUDA(expr)                                                              \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess)                                                                \
      return ::drn::fail("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                         __LINE__);                                                       \
  } while (0)

#define DRN_REQUIRE(cond, ...)                  \
  do {                                          \
    if (!(cond)) return ::drn::fail(__VA_ARGS__); \
  } while (0)

int num_sms();
int current_device();   // -1 on error

// "done once per device" latch for per-function attributes (cudaFuncSetAttribute is per device; the
// library may be used on several GPUs from one process).
struct PerDeviceOnce {
  bool done[64] = {};
  bool need() {
    const int d = current_device();
    return d < 0 || d >= 64 || !done[d];
  }
  void mark() {
    const int d = current_device();
    if (d >= 0 && d < 64) done[d] = true;
  }
};

// Programmatic dependent launch: the kernel may be scheduled while its predecessor in the stream is
// still draining; it must execute pdl_wait() (common.cuh) before touching memory the predecessor
// produced.  Inside a captured graph this becomes a programmatic edge.  DRN_PDL=0 turns it off.
int pdl_mask();   // bit 0: element-wise / pooling / optimizer kernels, bit 1: tcgen05 convolution kernels
// Per-launch override (thread-local, set by a launcher around ONE launch): -1 = follow pdl_mask(),
// 0 / 1 = without / with the programmatic edge.  Used by the size-based policy (DRN_PDL_SMALL_TILES).
int& pdl_override();

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_kind(int kind_bit, void (*kernel)(KArgs...), dim3 grid, dim3 block,
                                   size_t smem, cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  const int ov = pdl_override();
  cfg.numAttrs = (ov >= 0 ? ov != 0 : (pdl_mask() & kind_bit) != 0) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
// Launch with a (2,1,1) thread-block cluster (CTA pairs for tcgen05 cta_group::2).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_cluster2(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                   cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
// Scope guard of the size-based policy for element-wise kernels: programmatic edge when the tensor
// has at most DRN_PDL_SMALL_VECS 16-byte vectors (experiment; 0 = off).
struct PdlSmall {
  explicit PdlSmall(long long n_vec) {
    static const long long lim = getenv("DRN_PDL_SMALL_VECS") ? atoll(getenv("DRN_PDL_SMALL_VECS")) : 0;
    if (lim > 0 && n_vec <= lim) pdl_override() = 1;
  }
  ~PdlSmall() { pdl_override() = -1; }
};
#define launch_pdl(...) launch_pdl_kind(1, __VA_ARGS__)
#define launch_pdl_conv(...) launch_pdl_kind(2, __VA_ARGS__)

enum SwizzleMode { SWZ_NONE = 0, SWZ_32 = 32, SWZ_64 = 64, SWZ_128 = 128 };

// 2-D tiled map over a row-major [rows][cols] matrix of 2- or 4-byte elements.
// `row_stride_elems` is the distance between rows (>= cols).
int tmap_tiled_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t rows, uint64_t cols,
                  uint64_t row_stride_elems, uint32_t box_rows, uint32_t box_cols, int swizzle_bytes);

// im2col map over an NHWC bf16 tensor. Corners follow the CUDA driver convention:
// lower = -pad_lo, upper = pad_hi - (filter - 1) (dilation 1).
// `elem_strides` (optional) = {w, h, n} strides in ELEMENTS for a non-dense (possibly overlapping)
// view of the tensor; NULL = dense NHWC.
int tmap_im2col_nhwc(CUtensorMap* out, const void* base, int N, int H, int W, int C, int lower_h,
                     int lower_w, int upper_h, int upper_w, int stride_h, int stride_w,
                     uint32_t channels_per_pixel, uint32_t pixels_per_column, int swizzle_bytes,
                     const long long* elem_strides = nullptr);

// Optional view of the input tensor used by the *_view entry points (thread-local, set around the
// ordinary launchers): strides in elements and separate zero padding along W.
struct ConvView {
  long long stride_w, stride_h, stride_n;
  int pad_w_lo, pad_w_hi;
};
const ConvView*& current_view();
// Optional BatchNorm+ReLU prologue of the *_pre entry points: float2[C] (scale, shift) applied to the
// raw input before it is used as an operand (thread-local, set around the ordinary launchers).
const void*& current_pre();
// Optional "fetch the filter bank from the parameter-server shards first" request of
// drn_conv_igemm_ex (thread-local, set around the ordinary launcher); points at a drn::PullArgs.
const void*& current_pull();
// Optional second side input of the two-side epilogue modes (drn_conv_igemm_epi2); points at a
// conv_igemm.cu Epi2.
const void*& current_epi2();

// Deterministic mode (drn_set_deterministic / DRN_DETERMINISTIC=1) and the device address of ticket
// word `index` of common.cuh's g_det_tickets on the current device (nullptr when the mode is off).
bool deterministic();
unsigned int* det_ticket(int index);

}  // namespace drn
