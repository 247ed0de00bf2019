#include "host.cuh"

#include <stdarg.h>
#include <stdlib.h>

#include "common.cuh"

namespace drn {

__device__ unsigned int g_watchdog_flag = 0;
__device__ unsigned int* g_watchdog_host = nullptr;
static unsigned int* g_mailbox[64] = {};   // host side of the per-device mailbox
__device__ unsigned int g_det_tickets[kDetFamilies + kDetWgradUnits] = {};
static int g_deterministic = -1;           // drn_set_deterministic(); -1 = DRN_DETERMINISTIC / off

bool deterministic() {
  if (g_deterministic >= 0) return g_deterministic != 0;
  static const bool env = []() {
    const char* e = getenv("DRN_DETERMINISTIC");
    return e != nullptr && atoi(e) != 0;
  }();
  return env;
}

unsigned int* det_ticket(int index) {
  if (!deterministic()) return nullptr;
  static unsigned int* base[64] = {};
  const int dev = current_device();
  if (dev < 0 || dev >= 64) return nullptr;
  if (base[dev] == nullptr) {
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, g_det_tickets) != cudaSuccess) return nullptr;
    base[dev] = static_cast<unsigned int*>(p);
  }
  return base[dev] + index;
}

std::string& last_error() {
  static thread_local std::string s;
  return s;
}

int fail(const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  last_error() = buf;
  return 1;
}

const ConvView*& current_view() {
  static thread_local const ConvView* v = nullptr;
  return v;
}

const void*& current_pre() {
  static thread_local const void* v = nullptr;
  return v;
}

const void*& current_pull() {
  static thread_local const void* v = nullptr;
  return v;
}

const void*& current_epi2() {
  static thread_local const void* v = nullptr;
  return v;
}

static int g_pdl_override = -1;   // drn_set_pdl_mask(); -1 = environment / default
int pdl_mask() {
  static const int m = []() {
    const char* e = getenv("DRN_PDL");
    return e != nullptr ? atoi(e) : -1;
  }();
  if (m >= 0) return m;                       // DRN_PDL wins (experiments)
  // default 0: measured on ImageNet-sized steps, dependents that become resident early take SMs from
  // the overlapped filter-gradient stream; launch-bound CIFAR-sized steps gain ~4 % with
  // mask 2 (convolution kernels) - the trainer selects that through drn_set_pdl_mask()
  return g_pdl_override >= 0 ? g_pdl_override : 0;
}

int& pdl_override() {
  static thread_local int v = -1;
  return v;
}

int current_device() {
  int dev = -1;
  return cudaGetDevice(&dev) == cudaSuccess ? dev : -1;
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const int*, const int*,
                                   cuuint32_t, cuuint32_t, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn g_encode_tiled = nullptr;
static EncodeIm2colFn g_encode_im2col = nullptr;
static std::once_flag g_driver_once;

static void resolve_driver() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) ==
          cudaSuccess &&
      qres == cudaDriverEntryPointSuccess)
    g_encode_tiled = reinterpret_cast<EncodeTiledFn>(fn);
  fn = nullptr;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fn, cudaEnableDefault, &qres) ==
          cudaSuccess &&
      qres == cudaDriverEntryPointSuccess)
    g_encode_im2col = reinterpret_cast<EncodeIm2colFn>(fn);
}

static CUtensorMapSwizzle to_cu_swizzle(int bytes) {
  switch (bytes) {
    case 128: return CU_TENSOR_MAP_SWIZZLE_128B;
    case 64: return CU_TENSOR_MAP_SWIZZLE_64B;
    case 32: return CU_TENSOR_MAP_SWIZZLE_32B;
    default: return CU_TENSOR_MAP_SWIZZLE_NONE;
  }
}

int tmap_tiled_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t rows, uint64_t cols,
                  uint64_t row_stride_elems, uint32_t box_rows, uint32_t box_cols,
                  int swizzle_bytes) {
  std::call_once(g_driver_once, resolve_driver);
  DRN_REQUIRE(g_encode_tiled != nullptr, "cuTensorMapEncodeTiled entry point not available");
  DRN_REQUIRE(elem_bytes == 2 || elem_bytes == 4, "tmap_tiled_2d: elem_bytes must be 2 or 4");
  DRN_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "tmap_tiled_2d: base not 16B aligned");
  DRN_REQUIRE((row_stride_elems * elem_bytes) % 16 == 0,
              "tmap_tiled_2d: row stride (%llu B) must be a multiple of 16",
              (unsigned long long)(row_stride_elems * elem_bytes));
  DRN_REQUIRE(box_rows <= 256 && box_cols <= 256, "tmap_tiled_2d: box too large");
  if (swizzle_bytes)
    DRN_REQUIRE((int)(box_cols * elem_bytes) <= swizzle_bytes,
                "tmap_tiled_2d: inner box (%u B) exceeds swizzle span %d", box_cols * elem_bytes,
                swizzle_bytes);
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {row_stride_elems * (uint64_t)elem_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMapDataType dt =
      elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = g_encode_tiled(out, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, to_cu_swizzle(swizzle_bytes),
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DRN_REQUIRE(r == CUDA_SUCCESS,
              "cuTensorMapEncodeTiled failed (%d): rows=%llu cols=%llu stride=%llu box=%ux%u swz=%d",
              (int)r, (unsigned long long)rows, (unsigned long long)cols,
              (unsigned long long)row_stride_elems, box_rows, box_cols, swizzle_bytes);
  return 0;
}

int tmap_im2col_nhwc(CUtensorMap* out, const void* base, int N, int H, int W, int C, int lower_h,
                     int lower_w, int upper_h, int upper_w, int stride_h, int stride_w,
                     uint32_t channels_per_pixel, uint32_t pixels_per_column, int swizzle_bytes,
                     const long long* elem_strides) {
  std::call_once(g_driver_once, resolve_driver);
  DRN_REQUIRE(g_encode_im2col != nullptr, "cuTensorMapEncodeIm2col entry point not available");
  DRN_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "tmap_im2col: base not 16B aligned");
  DRN_REQUIRE((C * 2) % 16 == 0, "tmap_im2col: C*2 bytes must be a multiple of 16 (C=%d)", C);
  DRN_REQUIRE(pixels_per_column <= 1024 && channels_per_pixel <= 256, "tmap_im2col: box too large");
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t gstride[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  if (elem_strides != nullptr) {
    for (int i = 0; i < 3; ++i) {
      DRN_REQUIRE(elem_strides[i] > 0 && (elem_strides[i] * 2) % 16 == 0,
                  "tmap_im2col: view stride %d (%lld elements) must be a positive multiple of 8", i,
                  elem_strides[i]);
      gstride[i] = (cuuint64_t)elem_strides[i] * 2;
    }
  }
  int lower[2] = {lower_w, lower_h};
  int upper[2] = {upper_w, upper_h};
  cuuint32_t estr[4] = {1, (cuuint32_t)stride_w, (cuuint32_t)stride_h, 1};
  CUresult r = g_encode_im2col(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base),
                               gdim, gstride, lower, upper, channels_per_pixel, pixels_per_column,
                               estr, CU_TENSOR_MAP_INTERLEAVE_NONE, to_cu_swizzle(swizzle_bytes),
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  DRN_REQUIRE(r == CUDA_SUCCESS,
              "cuTensorMapEncodeIm2col failed (%d): NHWC=%d,%d,%d,%d lower=(%d,%d) upper=(%d,%d) "
              "stride=(%d,%d) cpp=%u ppc=%u swz=%d",
              (int)r, N, H, W, C, lower_h, lower_w, upper_h, upper_w, stride_h, stride_w,
              channels_per_pixel, pixels_per_column, swizzle_bytes);
  return 0;
}

}  // namespace drn

extern "C" {

const char* drn_last_error() { return drn::last_error().c_str(); }

// Gives the in-kernel watchdog a word of mapped pinned HOST memory to write its code to, once per
// device (allocates 4 bytes of pinned memory; call at start-up).  Without it the code lives in
// device memory only, which is unreadable once a trap has poisoned the context.
int drn_watchdog_arm() {
  const int dev = drn::current_device();
  if (dev < 0 || dev >= 64) return drn::fail("watchdog_arm: no current device");
  if (drn::g_mailbox[dev] != nullptr) return 0;
  unsigned int* host = nullptr;
  DRN_CHECK_CUDA(cudaHostAlloc(reinterpret_cast<void**>(&host), sizeof(unsigned int), cudaHostAllocMapped));
  *host = 0;
  unsigned int* dev_alias = nullptr;
  DRN_CHECK_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void**>(&dev_alias), host, 0));
  DRN_CHECK_CUDA(cudaMemcpyToSymbol(drn::g_watchdog_host, &dev_alias, sizeof(dev_alias)));
  drn::g_mailbox[dev] = host;
  return 0;
}

// 0 if no in-kernel watchdog has fired on the current device; otherwise the code of the
// barrier that timed out (mbarrier waits 0x10-0x160, collectives 0x200/0x201). Clears the flag.
// Works after the trap as well when drn_watchdog_arm() was called (host mailbox).
int drn_watchdog_status() {
  const int dev = drn::current_device();
  if (dev >= 0 && dev < 64 && drn::g_mailbox[dev] != nullptr) {
    volatile unsigned int* box = drn::g_mailbox[dev];
    const unsigned int code = *box;
    if (code) {
      *box = 0;
      return (int)code;
    }
  }
  unsigned int v = 0, z = 0;
  if (cudaMemcpyFromSymbol(&v, drn::g_watchdog_flag, sizeof(v)) != cudaSuccess) return -1;
  if (v) cudaMemcpyToSymbol(drn::g_watchdog_flag, &z, sizeof(z));
  return (int)v;
}

int drn_num_sms() { return drn::num_sms(); }

// Programmatic dependent launch policy: bit 0 element-wise kernels, bit 1 convolution kernels;
// -1 = back to the default.  The DRN_PDL environment variable overrides it.
int drn_set_pdl_mask(int mask) {
  drn::g_pdl_override = mask;
  return 0;
}

// Deterministic mode: 1 = every floating-point accumulation across thread blocks happens in a fixed
// order (bit-identical results from run to run, slower), 0 = off, -1 = back to DRN_DETERMINISTIC.
// Convolutions then take their BatchNorm statistics in a separate ordered pass instead of in the
// epilogue, the filter gradient reduces its split-K partial tiles in split order, the BatchNorm /
// loss / optimizer reductions take turns block by block.
int drn_set_deterministic(int on) {
  drn::g_deterministic = on;
  return 0;
}
int drn_deterministic() { return drn::deterministic() ? 1 : 0; }

}  // extern "C"
