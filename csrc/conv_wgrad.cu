// Filter-gradient ("wgrad") convolution for sm_100a on tcgen05/TMEM.
//
//   dW[k, r, s, c] += sum_{m in pixel range} dY[m, k] * X[pixel(m) + (r,s), c]
//
// GEMM view: M' = Cout (128 rows per CTA), N' = a slab of Cin (<= 256), K' = output pixels.
// Both operands are "MN-major" for the tensor core - the reduction index (pixel) is the slow
// axis in memory - which tcgen05.mma supports natively for 16-bit types, so neither tensor is
// transposed in memory:
//   A' = dY  tile, tiled TMA box   [64 pixels][<=64 out-channels], swizzled
//   B' = X   tile, im2col TMA box  [64 pixels][<=64 in-channels] of ONE filter tap, swizzled,
//        zero-filled at the image border, honouring the forward stride.
// Work is split over (tap, Cout tile, Cin slab, pixel range); every CTA reduces its pixel range
// into TMEM and adds the fp32 tile into dW with vector red.global (split-K).
//
// Replaces cudnnConvolutionBackwardFilter in the reference's autodiff graph (op K3).
#include <stdlib.h>

#include "common.cuh"
#include "host.cuh"
#include "../include/drn_b200.h"

namespace drn {
namespace {

constexpr int kPix = 64;          // pixels (GEMM-K) per pipeline stage
constexpr int kThreads = 192;     // warp0 TMA, warp1 MMA (+TMEM alloc), warps 2-5 epilogue
constexpr int kPreThreads = 128;  // PRE variant: warps 6-9 apply BatchNorm+ReLU to the X tiles in smem
constexpr int kMaxStages = 8;

struct WgradParams {
  int chunks_total;     // ceil(M / 64)
  int chunks_per_split;
  int splits;
  int co_tiles, col_tiles;
  int cols_per_cta;     // columns of the flattened (tap, channel) axis handled by one CTA (<= 256)
  int cols_total;       // R*S*C
  int S;
  int PQ, Q;
  int stride, pad, pad_w;
  int Kout, C;
  float* dw;
  const float2* pre_ss;  // PRE: (scale, shift) per input channel, X is the RAW tensor
  unsigned int* det_sem; // deterministic mode: one semaphore per (Cout tile, column tile), else nullptr
  int H_in, W_in;        // input plane (PRE: taps that fall into the zero padding stay zero)
  int taps;              // PRE: 1 when a box can contain padding (filter larger than 1x1 or padded)
  int debug;             // DRN_WGRAD_DEBUG (timing experiments, wrong results): 1 = no red.global, 2 = no MMA
};

template <int CWA, int CWB>
struct WCfg {
  static constexpr int kMaxCols = 256;
  static constexpr int kRBA = CWA * 2, kRBB = CWB * 2;
  static constexpr int kABoxes = 128 / CWA, kMaxBBoxes = kMaxCols / CWB;
  static constexpr int kABoxBytes = kPix * kRBA, kBBoxBytes = kPix * kRBB;
  static constexpr int kABytes = kABoxes * kABoxBytes;          // 16 KB
  static constexpr int kBBytes = kMaxBBoxes * kBBoxBytes;       // 32 KB
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = 4;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 1024;
  static constexpr int kTmemCols = 256;
};

__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c),
               "f"(d)
               : "memory");
}

// One CTA: rows = 128 output channels, columns = a <=256-wide window of the flattened
// (filter tap, input channel) axis - which is exactly the memory order of dW[k][r][s][c] - and a
// range of output pixels as the reduction.  Every column box of CWB channels is one im2col TMA
// load (its own tap), so the dY tile is fetched once for up to 16 taps.
template <int CWA, int CWB, bool PRE = false>
__global__ void __launch_bounds__(PRE ? kThreads + kPreThreads : kThreads, 1)
conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmapDY, const __grid_constant__ CUtensorMap tmapX,
                  const WgradParams p) {
  using C_ = WCfg<CWA, CWB>;
  constexpr int kStages = C_::kStages;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t sStages = smem_base;
  const uint32_t sBar = sStages + kStages * C_::kStageBytes;
  auto full_bar = [&](int s) { return sBar + 8u * s; };
  auto empty_bar = [&](int s) { return sBar + 8u * (kMaxStages + s); };
  const uint32_t tfull_bar = sBar + 8u * (2 * kMaxStages);
  const uint32_t sTmemPtr = sBar + 8u * (2 * kMaxStages + 1);
  auto ready_bar = [&](int s) { return sBar + 8u * (2 * kMaxStages + 2 + s); };  // PRE
  volatile uint32_t* tmem_ptr_gen =
      reinterpret_cast<volatile uint32_t*>(smem_gen + (sTmemPtr - smem_base));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  pdl_trigger();

  // decode the work unit
  int u = blockIdx.x;
  const int split = u % p.splits;
  u /= p.splits;
  const int col_tile = u % p.col_tiles;
  const int co_tile = u / p.col_tiles;
  const int n0 = col_tile * p.cols_per_cta;
  int ncols = p.cols_total - n0;
  if (ncols > p.cols_per_cta) ncols = p.cols_per_cta;
  const int nboxes = ncols / CWB;
  const int chunk_begin = split * p.chunks_per_split;
  int chunk_end = chunk_begin + p.chunks_per_split;
  if (chunk_end > p.chunks_total) chunk_end = p.chunks_total;
  const int n_chunks = chunk_end - chunk_begin;  // may be <= 0 for trailing splits

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmapDY);
    tma_prefetch_desc(&tmapX);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(full_bar(i), 1);
      mbar_init(empty_bar(i), 1);
    }
    mbar_init(tfull_bar, 1);
    if (PRE)
      for (int i = 0; i < kStages; ++i) mbar_init(ready_bar(i), kPreThreads / 32);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(sTmemPtr, C_::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;
  pdl_wait();

  if (n_chunks > 0) {
    if (warp == 0) {
      if (elect_one()) {
        int stage = 0;
        uint32_t phase = 0;
        // dY boxes that lie entirely beyond Cout are not fetched at all (their rows of the 128-row
        // UMMA tile then hold stale shared memory: those accumulator rows are never written out)
        int a_boxes = (p.Kout - co_tile * 128 + CWA - 1) / CWA;
        if (a_boxes > C_::kABoxes) a_boxes = C_::kABoxes;
        const uint32_t tx_bytes = a_boxes * C_::kABoxBytes + nboxes * C_::kBBoxBytes;
        // first box of this CTA's column window (constant over the pixel loop): no divisions inside
        const int tap0 = n0 / p.C, c00 = n0 - tap0 * p.C;
        const int r0 = tap0 / p.S, s0 = tap0 - r0 * p.S;
        // output-pixel cursor (image, row, column) of the chunk's first pixel, advanced incrementally
        const int Pn = p.PQ / p.Q;
        int img, op, oq;
        {
          const int m_first = chunk_begin * kPix;
          img = m_first / p.PQ;
          const int rem = m_first - img * p.PQ;
          op = rem / p.Q;
          oq = rem - op * p.Q;
        }
        for (int ch = chunk_begin; ch < chunk_end; ++ch) {
          const int m0 = ch * kPix;
          const int w0 = oq * p.stride - p.pad_w, h0 = op * p.stride - p.pad;
          mbar_wait(empty_bar(stage), phase ^ 1, 0x110 + stage);
          const uint32_t a_dst = sStages + stage * C_::kStageBytes;
          const uint32_t b_dst = a_dst + C_::kABytes;
          mbar_arrive_expect_tx(full_bar(stage), tx_bytes);
#pragma unroll
          for (int i = 0; i < C_::kABoxes; ++i)
            if (i < a_boxes)
              tma_load_2d(a_dst + i * C_::kABoxBytes, &tmapDY, full_bar(stage),
                          co_tile * 128 + i * CWA, m0);
          int c0 = c00, r = r0, s = s0;
          for (int i = 0; i < nboxes; ++i) {
            tma_load_im2col_4d(b_dst + i * C_::kBBoxBytes, &tmapX, full_bar(stage), c0, w0, h0, img,
                               static_cast<uint16_t>(s), static_cast<uint16_t>(r));
            c0 += CWB;
            if (c0 >= p.C) {
              c0 = 0;
              if (++s == p.S) {
                s = 0;
                ++r;
              }
            }
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
          oq += kPix;
          while (oq >= p.Q) {
            oq -= p.Q;
            if (++op == Pn) {
              op = 0;
              ++img;
            }
          }
        }
      }
    } else if (warp == 1) {
      if (elect_one()) {
        const uint32_t idesc = make_idesc_bf16(128, ncols, 1, 1);
        constexpr uint32_t kLayA = umma_layout_for_swizzle_bytes(C_::kRBA);
        constexpr uint32_t kLayB = umma_layout_for_swizzle_bytes(C_::kRBB);
        int stage = 0;
        uint32_t phase = 0;
        for (int ch = 0; ch < n_chunks; ++ch) {
          mbar_wait(PRE ? ready_bar(stage) : full_bar(stage), phase, 0x120 + stage);
          tc_fence_after();
          const uint32_t a_addr = sStages + stage * C_::kStageBytes;
          const uint32_t b_addr = a_addr + C_::kABytes;
#pragma unroll
          for (int kk = 0; kk < kPix / 16; ++kk) {
            // MN-major: 16 reduction rows further down = 16 * row_bytes; LBO = distance between
            // MN atoms (one TMA box), SBO = distance between 8-row groups.
            const uint64_t a_desc = make_smem_desc(a_addr + kk * 16 * C_::kRBA, C_::kABoxBytes,
                                                   8 * C_::kRBA, kLayA);
            const uint64_t b_desc = make_smem_desc(b_addr + kk * 16 * C_::kRBB, C_::kBBoxBytes,
                                                   8 * C_::kRBB, kLayB);
            if (!(p.debug & 2)) umma_f16(tmem_base, a_desc, b_desc, idesc, (ch | kk) != 0 ? 1u : 0u);
          }
          umma_commit(empty_bar(stage));
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(tfull_bar);
      }
    } else if (PRE && warp >= 6) {
      // X-tile transform: relu(x * scale + shift) in place, box by box (one box = 64 pixels x 64
      // channels of ONE filter tap; 128-byte rows, 128B swizzle: logical 16-byte chunk g of row r
      // sits at physical chunk g ^ (r & 7)).  Out-of-range pixel rows need no masking here: their
      // dY rows are zero-filled by TMA, so whatever the transform makes of them is multiplied by 0.
      if constexpr (PRE) {
        constexpr int kCPR = CWB / 8;                    // 16-byte chunks per row of a box
        constexpr int kR = kPix * kCPR / kPreThreads;    // rows per thread and box (4 / 2 / 1)
        constexpr int kRowStep = kPreThreads / kCPR;
        constexpr int kSwzShift = CWB == 64 ? 0 : (CWB == 32 ? 1 : 2);
        const int tt = threadIdx.x - kThreads;
        const int g = tt % kCPR;
        const int r0 = tt / kCPR;  // rows r0 + kRowStep * j
        uint32_t off[kR];
#pragma unroll
        for (int j = 0; j < kR; ++j) {
          const int row = r0 + kRowStep * j;
          off[j] = row * C_::kRBB + ((g ^ ((row >> kSwzShift) & (kCPR - 1))) << 4);
        }
        int stage = 0;
        uint32_t phase = 0;
        // first box of this CTA's column window: (tap row, tap column, channel); boxes wrap from tap to tap
        const int tap_first = n0 / p.C, c_first = n0 - tap_first * p.C;
        const int fr_first = tap_first / p.S, fs_first = tap_first - fr_first * p.S;
        for (int ch = 0; ch < n_chunks; ++ch) {
          // (input row, input column) under tap (0,0) of my rows' output pixels: a tap that falls into
          // the zero padding must stay zero after relu(0 * scale + shift).  (Rows beyond M need nothing:
          // their dY rows are zero-filled by TMA, so whatever the transform makes of them is multiplied by 0.)
          int ih0[kR], iw0[kR];
          if (p.taps) {
#pragma unroll
            for (int j = 0; j < kR; ++j) {
              const long long m = (long long)(chunk_begin + ch) * kPix + r0 + kRowStep * j;
              const int rem = (int)(m % p.PQ);
              const int op = rem / p.Q;
              ih0[j] = op * p.stride - p.pad;
              iw0[j] = (rem - op * p.Q) * p.stride - p.pad_w;
            }
          }
          mbar_wait(full_bar(stage), phase, 0x160 + stage);
          const uint32_t b_addr = sStages + stage * C_::kStageBytes + C_::kABytes;
          int c0 = c_first, fr = fr_first, fs = fs_first;
          for (int i = 0; i < nboxes; ++i) {
            const float4* sp = reinterpret_cast<const float4*>(p.pre_ss + c0 + g * 8);
            float4 q[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) q[k] = __ldg(sp + k);
            const uint32_t box = b_addr + i * C_::kBBoxBytes;
            uint4 v[kR];
#pragma unroll
            for (int j = 0; j < kR; ++j) v[j] = ld_shared_v4(box + off[j]);
#pragma unroll
            for (int j = 0; j < kR; ++j) {
              uint32_t w[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
#pragma unroll
              for (int k = 0; k < 4; ++k)
                w[k] = pack_relu_bf16x2(fmaf(bf16lo(w[k]), q[k].x, q[k].y), fmaf(bf16hi(w[k]), q[k].z, q[k].w));
              if (p.taps && (static_cast<unsigned>(ih0[j] + fr) >= static_cast<unsigned>(p.H_in) ||
                             static_cast<unsigned>(iw0[j] + fs) >= static_cast<unsigned>(p.W_in)))
                w[0] = w[1] = w[2] = w[3] = 0u;
              v[j] = make_uint4(w[0], w[1], w[2], w[3]);
            }
#pragma unroll
            for (int j = 0; j < kR; ++j) st_shared_v4(box + off[j], v[j].x, v[j].y, v[j].z, v[j].w);
            c0 += CWB;
            if (c0 >= p.C) {
              c0 = 0;
              if (++fs == p.S) {
                fs = 0;
                ++fr;
              }
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) mbar_arrive(ready_bar(stage));
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    } else if (warp < 6) {
      // epilogue: warps 2..5 -> TMEM lane quadrants 2,3,0,1
      const int quad = warp & 3;
      const int row = quad * 32 + lane;
      const int co = co_tile * 128 + row;
      mbar_wait(tfull_bar, 0, 0x140);
      tc_fence_after();
      if (threadIdx.x == 64) pdl_trigger_late();   // only the split-K reduction is left
      // deterministic mode: the split-K partial tiles of one output tile are added in split order
      unsigned int* sem = (p.det_sem != nullptr && p.splits > 1) ? p.det_sem + blockIdx.x / p.splits : nullptr;
      if (sem != nullptr) {
        if (threadIdx.x == 64) {
          const uint64_t t0 = globaltimer_ns();
          uint32_t it = 0;
          while (det_ld_acquire(sem) != (unsigned)split) {
            __nanosleep(64);
            if (((++it) & 0x3ff) == 0 && globaltimer_ns() - t0 > DRN_WATCHDOG_NS) watchdog_fire(0x148);
          }
        }
        named_bar_sync(1, 128);
      }
      float* dst_row = p.dw + (long long)co * p.cols_total + n0;
#pragma unroll 1
      for (int c0 = 0; c0 < ncols; c0 += 16) {
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + c0;
        uint32_t v[16];
        tmem_ld_32x16(taddr, v);
        tmem_ld_wait();
        if (co < p.Kout && !(p.debug & 1)) {
#pragma unroll
          for (int i = 0; i < 16; i += 4)
            red_add_v4(dst_row + c0 + i, __uint_as_float(v[i]), __uint_as_float(v[i + 1]),
                       __uint_as_float(v[i + 2]), __uint_as_float(v[i + 3]));
        }
      }
      if (sem != nullptr) {
        __threadfence();
        named_bar_sync(1, 128);
        if (threadIdx.x == 64) det_st_release(sem, split + 1 == p.splits ? 0u : (unsigned)(split + 1));
      }
    }
  } else if (p.det_sem != nullptr && p.splits > 1 && threadIdx.x == 64) {
    // a trailing split without pixels still takes (and passes on) its turn
    unsigned int* sem = p.det_sem + blockIdx.x / p.splits;
    while (det_ld_acquire(sem) != (unsigned)split) __nanosleep(64);
    det_st_release(sem, split + 1 == p.splits ? 0u : (unsigned)(split + 1));
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C_::kTmemCols);
  }
}

template <int CWA, int CWB>
int launch_wgrad(const drn_conv_desc& d, const void* x, const void* dy, float* dw, int split_k,
                 cudaStream_t stream) {
  using C_ = WCfg<CWA, CWB>;
  const long long M = 1LL * d.N * d.P * d.Q;
  WgradParams p;
  p.chunks_total = (int)((M + kPix - 1) / kPix);
  p.co_tiles = (d.K + 127) / 128;
  p.cols_total = d.R * d.S * d.C;
  int tiles = (p.cols_total + C_::kMaxCols - 1) / C_::kMaxCols;
  int cols = (p.cols_total + tiles - 1) / tiles;
  cols = (cols + CWB - 1) / CWB * CWB;
  if (cols % 16 != 0) cols = (cols + 15) / 16 * 16;
  p.cols_per_cta = cols;
  p.col_tiles = (p.cols_total + cols - 1) / cols;
  const int units = p.col_tiles * p.co_tiles;
  int splits = split_k;
  if (splits <= 0) {
    // keep at least 8 chunks (512 pixels) per CTA
    // measured on B200 (bench/sweep_wgrad2.sh): one round of CTAs is best for every ResNet shape -
    // fewer split-K partial tiles to reduce with red.global and no second pipeline fill.
    static const int waves_env = getenv("DRN_WGRAD_WAVES") ? atoi(getenv("DRN_WGRAD_WAVES")) : 0;
    const int waves = waves_env > 0 ? waves_env : 1;
    // round DOWN: one CTA per SM is resident (192 KB of stages), so a grid of waves*SMs + a few
    // CTAs costs a whole extra round
    static const int sms_env = getenv("DRN_WGRAD_SMS") ? atoi(getenv("DRN_WGRAD_SMS")) : 0;
    splits = (waves * (sms_env > 0 ? sms_env : num_sms())) / units;
    int max_splits = p.chunks_total / 8;
    if (max_splits < 1) max_splits = 1;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
  }
  if (splits > p.chunks_total) splits = p.chunks_total;
  static const int debug_env = getenv("DRN_WGRAD_DEBUG") ? atoi(getenv("DRN_WGRAD_DEBUG")) : 0;
  p.debug = debug_env;
  p.det_sem = det_ticket(kDetFamilies);
  if (p.det_sem != nullptr && units > kDetWgradUnits) splits = 1;
  p.splits = splits;
  p.chunks_per_split = (p.chunks_total + splits - 1) / splits;
  p.S = d.S;
  p.PQ = d.P * d.Q;
  p.Q = d.Q;
  p.stride = d.stride;
  const ConvView* view = current_view();
  const int pad_w_lo = view ? view->pad_w_lo : d.pad_lo, pad_w_hi = view ? view->pad_w_hi : d.pad_hi;
  long long vstr[3] = {0, 0, 0};
  if (view) { vstr[0] = view->stride_w; vstr[1] = view->stride_h; vstr[2] = view->stride_n; }
  p.pad = d.pad_lo;
  p.pad_w = pad_w_lo;
  p.Kout = d.K;
  p.C = d.C;
  p.dw = dw;
  p.pre_ss = reinterpret_cast<const float2*>(current_pre());
  p.H_in = d.H;
  p.W_in = d.W;
  p.taps = (d.R * d.S > 1 || d.pad_lo != 0 || d.pad_hi != 0) ? 1 : 0;

  CUtensorMap tDY, tX;
  if (int e = tmap_tiled_2d(&tDY, dy, 2, (uint64_t)M, d.K, d.K, kPix, CWA, C_::kRBA)) return e;
  if (int e = tmap_im2col_nhwc(&tX, x, d.N, d.H, d.W, d.C, -d.pad_lo, -pad_w_lo,
                               d.pad_hi - (d.R - 1), pad_w_hi - (d.S - 1), d.stride, d.stride, CWB,
                               kPix, C_::kRBB, view ? vstr : nullptr))
    return e;
  static PerDeviceOnce attr_set;
  if (attr_set.need()) {
    DRN_CHECK_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel<CWA, CWB>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        C_::kSmemBytes));
    attr_set.mark();
  }
  const int grid = units * splits;
  if (p.pre_ss != nullptr) {
    DRN_REQUIRE(view == nullptr, "conv_wgrad: the BatchNorm prologue does not combine with an input view");
    static PerDeviceOnce attr_set_pre;
    if (attr_set_pre.need()) {
      DRN_CHECK_CUDA(cudaFuncSetAttribute(conv_wgrad_kernel<CWA, CWB, true>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, C_::kSmemBytes));
      attr_set_pre.mark();
    }
    DRN_CHECK_CUDA(launch_pdl_conv(conv_wgrad_kernel<CWA, CWB, true>, dim3(grid), dim3(kThreads + kPreThreads),
                                   C_::kSmemBytes, stream, tDY, tX, p));
    return 0;
  }
  DRN_CHECK_CUDA(launch_pdl_conv(conv_wgrad_kernel<CWA, CWB>, dim3(grid), dim3(kThreads), C_::kSmemBytes, stream,
                            tDY, tX, p));
  return 0;
}

template <int CWA>
int dispatch_b(const drn_conv_desc& d, const void* x, const void* dy, float* dw, int split_k,
               cudaStream_t st) {
  if (d.C == 16) return launch_wgrad<CWA, 16>(d, x, dy, dw, split_k, st);
  if (d.C == 32) return launch_wgrad<CWA, 32>(d, x, dy, dw, split_k, st);
  return launch_wgrad<CWA, 64>(d, x, dy, dw, split_k, st);
}

}  // namespace
}  // namespace drn

extern "C" int drn_conv_wgrad(const drn_conv_desc* desc, const void* x, const void* dy, float* dw,
                              int split_k, void* stream) {
  using namespace drn;
  const drn_conv_desc& d = *desc;
  DRN_REQUIRE(d.C == 16 || d.C == 32 || d.C % 64 == 0,
              "conv_wgrad: C must be 16, 32 or a multiple of 64 (got %d)", d.C);
  DRN_REQUIRE(d.K % 16 == 0, "conv_wgrad: K must be a multiple of 16 (got %d)", d.K);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (d.K == 16) return dispatch_b<16>(d, x, dy, dw, split_k, st);
  if (d.K == 32) return dispatch_b<32>(d, x, dy, dw, split_k, st);
  DRN_REQUIRE(d.K % 64 == 0, "conv_wgrad: K must be 16, 32 or a multiple of 64 (got %d)", d.K);
  return dispatch_b<64>(d, x, dy, dw, split_k, st);
}

extern "C" int drn_conv_wgrad_view(const drn_conv_desc* desc, const drn_conv_view* view, const void* x,
                                   const void* dy, float* dw, int split_k, void* stream) {
  drn::ConvView v{view->stride_w, view->stride_h, view->stride_n, view->pad_w_lo, view->pad_w_hi};
  drn::current_view() = &v;
  const int rc = drn_conv_wgrad(desc, x, dy, dw, split_k, stream);
  drn::current_view() = nullptr;
  return rc;
}

// dw += dY (x) relu(x * scale + shift): filter gradient of a convolution whose input is the
// BatchNorm+ReLU of the RAW tensor x - the activation is rebuilt in shared memory (see
// drn_conv_igemm_pre), so the forward pass never has to store it.  Any filter size / stride.
extern "C" int drn_conv_wgrad_pre(const drn_conv_desc* desc, const void* x, const void* pre_scale_shift,
                                  const void* dy, float* dw, int split_k, void* stream) {
  if (pre_scale_shift == nullptr || (reinterpret_cast<uintptr_t>(pre_scale_shift) & 15) != 0)
    return drn::fail("conv_wgrad_pre: pre_scale_shift must be a 16-byte aligned float2[C]");
  drn::current_pre() = pre_scale_shift;
  const int rc = drn_conv_wgrad(desc, x, dy, dw, split_k, stream);
  drn::current_pre() = nullptr;
  return rc;
}
