// Optimizer and weight-layout kernels.
//
// drn_sgd_momentum: the reference's update (resnet_model.py:85-99): L2 weight decay on EVERY
//   trainable tensor (BN gamma/beta and the dense bias included), then
//   m = mu*m + g ; w = w - lr*m   (tf.train.MomentumOptimizer, no Nesterov), on fp32 master
//   weights kept in one flat arena; also writes the bf16 working copy the convolutions read,
//   clears the gradient (wgrad accumulates with atomics) and returns sum(w^2) for the reported
//   cost.  Hyper-parameters live in device memory so a captured CUDA graph follows the schedule.
//   Ops K17, K19.
// drn_weight_prep_dgrad: builds the flipped + transposed bf16 filter bank that turns the
//   implicit-GEMM forward kernel into the data-gradient kernel.
// drn_upsample_zero: zero-insertion of dy for the data gradient of strided convolutions.
#include "common.cuh"
#include "host.cuh"
#include "../include/drn_b200.h"

namespace drn {
namespace {

__global__ void sgd_momentum_kernel(float4* __restrict__ w, float4* __restrict__ g,
                                    float4* __restrict__ m, uint2* __restrict__ shadow,
                                    long long n4, const float* __restrict__ hyper,
                                    float* __restrict__ l2_sum, int zero_grad, unsigned int* det) {
  pdl_trigger();
  pdl_wait();
  const float lr = hyper[0], mu = hyper[1], wd = hyper[2], gs = hyper[3];
  float l2 = 0.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 wv = w[i], gv = g[i], mv = m[i];
    l2 += wv.x * wv.x + wv.y * wv.y + wv.z * wv.z + wv.w * wv.w;
    mv.x = fmaf(mu, mv.x, fmaf(wd, wv.x, gv.x * gs));
    mv.y = fmaf(mu, mv.y, fmaf(wd, wv.y, gv.y * gs));
    mv.z = fmaf(mu, mv.z, fmaf(wd, wv.z, gv.z * gs));
    mv.w = fmaf(mu, mv.w, fmaf(wd, wv.w, gv.w * gs));
    wv.x = fmaf(-lr, mv.x, wv.x);
    wv.y = fmaf(-lr, mv.y, wv.y);
    wv.z = fmaf(-lr, mv.z, wv.z);
    wv.w = fmaf(-lr, mv.w, wv.w);
    w[i] = wv;
    m[i] = mv;
    if (zero_grad) g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (shadow) shadow[i] = make_uint2(pack_bf16x2(wv.x, wv.y), pack_bf16x2(wv.z, wv.w));
  }
  if (l2_sum != nullptr) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) l2 += __shfl_xor_sync(0xffffffffu, l2, o);
    __shared__ float s[32];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = l2;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += s[i];
      if (det != nullptr) {     // deterministic mode: the blocks add in index order
        while (det_ld_acquire(det) != blockIdx.x) __nanosleep(64);
      }
      atomicAdd(l2_sum, t);
      if (det != nullptr) {
        __threadfence();
        det_st_release(det, blockIdx.x + 1 == gridDim.x ? 0u : blockIdx.x + 1);
      }
    }
  }
}

// dst[c][R-1-r][S-1-s][k] = src[k][r][s][c]     (32x32 smem tile transpose per filter tap)
__global__ void weight_prep_dgrad_kernel(const __nv_bfloat16* __restrict__ src,
                                         __nv_bfloat16* __restrict__ dst, int K, int RS, int C) {
  __shared__ __nv_bfloat16 tile[32][33];
  const int tap = blockIdx.z;
  const int c0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int k = k0 + j, c = c0 + threadIdx.x;
    if (k < K && c < C) tile[j][threadIdx.x] = src[((long long)k * RS + tap) * C + c];
  }
  __syncthreads();
  const int ftap = RS - 1 - tap;  // flipping r and s together reverses the linear tap index
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int c = c0 + j, k = k0 + threadIdx.x;
    if (k < K && c < C) dst[((long long)c * RS + ftap) * K + k] = tile[threadIdx.x][j];
  }
}

// All layers in one launch: block -> (layer, tap, k-tile, c-tile) through a table of tile offsets.
struct PrepEntry {
  long long src_off, dst_off;  // element offsets into the bf16 weight arena / the dgrad-bank arena
  int K, RS, C, tile_begin;    // tile_begin = first block index of this layer
};
__global__ void __launch_bounds__(256)
weight_prep_dgrad_batched_kernel(const __nv_bfloat16* __restrict__ src_base,
                                 __nv_bfloat16* __restrict__ dst_base,
                                 const PrepEntry* __restrict__ table, int n_layers) {
  pdl_trigger();
  pdl_wait();
  // 64 x 64 tiles moved with 4-byte accesses: a warp reads / writes 128 contiguous bytes of a row
  // (the 32 x 32 tiles with 2-byte accesses of the first version moved 64-byte segments: 100 us
  // for the 23.5 M weights of ResNet-50 where ~30 us of traffic is needed)
  __shared__ uint32_t tile[64][33];
  // binary search: last layer whose tile_begin <= blockIdx.x
  int lo = 0, hi = n_layers - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (table[mid].tile_begin <= (int)blockIdx.x) lo = mid; else hi = mid - 1;
  }
  const PrepEntry e = table[lo];
  int t = blockIdx.x - e.tile_begin;
  const int ct = (e.C + 63) / 64, kt = (e.K + 63) / 64;
  const int c0 = (t % ct) * 64;
  t /= ct;
  const int k0 = (t % kt) * 64;
  const int tap = t / kt;
  const __nv_bfloat16* src = src_base + e.src_off;
  __nv_bfloat16* dst = dst_base + e.dst_off;
  const int tx = threadIdx.x, ty = threadIdx.y;   // (32, 8)
  for (int j = ty; j < 64; j += 8) {
    const int k = k0 + j, c = c0 + 2 * tx;
    uint32_t v = 0u;
    if (k < e.K && c < e.C)     // C is even: a pair is in range or out of range as a whole
      v = *reinterpret_cast<const uint32_t*>(src + ((long long)k * e.RS + tap) * e.C + c);
    tile[j][tx] = v;
  }
  __syncthreads();
  const int ftap = e.RS - 1 - tap;  // flipping r and s together reverses the linear tap index
  const unsigned short* t16 = reinterpret_cast<const unsigned short*>(&tile[0][0]);
  for (int j = ty; j < 64; j += 8) {
    const int c = c0 + j, k = k0 + 2 * tx;
    if (c < e.C && k < e.K) {   // K is even as well
      const uint32_t lo16 = t16[(2 * tx) * 66 + j], hi16 = t16[(2 * tx + 1) * 66 + j];
      *reinterpret_cast<uint32_t*>(dst + ((long long)c * e.RS + ftap) * e.K + k) = lo16 | (hi16 << 16);
    }
  }
}

// out[n, p*s, q*s, :] = dy[n, p, q, :], zeros elsewhere. 8 channels per thread.
__global__ void upsample_zero_kernel(const uint4* __restrict__ dy, uint4* __restrict__ out, int N,
                                     int P, int Q, int U, int V, int c8, int stride) {
  pdl_trigger();
  pdl_wait();
  const long long total = (long long)N * U * V * c8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % c8);
    long long t = i / c8;
    const int v = (int)(t % V);
    t /= V;
    const int u = (int)(t % U);
    const long long n = t / U;
    uint4 val = make_uint4(0, 0, 0, 0);
    if (u % stride == 0 && v % stride == 0) {
      const int p = u / stride, q = v / stride;
      if (p < P && q < Q) val = dy[((n * P + p) * Q + q) * c8 + cg];
    }
    out[i] = val;
  }
}

__global__ void f32_to_bf16_kernel(const float4* __restrict__ src, uint2* __restrict__ dst,
                                   long long n4) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (long long)gridDim.x * blockDim.x) {
    const float4 v = src[i];
    dst[i] = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
}

__global__ void mul_inplace_kernel(float* __restrict__ a, const float* __restrict__ b, long long n) {
  pdl_trigger();
  pdl_wait();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    a[i] *= b[i];
}

int ew_grid(long long n, int threads) {
  long long blocks = (n + threads - 1) / threads;
  const long long cap = (long long)num_sms() * 8;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}


// Filter banks of the four parity phases of a 3x3 / stride-2 / pad-1 data gradient.
// dx[2i+a][2j+b] only receives the filter rows r with (a + 1 - r) even: a = 0 -> {r = 1};
// a = 1 -> {r = 2 (dy row i), r = 0 (dy row i + 1)}, likewise for columns.  Bank (a,b) is the filter
// of a stride-1 convolution over dy with (1+a) x (1+b) taps, stored [C][1+a][1+b][K]:
//   bank[c][tr][ts][k] = w[k][rmap(a,tr)][rmap(b,ts)][c],  rmap(0,0) = 1, rmap(1,0) = 2, rmap(1,1) = 0.
__global__ void weight_prep_dgrad_s2_kernel(const __nv_bfloat16* __restrict__ w,
                                            __nv_bfloat16* __restrict__ banks, int K, int C) {
  pdl_trigger();
  pdl_wait();
  const long long total = 9LL * K * C;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    // decode the destination: phase banks are laid out back to back with 1, 2, 2, 4 taps
    const long long ck = (long long)C * K;
    long long rem = idx;
    int phase = 0, taps = 1;
    const int ntaps[4] = {1, 2, 2, 4};
#pragma unroll
    for (int ph = 0; ph < 4; ++ph) {
      if (rem >= 0 && rem < ntaps[ph] * ck) {
        phase = ph;
        taps = ntaps[ph];
        break;
      }
      rem -= ntaps[ph] * ck;
    }
    const int a = phase >> 1, b = phase & 1;
    const int k = (int)(rem % K);
    long long t = rem / K;
    const int tap = (int)(t % taps);
    const int c = (int)(t / taps);
    const int tr = tap / (1 + b), ts = tap % (1 + b);
    const int r = a == 0 ? 1 : (tr == 0 ? 2 : 0);
    const int s_ = b == 0 ? 1 : (ts == 0 ? 2 : 0);
    banks[idx] = w[((long long)k * 9 + r * 3 + s_) * C + c];
  }
}

}  // namespace
}  // namespace drn

using namespace drn;

// n must be a multiple of 4 (pad the arena). hyper = device float[4] {lr, momentum, wd, grad_scale}.
extern "C" int drn_sgd_momentum(float* master, float* grad, float* mom, void* shadow_bf16,
                                long long n, const float* hyper, float* l2_sum, int zero_grad,
                                void* stream) {
  DRN_REQUIRE(n % 4 == 0, "sgd_momentum: n must be a multiple of 4");
  const long long n4 = n / 4;
  unsigned int* det = l2_sum != nullptr ? det_ticket(kDetOptim) : nullptr;
  int grid = ew_grid(n4, 256);
  if (det != nullptr && grid > 2 * num_sms()) grid = 2 * num_sms();   // the blocks take turns: keep them few
  DRN_CHECK_CUDA(launch_pdl(sgd_momentum_kernel, dim3(grid), dim3(256), 0, (cudaStream_t)stream,
                            (float4*)master, (float4*)grad, (float4*)mom, (uint2*)shadow_bf16, n4, hyper, l2_sum,
                            zero_grad, det));
  return 0;
}

extern "C" int drn_weight_prep_dgrad(const void* w_krsc, void* w_dgrad, int K, int R, int S, int C,
                                     void* stream) {
  dim3 grid((C + 31) / 32, (K + 31) / 32, R * S), block(32, 8);
  weight_prep_dgrad_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)w_krsc, (__nv_bfloat16*)w_dgrad, K, R * S, C);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int drn_upsample_zero(const void* dy, void* out, int N, int P, int Q, int U, int V, int C,
                                 int stride, void* stream) {
  DRN_REQUIRE(C % 8 == 0, "upsample_zero: C must be a multiple of 8");
  const long long total = (long long)N * U * V * (C / 8);
  DRN_CHECK_CUDA(launch_pdl(upsample_zero_kernel, dim3(ew_grid(total, 256)), dim3(256), 0, (cudaStream_t)stream,
                            (const uint4*)dy, (uint4*)out, N, P, Q, U, V, C / 8, stride));
  return 0;
}

extern "C" int drn_f32_to_bf16(const float* src, void* dst, long long n, void* stream) {
  DRN_REQUIRE(n % 4 == 0, "f32_to_bf16: n must be a multiple of 4");
  f32_to_bf16_kernel<<<ew_grid(n / 4, 256), 256, 0, (cudaStream_t)stream>>>((const float4*)src,
                                                                            (uint2*)dst, n / 4);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// a[i] *= b[i]  (used to keep the structurally-zero taps of the space-to-depth stem filter at zero)
extern "C" int drn_mul_inplace_f32(float* a, const float* b, long long n, void* stream) {
  DRN_CHECK_CUDA(launch_pdl(mul_inplace_kernel, dim3(ew_grid(n, 256)), dim3(256), 0, (cudaStream_t)stream,
                            a, b, n));
  return 0;
}

// Flipped/transposed dgrad banks of ALL layers in one launch. `table` is a device array of
// n_layers entries {int64 src_off, int64 dst_off, int32 K, RS, C, tile_begin} (32 bytes each) with
// tile_begin the running sum of RS * ceil(K/32) * ceil(C/32); total_tiles is the grand total.
extern "C" int drn_weight_prep_dgrad_batched(const void* w_arena, void* dgrad_arena, const void* table,
                                             int n_layers, int total_tiles, void* stream) {
  static_assert(sizeof(PrepEntry) == 32, "PrepEntry layout is part of the ABI");
  DRN_REQUIRE(n_layers > 0 && total_tiles > 0, "weight_prep_dgrad_batched: empty table");
  DRN_CHECK_CUDA(launch_pdl(weight_prep_dgrad_batched_kernel, dim3(total_tiles), dim3(dim3(32, 8)), 0, (cudaStream_t)stream,
                            (const __nv_bfloat16*)w_arena, (__nv_bfloat16*)dgrad_arena, (const PrepEntry*)table, n_layers));
  return 0;
}

// Four phase banks (9*K*C elements, phases (0,0) (0,1) (1,0) (1,1) back to back) for drn_conv_dgrad_s2.
extern "C" int drn_weight_prep_dgrad_s2(const void* w_krsc, void* banks, int K, int C, void* stream) {
  const long long total = 9LL * K * C;
  DRN_CHECK_CUDA(launch_pdl(weight_prep_dgrad_s2_kernel, dim3(ew_grid(total, 256)), dim3(256), 0,
                            (cudaStream_t)stream, (const __nv_bfloat16*)w_krsc, (__nv_bfloat16*)banks, K, C));
  return 0;
}
