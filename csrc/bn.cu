// BatchNorm (+ReLU) kernels around the convolution epilogues. NHWC bf16 activations, fp32 statistics.
//
// Forward (training):  the producing convolution accumulates per-channel sum / sum-of-squares
//   (conv_igemm.cu) -> drn_bn_finalize turns them into (scale, shift), saves (mean, invstd) and
//   updates the moving averages -> drn_bn_apply writes a = relu(y*scale + shift).
// Backward:  dm = da * [a > 0] and the reductions s1 = sum dm, s2 = sum dm*y come from the dgrad
//   epilogue (or drn_bn_bwd_mask_reduce) -> drn_bn_bwd_finalize emits dgamma/dbeta and the three
//   per-channel coefficients of  dy = A*dm + B*y + C  -> drn_bn_bwd_apply (+ optional addend,
//   the identity-shortcut gradient).
//
// Numerics follow tf.layers.batch_normalization(fused=True) as used by the reference
// (resnet_model_official.py:41-50): biased variance to normalise, unbiased variance into the
// moving average, decay 0.997, eps 1e-5.  Ops K5-K9.
#include <stdlib.h>

#include "common.cuh"
#include "host.cuh"
#include "../include/drn_b200.h"

namespace drn {
namespace {

constexpr int kThreads = 256;

struct alignas(16) BF8 {
  uint32_t v[4];
};
__device__ __forceinline__ void unpack8(const BF8& p, float (&f)[8]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = bf16lo(p.v[i]);
    f[2 * i + 1] = bf16hi(p.v[i]);
  }
}
__device__ __forceinline__ BF8 pack8(const float (&f)[8]) {
  BF8 p;
#pragma unroll
  for (int i = 0; i < 4; ++i) p.v[i] = pack_bf16x2(f[2 * i], f[2 * i + 1]);
  return p;
}

// ------------------------------------------------------------------------------------------
__global__ void bn_finalize_kernel(const float* __restrict__ stats, int stats_ld, float inv_count,
                                   float unbias, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float* __restrict__ run_mean,
                                   float* __restrict__ run_var, float one_minus_decay, float eps,
                                   float2* __restrict__ scale_shift, float2* __restrict__ mean_invstd,
                                   int C) {
  pdl_trigger();
  pdl_wait();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float mean = stats[c] * inv_count;
  float var = fmaf(-mean, mean, stats[stats_ld + c] * inv_count);
  var = fmaxf(var, 0.f);
  const float invstd = rsqrtf(var + eps);
  const float sc = gamma[c] * invstd;
  scale_shift[c] = make_float2(sc, fmaf(-mean, sc, beta[c]));
  mean_invstd[c] = make_float2(mean, invstd);
  if (run_mean != nullptr) {
    // moving -= (moving - batch) * (1 - decay)   (assign_moving_average, zero_debias=False)
    run_mean[c] -= (run_mean[c] - mean) * one_minus_decay;
    run_var[c] -= (run_var[c] - var * unbias) * one_minus_decay;
  }
}

__global__ void bn_inference_params_kernel(const float* __restrict__ gamma,
                                           const float* __restrict__ beta,
                                           const float* __restrict__ run_mean,
                                           const float* __restrict__ run_var, float eps,
                                           float2* __restrict__ scale_shift, int C) {
  pdl_trigger();
  pdl_wait();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float sc = gamma[c] * rsqrtf(run_var[c] + eps);
  scale_shift[c] = make_float2(sc, fmaf(-run_mean[c], sc, beta[c]));
}

// a = [relu](y*scale + shift [+ addend]); 8 channels per thread.
// The grid stride is a multiple of c8 (a power of two <= 256), so a thread always works on the
// same 8 channels and keeps their coefficients in registers.
__global__ void bn_apply_kernel(const BF8* __restrict__ y, const float2* __restrict__ scale_shift,
                                const BF8* __restrict__ addend, BF8* __restrict__ out,
                                long long n_vec, int c8, int relu) {
  pdl_trigger();
  pdl_wait();
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int cg = (int)(i % c8);
  float2 ss[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) ss[k] = __ldg(scale_shift + cg * 8 + k);
  for (; i + stride < n_vec; i += 2 * stride) {
    const BF8 y0 = y[i], y1 = y[i + stride];
    BF8 a0 = y0, a1 = y1;
    if (addend != nullptr) {
      a0 = addend[i];
      a1 = addend[i + stride];
    }
    float f0[8], f1[8];
    unpack8(y0, f0);
    unpack8(y1, f1);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      f0[k] = fmaf(f0[k], ss[k].x, ss[k].y);
      f1[k] = fmaf(f1[k], ss[k].x, ss[k].y);
    }
    if (addend != nullptr) {
      float t0[8], t1[8];
      unpack8(a0, t0);
      unpack8(a1, t1);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        f0[k] += t0[k];
        f1[k] += t1[k];
      }
    }
    if (relu) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        f0[k] = fmaxf(f0[k], 0.f);
        f1[k] = fmaxf(f1[k], 0.f);
      }
    }
    out[i] = pack8(f0);
    out[i + stride] = pack8(f1);
  }
  for (; i < n_vec; i += stride) {
    float f[8];
    unpack8(y[i], f);
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = fmaf(f[k], ss[k].x, ss[k].y);
    if (addend != nullptr) {
      float ad[8];
      unpack8(addend[i], ad);
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] += ad[k];
    }
    if (relu) {
#pragma unroll
      for (int k = 0; k < 8; ++k) f[k] = fmaxf(f[k], 0.f);
    }
    out[i] = pack8(f);
  }
}

// per-channel sum and sum of squares of a [M][C] tensor (standalone producer of BN statistics);
// with a second tensor y2: (sum y, sum y*y2) - the two reductions of a BatchNorm backward from the
// masked gradient and the BatchNorm input.  det: ticket of the deterministic mode (common.cuh) or nullptr.
__global__ void bn_stats_kernel(const BF8* __restrict__ y, const BF8* __restrict__ y2, float* __restrict__ stats,
                                int stats_ld, long long M, int c8, unsigned int* det) {
  pdl_trigger();
  pdl_wait();
  __shared__ float s_acc[kThreads * 16];
  const int cg = threadIdx.x % c8;
  const int pl = threadIdx.x / c8;
  const int lanes = kThreads / c8;
  float s1[8], s2[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s1[k] = s2[k] = 0.f;
  for (long long m = (long long)blockIdx.x * lanes + pl; m < M; m += (long long)gridDim.x * lanes) {
    float f[8], f2[8];
    unpack8(y[m * c8 + cg], f);
    if (y2 != nullptr) unpack8(y2[m * c8 + cg], f2);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      s1[k] += f[k];
      s2[k] = fmaf(f[k], y2 != nullptr ? f2[k] : f[k], s2[k]);
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    s_acc[threadIdx.x * 16 + k] = s1[k];
    s_acc[threadIdx.x * 16 + 8 + k] = s2[k];
  }
  __syncthreads();
  det_turn_begin(det, blockIdx.x);
  // thread t < c8*16 reduces one (channel, which) pair over the pixel lanes
  for (int t = threadIdx.x; t < c8 * 16; t += blockDim.x) {
    const int g = t / 16, k = t % 16;
    float a = 0.f;
    for (int l = 0; l < lanes; ++l) a += s_acc[(l * c8 + g) * 16 + k];
    const int ch = g * 8 + (k & 7);
    atomicAdd(stats + (k >> 3) * stats_ld + ch, a);
  }
  det_turn_end(det, blockIdx.x + 1 == gridDim.x ? 0u : blockIdx.x + 1);
}

// dm = (da [+ addend]) * [y*scale+shift > 0];  stats += (sum dm, sum dm*y).
// addend may be a stride-2 "compact" tensor [N, H/2.., W/2.., C] that is added at even pixels
// (the gradient of a 1x1/stride-2 projection shortcut), or a per-image vector [N][C] scaled by
// `bcast_scale` (the gradient of the global average pool).
template <int KIND>
__global__ void bn_bwd_mask_reduce_kernel(const BF8* __restrict__ da, const BF8* __restrict__ addend,
                                          int H, int W, int Ha, int Wa,
                                          float bcast_scale, const BF8* __restrict__ y,
                                          const float2* __restrict__ scale_shift,
                                          BF8* __restrict__ dm, float* __restrict__ stats,
                                          int stats_ld, long long M, int c8, unsigned int* det) {
  pdl_trigger();
  pdl_wait();
  __shared__ float s_acc[kThreads * 16];
  const int cg = threadIdx.x % c8;
  const int pl = threadIdx.x / c8;
  const int lanes = kThreads / c8;
  float2 ss[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) ss[k] = scale_shift[cg * 8 + k];
  float s1[8], s2[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s1[k] = s2[k] = 0.f;

  // raw loads of one pixel's 8 channels: gradient, optional addend, pre-activation
  struct Px {
    BF8 g, a, y;
    bool has_g, has_a;
  };
  auto fetch = [&](long long m) {
    Px p;
    p.has_g = true;
    p.has_a = false;
    if (KIND == 4) {  // da = four parity-phase tensors [4][N][H/2][W/2][C] of a strided dgrad
      // 32-bit index arithmetic (the launcher checks M < 2^31): 64-bit div/mod would cost more
      // instructions than the rest of the pixel
      const unsigned mi = (unsigned)m, w = mi % (unsigned)W, t = mi / (unsigned)W;
      const unsigned h = t % (unsigned)H, n = t / (unsigned)H;
      const long long ph = (long long)((h & 1) * 2 + (w & 1)) * (M >> 2);
      p.g = da[(ph + ((long long)n * (H >> 1) + (h >> 1)) * (W >> 1) + (w >> 1)) * c8 + cg];
    } else if (da != nullptr) {
      p.g = da[m * c8 + cg];
    } else {
      p.has_g = false;
    }
    p.y = y[m * c8 + cg];
    if (KIND == 1) {
      p.a = addend[m * c8 + cg];
      p.has_a = true;
    } else if (KIND == 2) {  // compact stride-2 tensor added at even pixels
      const unsigned mi = (unsigned)m, w = mi % (unsigned)W, t = mi / (unsigned)W;
      const unsigned h = t % (unsigned)H, n = t / (unsigned)H;
      if (((h | w) & 1) == 0 && (int)(h >> 1) < Ha && (int)(w >> 1) < Wa) {
        p.a = addend[(((long long)n * Ha + (h >> 1)) * Wa + (w >> 1)) * c8 + cg];
        p.has_a = true;
      }
    } else if (KIND == 3) {  // per-image vector
      p.a = addend[(long long)((unsigned)m / (unsigned)(H * W)) * c8 + cg];
      p.has_a = true;
    }
    return p;
  };
  auto finish = [&](long long m, const Px& p) {
    float g[8], yy[8];
    if (p.has_g) {
      unpack8(p.g, g);
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = 0.f;
    }
    unpack8(p.y, yy);
    if (p.has_a) {
      float a[8];
      unpack8(p.a, a);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = KIND == 3 ? fmaf(a[k], bcast_scale, g[k]) : g[k] + a[k];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] = fmaf(yy[k], ss[k].x, ss[k].y) > 0.f ? g[k] : 0.f;
    const BF8 packed = pack8(g);
    dm[m * c8 + cg] = packed;
    float gq[8];
    unpack8(packed, gq);  // reduce what was stored (bf16-rounded), like the fused epilogue does
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      s1[k] += gq[k];
      s2[k] = fmaf(gq[k], yy[k], s2[k]);
    }
  };
  const long long step = (long long)gridDim.x * lanes;
  long long m = (long long)blockIdx.x * lanes + pl;
  for (; m + step < M; m += 2 * step) {   // two pixels in flight
    const Px p0 = fetch(m), p1 = fetch(m + step);
    finish(m, p0);
    finish(m + step, p1);
  }
  if (m < M) finish(m, fetch(m));
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    s_acc[threadIdx.x * 16 + k] = s1[k];
    s_acc[threadIdx.x * 16 + 8 + k] = s2[k];
  }
  __syncthreads();
  det_turn_begin(det, blockIdx.x);
  for (int t = threadIdx.x; t < c8 * 16; t += blockDim.x) {
    const int g = t / 16, k = t % 16;
    float a = 0.f;
    for (int l = 0; l < lanes; ++l) a += s_acc[(l * c8 + g) * 16 + k];
    atomicAdd(stats + (k >> 3) * stats_ld + g * 8 + (k & 7), a);
  }
  det_turn_end(det, blockIdx.x + 1 == gridDim.x ? 0u : blockIdx.x + 1);
}

// Post-activation (v1) blocks: out = relu(bn(y) + shortcut).  Backward of the ReLU plus the two
// reductions the following BatchNorm backward needs, for up to two BatchNorms that feed the same
// sum (the main path's last BN and the projection shortcut's BN):
//   dm = g * [out > 0];  stats += (sum dm, sum dm*y);  stats2 += (sum dm, sum dm*y2)   (y2 optional)
__global__ void relu_bwd_reduce_kernel(const BF8* __restrict__ g, const BF8* __restrict__ out,
                                       const BF8* __restrict__ y, const BF8* __restrict__ y2,
                                       BF8* __restrict__ dm, float* __restrict__ stats,
                                       float* __restrict__ stats2, int stats_ld, long long M, int c8,
                                       unsigned int* det) {
  pdl_trigger();
  pdl_wait();
  __shared__ float s_acc[kThreads * 24];
  const int cg = threadIdx.x % c8;
  const int pl = threadIdx.x / c8;
  const int lanes = kThreads / c8;
  float s1[8], s2[8], s3[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s1[k] = s2[k] = s3[k] = 0.f;
  const long long step = (long long)gridDim.x * lanes;
  for (long long m = (long long)blockIdx.x * lanes + pl; m < M; m += step) {
    const long long i = m * c8 + cg;
    const BF8 gv = g[i], ov = out[i], yv = y[i];
    float gf[8], of[8], yf[8];
    unpack8(gv, gf);
    unpack8(ov, of);
    unpack8(yv, yf);
#pragma unroll
    for (int k = 0; k < 8; ++k) gf[k] = of[k] > 0.f ? gf[k] : 0.f;
    const BF8 packed = pack8(gf);
    dm[i] = packed;
    float gq[8];
    unpack8(packed, gq);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      s1[k] += gq[k];
      s2[k] = fmaf(gq[k], yf[k], s2[k]);
    }
    if (y2 != nullptr) {
      float y2f[8];
      unpack8(y2[i], y2f);
#pragma unroll
      for (int k = 0; k < 8; ++k) s3[k] = fmaf(gq[k], y2f[k], s3[k]);
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    s_acc[threadIdx.x * 24 + k] = s1[k];
    s_acc[threadIdx.x * 24 + 8 + k] = s2[k];
    s_acc[threadIdx.x * 24 + 16 + k] = s3[k];
  }
  __syncthreads();
  det_turn_begin(det, blockIdx.x);
  for (int t = threadIdx.x; t < c8 * 24; t += blockDim.x) {
    const int grp = t / 24, k = t % 24;
    float a = 0.f;
    for (int l = 0; l < lanes; ++l) a += s_acc[(l * c8 + grp) * 24 + k];
    const int c = grp * 8 + (k & 7);
    if (k < 8) {
      atomicAdd(stats + c, a);
      if (stats2 != nullptr) atomicAdd(stats2 + c, a);
    } else if (k < 16) {
      atomicAdd(stats + stats_ld + c, a);
    } else if (stats2 != nullptr) {
      atomicAdd(stats2 + stats_ld + c, a);
    }
  }
  det_turn_end(det, blockIdx.x + 1 == gridDim.x ? 0u : blockIdx.x + 1);
}

// dgamma, dbeta and the coefficients of dy = A*dm + B*y + C
__global__ void bn_bwd_finalize_kernel(const float* __restrict__ stats, int stats_ld,
                                       float inv_count, const float* __restrict__ gamma,
                                       const float2* __restrict__ mean_invstd,
                                       float* __restrict__ dgamma, float* __restrict__ dbeta,
                                       float4* __restrict__ coef, int C) {
  pdl_trigger();
  pdl_wait();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float s1 = stats[c], s2 = stats[stats_ld + c];
  const float2 mi = mean_invstd[c];
  const float dg = mi.y * (s2 - mi.x * s1);  // sum dm * xhat
  if (dgamma != nullptr) {
    dgamma[c] += dg;
    dbeta[c] += s1;
  }
  const float a = gamma[c] * mi.y;
  const float b = -a * mi.y * dg * inv_count;
  const float cc = -a * s1 * inv_count - b * mi.x;
  coef[c] = make_float4(a, b, cc, 0.f);
}

__global__ void bn_bwd_apply_kernel(const BF8* __restrict__ dm, const BF8* __restrict__ y,
                                    const float4* __restrict__ coef, const BF8* __restrict__ addend,
                                    BF8* __restrict__ dy, long long n_vec, int c8) {
  pdl_trigger();
  pdl_wait();
  const long long stride = (long long)gridDim.x * blockDim.x;  // multiple of c8: channels stay fixed
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int cg = (int)(i % c8);
  float4 cf[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) cf[k] = __ldg(coef + cg * 8 + k);
  // two independent vectors per iteration: more loads in flight per thread
  for (; i + stride < n_vec; i += 2 * stride) {
    const BF8 d0 = dm[i], d1 = dm[i + stride];
    const BF8 y0 = y[i], y1 = y[i + stride];
    BF8 a0 = d0, a1 = d1;
    if (addend != nullptr) {
      a0 = addend[i];
      a1 = addend[i + stride];
    }
    float g0[8], g1[8], t0[8], t1[8];
    unpack8(d0, g0);
    unpack8(d1, g1);
    unpack8(y0, t0);
    unpack8(y1, t1);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      g0[k] = fmaf(cf[k].x, g0[k], fmaf(cf[k].y, t0[k], cf[k].z));
      g1[k] = fmaf(cf[k].x, g1[k], fmaf(cf[k].y, t1[k], cf[k].z));
    }
    if (addend != nullptr) {
      unpack8(a0, t0);
      unpack8(a1, t1);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        g0[k] += t0[k];
        g1[k] += t1[k];
      }
    }
    dy[i] = pack8(g0);
    dy[i + stride] = pack8(g1);
  }
  for (; i < n_vec; i += stride) {
    float g[8], yy[8];
    unpack8(dm[i], g);
    unpack8(y[i], yy);
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] = fmaf(cf[k].x, g[k], fmaf(cf[k].y, yy[k], cf[k].z));
    if (addend != nullptr) {
      float ad[8];
      unpack8(addend[i], ad);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] += ad[k];
    }
    dy[i] = pack8(g);
  }
}


// ---- fused variants: the per-channel "finalize" arithmetic is redone by every thread for its own
// 8 channels (a few dozen flops) instead of being a separate tiny launch; the threads of block 0
// publish the per-channel results other kernels need (scale/shift, mean/invstd, moving averages,
// dgamma/dbeta). Saves ~100 launches per ResNet-50 step.
__global__ void bn_finalize_apply_kernel(const BF8* __restrict__ y, const float* __restrict__ stats,
                                         int stats_ld, float inv_count, float unbias,
                                         const float* __restrict__ gamma, const float* __restrict__ beta,
                                         float* __restrict__ run_mean, float* __restrict__ run_var,
                                         float one_minus_decay, float eps,
                                         float2* __restrict__ scale_shift, float2* __restrict__ mean_invstd,
                                         BF8* __restrict__ out, long long n_vec, int c8, int relu) {
  pdl_trigger();
  pdl_wait();
  // The LAST block only publishes the per-channel results (one channel per thread, so the
  // read-modify-writes of the moving averages are parallel instead of eight dependent round trips
  // in front of a worker's data loop); all other blocks stream the tensor.
  if (blockIdx.x == gridDim.x - 1) {
    for (int c = threadIdx.x; c < c8 * 8; c += blockDim.x) {
      const float mean = __ldg(stats + c) * inv_count;
      const float var = fmaxf(fmaf(-mean, mean, __ldg(stats + stats_ld + c) * inv_count), 0.f);
      const float invstd = rsqrtf(var + eps);
      const float sc = __ldg(gamma + c) * invstd;
      scale_shift[c] = make_float2(sc, fmaf(-mean, sc, __ldg(beta + c)));
      mean_invstd[c] = make_float2(mean, invstd);
      if (run_mean != nullptr) {
        run_mean[c] -= (run_mean[c] - mean) * one_minus_decay;
        run_var[c] -= (run_var[c] - var * unbias) * one_minus_decay;
      }
    }
    return;
  }
  const long long stride = (long long)(gridDim.x - 1) * blockDim.x;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int cg = (int)(i % c8);
  // first pair of data loads before the statistics round trip (see the backward kernel)
  bool first = i + stride < n_vec;
  BF8 y0, y1;
  if (first) {
    y0 = y[i];
    y1 = y[i + stride];
  }
  float2 ss[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cg * 8 + k;
    const float mean = __ldg(stats + c) * inv_count;
    const float var = fmaxf(fmaf(-mean, mean, __ldg(stats + stats_ld + c) * inv_count), 0.f);
    const float invstd = rsqrtf(var + eps);
    const float sc = __ldg(gamma + c) * invstd;
    ss[k] = make_float2(sc, fmaf(-mean, sc, __ldg(beta + c)));
  }
  for (; i + stride < n_vec; i += 2 * stride) {
    if (!first) {
      y0 = y[i];
      y1 = y[i + stride];
    }
    first = false;
    float f0[8], f1[8];
    unpack8(y0, f0);
    unpack8(y1, f1);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      f0[k] = fmaf(f0[k], ss[k].x, ss[k].y);
      f1[k] = fmaf(f1[k], ss[k].x, ss[k].y);
      if (relu) {
        f0[k] = fmaxf(f0[k], 0.f);
        f1[k] = fmaxf(f1[k], 0.f);
      }
    }
    out[i] = pack8(f0);
    out[i + stride] = pack8(f1);
  }
  for (; i < n_vec; i += stride) {
    float f[8];
    unpack8(y[i], f);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      f[k] = fmaf(f[k], ss[k].x, ss[k].y);
      if (relu) f[k] = fmaxf(f[k], 0.f);
    }
    out[i] = pack8(f);
  }
}

template <int POL>
__device__ __forceinline__ BF8 ld_vec(const BF8* p) {
  if (POL == 0) return *p;
  BF8 r;
  // streaming read: the tensor is dead (or far from its next use) after this kernel
  asm volatile("ld.global.cs.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3])
               : "l"(p));
  return r;
}
template <int POL>
__device__ __forceinline__ void st_vec(BF8* p, const BF8& r) {
  if (POL == 0) {
    *p = r;
    return;
  }
  asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(r.v[0]), "r"(r.v[1]),
               "r"(r.v[2]), "r"(r.v[3])
               : "memory");
}

// U vectors in flight per thread; POL = cache policy of the streams (0 default, 1 streaming)
template <int U, int POL>
__global__ void bn_bwd_finalize_apply_kernel(const BF8* __restrict__ dm, const BF8* __restrict__ y,
                                             const float* __restrict__ stats, int stats_ld,
                                             float inv_count, const float* __restrict__ gamma,
                                             const float2* __restrict__ mean_invstd,
                                             float* __restrict__ dgamma, float* __restrict__ dbeta,
                                             const BF8* __restrict__ addend, BF8* __restrict__ dy,
                                             long long n_vec, int c8) {
  pdl_trigger();
  pdl_wait();
  if (blockIdx.x == gridDim.x - 1) {   // publisher block, see bn_finalize_apply_kernel
    if (dgamma != nullptr) {
      for (int c = threadIdx.x; c < c8 * 8; c += blockDim.x) {
        const float s1 = __ldg(stats + c), s2 = __ldg(stats + stats_ld + c);
        const float2 mi = __ldg(mean_invstd + c);
        dgamma[c] += mi.y * (s2 - mi.x * s1);
        dbeta[c] += s1;
      }
    }
    return;
  }
  const long long stride = (long long)(gridDim.x - 1) * blockDim.x;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int cg = (int)(i % c8);
  // the first batch of data loads does not depend on the coefficients: issue it before the
  // statistics round trip so the two latencies overlap (small tensors are latency-bound)
  const bool has_add = addend != nullptr;
  BF8 d[U], t[U], a[U];
  bool first = i + (U - 1) * stride < n_vec;
  if (first) {
#pragma unroll
    for (int u = 0; u < U; ++u) d[u] = ld_vec<POL>(dm + i + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u) t[u] = ld_vec<0>(y + i + u * stride);
    if (has_add) {
#pragma unroll
      for (int u = 0; u < U; ++u) a[u] = ld_vec<POL>(addend + i + u * stride);
    }
  }
  float ca[8], cb[8], cc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cg * 8 + k;
    const float s1 = __ldg(stats + c), s2 = __ldg(stats + stats_ld + c);
    const float2 mi = __ldg(mean_invstd + c);
    const float dg = mi.y * (s2 - mi.x * s1);
    ca[k] = __ldg(gamma + c) * mi.y;
    cb[k] = -ca[k] * mi.y * dg * inv_count;
    cc[k] = -ca[k] * s1 * inv_count - cb[k] * mi.x;
  }
  for (; i + (U - 1) * stride < n_vec; i += U * stride) {
    if (!first) {
#pragma unroll
      for (int u = 0; u < U; ++u) d[u] = ld_vec<POL>(dm + i + u * stride);
#pragma unroll
      for (int u = 0; u < U; ++u) t[u] = ld_vec<0>(y + i + u * stride);
      if (has_add) {
#pragma unroll
        for (int u = 0; u < U; ++u) a[u] = ld_vec<POL>(addend + i + u * stride);
      }
    }
    first = false;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float g[8], yy[8];
      unpack8(d[u], g);
      unpack8(t[u], yy);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = fmaf(ca[k], g[k], fmaf(cb[k], yy[k], cc[k]));
      if (has_add) {
        unpack8(a[u], yy);
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] += yy[k];
      }
      st_vec<POL>(dy + i + u * stride, pack8(g));
    }
  }
  for (; i < n_vec; i += stride) {
    float g[8], yy[8];
    unpack8(dm[i], g);
    unpack8(y[i], yy);
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] = fmaf(ca[k], g[k], fmaf(cb[k], yy[k], cc[k]));
    if (has_add) {
      float ad[8];
      unpack8(addend[i], ad);
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] += ad[k];
    }
    dy[i] = pack8(g);
  }
}

int check_c(int C, const char* who) {
  const int c8 = C / 8;
  if (C % 8 != 0 || c8 > kThreads || (c8 & (c8 - 1)) != 0)
    return fail("%s: C=%d must be 8 * a power of two <= 2048", who, C);
  return 0;
}

// Grid for the fused finalize+apply kernels: every thread first derives the coefficients of its 8
// channels (a few dozen dependent loads and flops), so small tensors want FEW threads with several
// vectors each rather than a full machine of threads with two.
int fused_grid(long long n_vec) {
  static const int vpt = getenv("DRN_BN_VPT") ? atoi(getenv("DRN_BN_VPT")) : 8;
  static const int bps = getenv("DRN_BN_BPS") ? atoi(getenv("DRN_BN_BPS")) : 8;
  long long blocks = (n_vec + (long long)kThreads * vpt - 1) / ((long long)kThreads * vpt);
  const long long lo = num_sms(), hi = (long long)num_sms() * bps;
  if (blocks < lo) blocks = lo;
  if (blocks > hi) blocks = hi;
  const long long all = (n_vec + kThreads - 1) / kThreads;
  if (blocks > all) blocks = all;
  return (int)(blocks > 0 ? blocks : 1);
}

int ew_grid(long long n_vec) {
  long long blocks = (n_vec + kThreads - 1) / kThreads;
  const long long cap = (long long)num_sms() * 8;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

}  // namespace
}  // namespace drn

using namespace drn;

extern "C" int drn_bn_finalize(const float* stats, int stats_ld, long long count, const float* gamma,
                               const float* beta, float* run_mean, float* run_var, float decay,
                               float eps, void* scale_shift, void* mean_invstd, int C,
                               void* stream) {
  DRN_REQUIRE(count > 0, "bn_finalize: count must be positive");
  const float unbias = count > 1 ? (float)((double)count / (double)(count - 1)) : 1.f;
  DRN_CHECK_CUDA(launch_pdl(bn_finalize_kernel, dim3((C + 127) / 128), dim3(128), 0, (cudaStream_t)stream,
                            stats, stats_ld, (float)(1.0 / (double)count), unbias, gamma, beta, run_mean, run_var,
      1.f - decay, eps, (float2*)scale_shift, (float2*)mean_invstd, C));
  return 0;
}

extern "C" int drn_bn_inference_params(const float* gamma, const float* beta, const float* run_mean,
                                       const float* run_var, float eps, void* scale_shift, int C,
                                       void* stream) {
  DRN_CHECK_CUDA(launch_pdl(bn_inference_params_kernel, dim3((C + 127) / 128), dim3(128), 0, (cudaStream_t)stream,
                            gamma, beta, run_mean, run_var, eps, (float2*)scale_shift, C));
  return 0;
}

extern "C" int drn_bn_apply(const void* y, const void* scale_shift, const void* addend, void* out,
                            long long M, int C, int relu, void* stream) {
  if (int e = check_c(C, "bn_apply")) return e;
  const long long n_vec = M * (C / 8);
  DRN_CHECK_CUDA(launch_pdl(bn_apply_kernel, dim3(ew_grid(n_vec)), dim3(kThreads), 0, (cudaStream_t)stream,
                            (const BF8*)y, (const float2*)scale_shift, (const BF8*)addend, (BF8*)out, n_vec, C / 8, relu));
  return 0;
}

// grid of the reduction kernels: in the deterministic mode one resident block per SM (the blocks
// take turns in index order)
static int reduce_grid(long long M, int C, int per_sm) {
  const int lanes = kThreads / (C / 8);
  long long blocks = (M + lanes - 1) / lanes;
  const long long cap = (long long)num_sms() * (deterministic() ? 1 : per_sm);
  return (int)(blocks > cap ? cap : blocks);
}

extern "C" int drn_bn_stats(const void* y, float* stats, int stats_ld, long long M, int C,
                            void* stream) {
  if (int e = check_c(C, "bn_stats")) return e;
  DRN_CHECK_CUDA(launch_pdl(bn_stats_kernel, dim3(reduce_grid(M, C, 4)), dim3(kThreads), 0, (cudaStream_t)stream,
                            (const BF8*)y, (const BF8*)nullptr, stats, stats_ld, M, C / 8, det_ticket(kDetBnStats)));
  return 0;
}

// stats[0][c] += sum_m a[m][c];  stats[1][c] += sum_m a[m][c] * y[m][c]
extern "C" int drn_bn_pair_reduce(const void* a, const void* y, float* stats, int stats_ld, long long M, int C,
                                  void* stream) {
  if (int e = check_c(C, "bn_pair_reduce")) return e;
  DRN_REQUIRE(a != nullptr && y != nullptr, "bn_pair_reduce: two tensors needed");
  DRN_CHECK_CUDA(launch_pdl(bn_stats_kernel, dim3(reduce_grid(M, C, 4)), dim3(kThreads), 0, (cudaStream_t)stream,
                            (const BF8*)a, (const BF8*)y, stats, stats_ld, M, C / 8, det_ticket(kDetPairReduce)));
  return 0;
}

extern "C" int drn_bn_bwd_mask_reduce(const void* da, const void* addend, int addend_kind, int N,
                                      int H, int W, int Ha, int Wa, float bcast_scale,
                                      const void* y, const void* scale_shift, void* dm, float* stats,
                                      int stats_ld, int C, void* stream) {
  if (int e = check_c(C, "bn_bwd_mask_reduce")) return e;
  DRN_REQUIRE(addend_kind >= 0 && addend_kind <= 4, "bn_bwd_mask_reduce: bad addend_kind");
  DRN_REQUIRE(addend_kind == 0 || addend_kind == 4 || addend != nullptr, "bn_bwd_mask_reduce: addend missing");
  DRN_REQUIRE(addend_kind != 4 || (da != nullptr && H % 2 == 0 && W % 2 == 0),
              "bn_bwd_mask_reduce: phase-compact gradients need even H and W");
  DRN_REQUIRE(da != nullptr || addend_kind != 0, "bn_bwd_mask_reduce: no gradient given");
  const long long M = (long long)N * H * W;
  DRN_REQUIRE(M < (1LL << 31), "bn_bwd_mask_reduce: more than 2^31 pixels");
  const int blocks = reduce_grid(M, C, 8);
  auto kern = addend_kind == 0   ? bn_bwd_mask_reduce_kernel<0>
              : addend_kind == 1 ? bn_bwd_mask_reduce_kernel<1>
              : addend_kind == 2 ? bn_bwd_mask_reduce_kernel<2>
              : addend_kind == 3 ? bn_bwd_mask_reduce_kernel<3>
                                 : bn_bwd_mask_reduce_kernel<4>;
  DRN_CHECK_CUDA(launch_pdl(kern, dim3((int)blocks), dim3(kThreads), 0, (cudaStream_t)stream,
                            (const BF8*)da, (const BF8*)addend, H, W, Ha, Wa, bcast_scale, (const BF8*)y,
                            (const float2*)scale_shift, (BF8*)dm, stats, stats_ld, M, C / 8,
                            det_ticket(kDetMaskReduce)));
  return 0;
}

// v1 (post-activation) residual blocks: backward of out = relu(bn(y) + shortcut), see the kernel.
extern "C" int drn_relu_bwd_reduce(const void* g, const void* out, const void* y, const void* y2,
                                   void* dm, float* stats, float* stats2, int stats_ld, long long M,
                                   int C, void* stream) {
  if (int e = check_c(C, "relu_bwd_reduce")) return e;
  DRN_REQUIRE((y2 == nullptr) == (stats2 == nullptr), "relu_bwd_reduce: y2 and stats2 go together");
  const int blocks = reduce_grid(M, C, 8);
  DRN_CHECK_CUDA(launch_pdl(relu_bwd_reduce_kernel, dim3((int)blocks), dim3(kThreads), 0, (cudaStream_t)stream,
                            (const BF8*)g, (const BF8*)out, (const BF8*)y, (const BF8*)y2, (BF8*)dm, stats,
                            stats2, stats_ld, M, C / 8, det_ticket(kDetReluBwd)));
  return 0;
}

extern "C" int drn_bn_bwd_finalize(const float* stats, int stats_ld, long long count,
                                   const float* gamma, const void* mean_invstd, float* dgamma,
                                   float* dbeta, void* coef, int C, void* stream) {
  DRN_CHECK_CUDA(launch_pdl(bn_bwd_finalize_kernel, dim3((C + 127) / 128), dim3(128), 0, (cudaStream_t)stream,
                            stats, stats_ld, (float)(1.0 / (double)count), gamma, (const float2*)mean_invstd, dgamma,
      dbeta, (float4*)coef, C));
  return 0;
}

extern "C" int drn_bn_bwd_apply(const void* dm, const void* y, const void* coef, const void* addend,
                                void* dy, long long M, int C, void* stream) {
  if (int e = check_c(C, "bn_bwd_apply")) return e;
  const long long n_vec = M * (C / 8);
  DRN_CHECK_CUDA(launch_pdl(bn_bwd_apply_kernel, dim3(ew_grid(n_vec)), dim3(kThreads), 0, (cudaStream_t)stream,
                            (const BF8*)dm, (const BF8*)y, (const float4*)coef, (const BF8*)addend, (BF8*)dy, n_vec,
      C / 8));
  return 0;
}

// bn_finalize + bn_apply in one launch (training forward).
extern "C" int drn_bn_finalize_apply(const void* y, const float* stats, int stats_ld, long long count,
                                     const float* gamma, const float* beta, float* run_mean,
                                     float* run_var, float decay, float eps, void* scale_shift,
                                     void* mean_invstd, void* out, int C, int relu, void* stream) {
  if (int e = check_c(C, "bn_finalize_apply")) return e;
  DRN_REQUIRE(count > 0, "bn_finalize_apply: count must be positive");
  const float unbias = count > 1 ? (float)((double)count / (double)(count - 1)) : 1.f;
  const long long n_vec = count * (C / 8);
  PdlSmall pdl_small(n_vec);
  DRN_CHECK_CUDA(launch_pdl(bn_finalize_apply_kernel, dim3(fused_grid(n_vec) + 1), dim3(kThreads), 0, (cudaStream_t)stream,
                            (const BF8*)y, stats, stats_ld, (float)(1.0 / (double)count), unbias, gamma, beta, run_mean,
      run_var, 1.f - decay, eps, (float2*)scale_shift, (float2*)mean_invstd, (BF8*)out, n_vec, C / 8,
      relu));
  return 0;
}

// bn_bwd_finalize + bn_bwd_apply in one launch.
extern "C" int drn_bn_bwd_finalize_apply(const void* dm, const void* y, const float* stats,
                                         int stats_ld, long long count, const float* gamma,
                                         const void* mean_invstd, float* dgamma, float* dbeta,
                                         const void* addend, void* dy, int C, void* stream) {
  if (int e = check_c(C, "bn_bwd_finalize_apply")) return e;
  const long long n_vec = count * (C / 8);
  static const int variant = getenv("DRN_BN_VARIANT") ? atoi(getenv("DRN_BN_VARIANT")) : 0;
  const int blocks = fused_grid(n_vec) + 1;   // + the publisher block
  PdlSmall pdl_small(n_vec);
  auto kern = variant == 1   ? bn_bwd_finalize_apply_kernel<2, 1>
              : variant == 2 ? bn_bwd_finalize_apply_kernel<4, 0>
              : variant == 3 ? bn_bwd_finalize_apply_kernel<4, 1>
              : variant == 4 ? bn_bwd_finalize_apply_kernel<1, 0>
                             : bn_bwd_finalize_apply_kernel<2, 0>;
  DRN_CHECK_CUDA(launch_pdl(kern, dim3((unsigned)blocks), dim3(kThreads), 0, (cudaStream_t)stream,
                            (const BF8*)dm, (const BF8*)y, stats, stats_ld, (float)(1.0 / (double)count), gamma,
                            (const float2*)mean_invstd, dgamma, dbeta, (const BF8*)addend, (BF8*)dy, n_vec, C / 8));
  return 0;
}
