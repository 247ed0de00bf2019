// Pooling kernels (NHWC bf16, 8 channels per thread).
//   - 3x3/stride-2 max pool with TF "SAME" padding (reference resnet_model_official.py:314-316;
//     112 -> 56 pads 0 before / 1 after) + its backward through a stored arg-max   (op K11)
//   - global average pool fused with the final BatchNorm+ReLU (resnet_model_official.py:266-271,
//     334-339)                                                                     (ops K8/K9/K12)
//     (its backward is drn_bn_bwd_mask_reduce with addend_kind = 3).
#include "common.cuh"
#include "host.cuh"
#include "../include/drn_b200.h"

namespace drn {
namespace {

struct alignas(16) BF8 {
  uint32_t v[4];
};
struct alignas(8) U8x8 {
  uint8_t v[8];
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

__global__ void maxpool_fwd_kernel(const BF8* __restrict__ x, BF8* __restrict__ out,
                                   U8x8* __restrict__ argmax, int N, int H, int W, int c8, int P,
                                   int Q, int ksize, int stride, int pad_lo) {
  pdl_trigger();
  pdl_wait();
  const long long total = (long long)N * P * Q * c8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % c8);
    long long t = i / c8;
    const int q = (int)(t % Q);
    t /= Q;
    const int p = (int)(t % P);
    const long long n = t / P;
    float best[8];
    uint8_t arg[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      best[k] = -INFINITY;
      arg[k] = 0xff;
    }
    for (int r = 0; r < ksize; ++r) {
      const int h = p * stride - pad_lo + r;
      if (h < 0 || h >= H) continue;
      for (int s = 0; s < ksize; ++s) {
        const int w = q * stride - pad_lo + s;
        if (w < 0 || w >= W) continue;
        float f[8];
        unpack8(x[((n * H + h) * W + w) * c8 + cg], f);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          if (f[k] > best[k]) {  // strict: first maximum wins, like cuDNN/TF
            best[k] = f[k];
            arg[k] = (uint8_t)(r * ksize + s);
          }
        }
      }
    }
    out[i] = pack8(best);
    U8x8 a;
#pragma unroll
    for (int k = 0; k < 8; ++k) a.v[k] = arg[k];
    argmax[i] = a;
  }
}

// 3x3 / stride 2 specialisation (the only pooling the networks use): the nine loads are issued
// together (out-of-range taps read a clamped address and are masked afterwards) so that each thread
// has 144 bytes in flight instead of 16.
__global__ void maxpool3x3s2_fwd_kernel(const BF8* __restrict__ x, BF8* __restrict__ out,
                                        U8x8* __restrict__ argmax, int N, int H, int W, int c8, int P,
                                        int Q, int pad_lo) {
  pdl_trigger();
  pdl_wait();
  const long long total = (long long)N * P * Q * c8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    // 32-bit index arithmetic (launcher checks the element count): 64-bit div/mod dominates otherwise
    const unsigned iu = (unsigned)i;
    const int cg = (int)(iu % (unsigned)c8);
    unsigned t = iu / (unsigned)c8;
    const int q = (int)(t % (unsigned)Q);
    t /= (unsigned)Q;
    const int p = (int)(t % (unsigned)P);
    const long long n = t / (unsigned)P;
    const int h0 = p * 2 - pad_lo, w0 = q * 2 - pad_lo;
    BF8 v[9];
    bool ok[9];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int h = h0 + r, w = w0 + s;
        ok[r * 3 + s] = h >= 0 && h < H && w >= 0 && w < W;
        const int hc = min(max(h, 0), H - 1), wc = min(max(w, 0), W - 1);
        v[r * 3 + s] = x[((n * H + hc) * W + wc) * c8 + cg];
      }
    }
    float best[8];
    uint32_t arg[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      best[k] = -INFINITY;
      arg[k] = 0xff;
    }
#pragma unroll
    for (int j = 0; j < 9; ++j) {
      float f[8];
      unpack8(v[j], f);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (ok[j] && f[k] > best[k]) {
          best[k] = f[k];
          arg[k] = j;
        }
      }
    }
    out[i] = pack8(best);
    U8x8 a;
#pragma unroll
    for (int k = 0; k < 8; ++k) a.v[k] = (uint8_t)arg[k];
    argmax[i] = a;
  }
}

// gather formulation: each input pixel sums the gradients of the windows that chose it
__global__ void maxpool_bwd_kernel(const BF8* __restrict__ dy, const U8x8* __restrict__ argmax,
                                   BF8* __restrict__ dx, int N, int H, int W, int c8, int P, int Q,
                                   int ksize, int stride, int pad_lo) {
  pdl_trigger();
  pdl_wait();
  const long long total = (long long)N * H * W * c8;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int cg = (int)(i % c8);
    long long t = i / c8;
    const int w = (int)(t % W);
    t /= W;
    const int h = (int)(t % H);
    const long long n = t / H;
    float g[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) g[k] = 0.f;
    // windows p with p*stride <= h+pad_lo <= p*stride + ksize-1 (at most ceil(ksize/stride)^2 of them)
    const int hh = h + pad_lo, ww = w + pad_lo;
    int p_lo = (hh - (ksize - 1) + stride - 1) / stride, p_hi = hh / stride;
    int q_lo = (ww - (ksize - 1) + stride - 1) / stride, q_hi = ww / stride;
    if (hh - (ksize - 1) < 0) p_lo = 0;
    if (ww - (ksize - 1) < 0) q_lo = 0;
    if (p_hi > P - 1) p_hi = P - 1;
    if (q_hi > Q - 1) q_hi = Q - 1;
    for (int p = p_lo; p <= p_hi; ++p) {
      const int r = hh - p * stride;
      for (int q = q_lo; q <= q_hi; ++q) {
        const int s2 = ww - q * stride;
        const long long o = ((n * P + p) * Q + q) * c8 + cg;
        const U8x8 a = argmax[o];
        float d[8];
        unpack8(dy[o], d);
        const uint8_t me = (uint8_t)(r * ksize + s2);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (a.v[k] == me) g[k] += d[k];
      }
    }
    dx[i] = pack8(g);
  }
}

// 3x3 / stride 2 specialisation: one thread per 2x2 input patch and 8 channels.  The patch
// (rows 2i-pad+{0,1}, columns 2j-pad+{0,1}) is touched by exactly the four windows (i|i-1, j|j-1),
// so each window's gradient and arg-max are loaded once per patch instead of once per pixel.
__global__ void maxpool3x3s2_bwd_kernel(const BF8* __restrict__ dy, const U8x8* __restrict__ argmax,
                                        BF8* __restrict__ dx, int N, int H, int W, int c8, int P,
                                        int Q, int pad_lo, int PI, int QI) {
  pdl_trigger();
  pdl_wait();
  const long long total = (long long)N * PI * QI * c8;
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const unsigned iu = (unsigned)idx;
    const int cg = (int)(iu % (unsigned)c8);
    unsigned t = iu / (unsigned)c8;
    const int j = (int)(t % (unsigned)QI);
    t /= (unsigned)QI;
    const int i = (int)(t % (unsigned)PI);
    const long long n = t / (unsigned)PI;
    float g[4][8];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int k = 0; k < 8; ++k) g[a][k] = 0.f;
    // window (i - di, j - dj) reaches the patch through taps r in {2*di + (0|1)}, s likewise
#pragma unroll
    for (int di = 0; di < 2; ++di) {
#pragma unroll
      for (int dj = 0; dj < 2; ++dj) {
        const int p = i - di, q = j - dj;
        if (p < 0 || p >= P || q < 0 || q >= Q) continue;
        const long long o = ((n * P + p) * Q + q) * c8 + cg;
        const U8x8 a = argmax[o];
        float d[8];
        unpack8(dy[o], d);
#pragma unroll
        for (int er = 0; er < 2 - di; ++er) {
#pragma unroll
          for (int es = 0; es < 2 - dj; ++es) {
            const uint8_t me = (uint8_t)((2 * di + er) * 3 + (2 * dj + es));
#pragma unroll
            for (int k = 0; k < 8; ++k)
              if (a.v[k] == me) g[er * 2 + es][k] += d[k];
          }
        }
      }
    }
#pragma unroll
    for (int er = 0; er < 2; ++er) {
#pragma unroll
      for (int es = 0; es < 2; ++es) {
        const int h = 2 * i - pad_lo + er, w = 2 * j - pad_lo + es;
        if (h >= 0 && h < H && w >= 0 && w < W) dx[((n * H + h) * W + w) * c8 + cg] = pack8(g[er * 2 + es]);
      }
    }
  }
}

// out[n][c] = mean_{hw} relu(y*scale + shift)      (one block per image)
__global__ void bnrelu_avgpool_kernel(const BF8* __restrict__ y,
                                      const float2* __restrict__ scale_shift, BF8* __restrict__ out,
                                      int HW, int c8, float inv_hw) {
  pdl_trigger();
  pdl_wait();
  __shared__ float s_acc[256 * 8];
  const int n = blockIdx.x;
  const int cg = threadIdx.x % c8;
  const int pl = threadIdx.x / c8;
  const int lanes = blockDim.x / c8;
  float2 ss[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) ss[k] = scale_shift[cg * 8 + k];
  float acc[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] = 0.f;
  for (int px = pl; px < HW; px += lanes) {
    float f[8];
    unpack8(y[((long long)n * HW + px) * c8 + cg], f);
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] += fmaxf(fmaf(f[k], ss[k].x, ss[k].y), 0.f);
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) s_acc[threadIdx.x * 8 + k] = acc[k];
  __syncthreads();
  if (pl == 0) {
    float r[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = 0.f;
    for (int l = 0; l < lanes; ++l)
#pragma unroll
      for (int k = 0; k < 8; ++k) r[k] += s_acc[(l * c8 + cg) * 8 + k];
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] *= inv_hw;
    out[(long long)n * c8 + cg] = pack8(r);
  }
}

}  // namespace
}  // namespace drn

using namespace drn;

static int grid_for(long long total) {
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

extern "C" int drn_maxpool_fwd(const void* x, void* out, void* argmax, int N, int H, int W, int C,
                               int ksize, int stride, int pad_lo, int P, int Q, void* stream) {
  DRN_REQUIRE(C % 8 == 0, "maxpool: C must be a multiple of 8");
  DRN_REQUIRE(ksize * ksize < 255, "maxpool: window too large");
  const long long total = (long long)N * P * Q * (C / 8);
  if (ksize == 3 && stride == 2 && total < (1LL << 31)) {
    DRN_CHECK_CUDA(launch_pdl(maxpool3x3s2_fwd_kernel, dim3(grid_for(total)), dim3(256), 0, (cudaStream_t)stream,
                            (const BF8*)x, (BF8*)out, (U8x8*)argmax, N, H, W, C / 8, P, Q, pad_lo));
    return 0;
  }
  DRN_CHECK_CUDA(launch_pdl(maxpool_fwd_kernel, dim3(grid_for(total)), dim3(256), 0, (cudaStream_t)stream,
                            (const BF8*)x, (BF8*)out, (U8x8*)argmax, N, H, W, C / 8, P, Q, ksize, stride, pad_lo));
  return 0;
}

extern "C" int drn_maxpool_bwd(const void* dy, const void* argmax, void* dx, int N, int H, int W,
                               int C, int ksize, int stride, int pad_lo, int P, int Q,
                               void* stream) {
  DRN_REQUIRE(C % 8 == 0, "maxpool: C must be a multiple of 8");
  if (ksize == 3 && stride == 2 && pad_lo >= 0 && pad_lo <= 1) {
    const int PI = (H - 1 + pad_lo) / 2 + 1, QI = (W - 1 + pad_lo) / 2 + 1;
    const long long patches = (long long)N * PI * QI * (C / 8);
    DRN_REQUIRE(patches < (1LL << 31), "maxpool_bwd: tensor too large for the 3x3/2 specialisation");
    DRN_CHECK_CUDA(launch_pdl(maxpool3x3s2_bwd_kernel, dim3(grid_for(patches)), dim3(256), 0, (cudaStream_t)stream,
                            (const BF8*)dy, (const U8x8*)argmax, (BF8*)dx, N, H, W, C / 8, P, Q, pad_lo, PI, QI));
    return 0;
  }
  const long long total = (long long)N * H * W * (C / 8);
  DRN_CHECK_CUDA(launch_pdl(maxpool_bwd_kernel, dim3(grid_for(total)), dim3(256), 0, (cudaStream_t)stream,
                            (const BF8*)dy, (const U8x8*)argmax, (BF8*)dx, N, H, W, C / 8, P, Q, ksize, stride, pad_lo));
  return 0;
}

extern "C" int drn_bnrelu_avgpool(const void* y, const void* scale_shift, void* out, int N, int HW,
                                  int C, void* stream) {
  const int c8 = C / 8;
  DRN_REQUIRE(C % 8 == 0 && c8 <= 256 && (c8 & (c8 - 1)) == 0,
              "bnrelu_avgpool: C=%d must be 8 * a power of two <= 2048", C);
  DRN_CHECK_CUDA(launch_pdl(bnrelu_avgpool_kernel, dim3(N), dim3(256), 0, (cudaStream_t)stream,
                            (const BF8*)y, (const float2*)scale_shift, (BF8*)out, HW, c8, 1.f / (float)HW));
  return 0;
}
