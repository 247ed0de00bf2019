// On-device input pipelines (the reference runs these on the CPU inside tf.data / queue runners).
//
// drn_cifar_augment   CIFAR binary records (uint8, CHW) -> zero-pad, random crop, random flip,
//                     per-image standardisation -> NHWC bf16 with channels padded to 16
//                     (reference resnet_cifar_main.py:157-200; legacy cifar_input.py:66-75).   K21, K13
// drn_imagenet_prep   uint8 HWC images -> crop 224x224 at a given offset, flip, scale to [0,1],
//                     subtract the VGG channel means / 255 (vgg_preprocessing.py:37-39,196-226,
//                     284-333) -> "space-to-depth" NHWC bf16 [B,112,112,16] that turns the 7x7/2
//                     stem into a 4x4/1 convolution for the implicit-GEMM kernel.              K22
// drn_resize_bilinear uint8 HWC bilinear resize with TF1 align_corners=False sampling
//                     (vgg_preprocessing.py:259-281).                                           K22
// Randomness is Philox4x32-10 keyed by (seed, step) and counted by the global image index, so a
// CPU oracle (data/preprocess.py) reproduces every crop and flip bit for bit.
#include "common.cuh"
#include "host.cuh"
#include "../include/drn_b200.h"

namespace drn {
namespace {

__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

// one block per image
__global__ void cifar_augment_kernel(const uint8_t* __restrict__ records, int record_stride,
                                     int image_offset, __nv_bfloat16* __restrict__ out, int B,
                                     int pad, int training, uint32_t seed_lo, uint32_t seed_hi,
                                     uint32_t step, long long index_base, int cpad) {
  constexpr int HW = 32, CH = 3, NPIX = HW * HW;
  __shared__ float s_img[CH * NPIX];
  __shared__ float s_red[2][8];
  const int b = blockIdx.x;
  const uint8_t* src = records + (long long)b * record_stride + image_offset;
  int ox = pad, oy = pad, flip = 0;
  if (training) {
    const unsigned long long idx = (unsigned long long)(index_base + b);
    const uint4 r = philox4x32_10(make_uint4((uint32_t)idx, (uint32_t)(idx >> 32), step, 0u),
                                  make_uint2(seed_lo, seed_hi));
    oy = (int)(r.x % (uint32_t)(2 * pad + 1));
    ox = (int)(r.y % (uint32_t)(2 * pad + 1));
    flip = (int)(r.z & 1u);
  }
  float s1 = 0.f, s2 = 0.f;
  for (int i = threadIdx.x; i < CH * NPIX; i += blockDim.x) {
    const int c = i / NPIX, rem = i % NPIX, y = rem / HW, x = rem % HW;
    const int xs = flip ? (HW - 1 - x) : x;  // flip acts on the cropped image
    const int sy = y + oy - pad, sx = xs + ox - pad;
    float v = 0.f;
    if (sy >= 0 && sy < HW && sx >= 0 && sx < HW) v = (float)src[c * NPIX + sy * HW + sx];
    s_img[i] = v;
    s1 += v;
    s2 = fmaf(v, v, s2);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    s2 += __shfl_xor_sync(0xffffffffu, s2, o);
  }
  if ((threadIdx.x & 31) == 0) {
    s_red[0][threadIdx.x >> 5] = s1;
    s_red[1][threadIdx.x >> 5] = s2;
  }
  __syncthreads();
  float t1 = 0.f, t2 = 0.f;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
    t1 += s_red[0][w];
    t2 += s_red[1][w];
  }
  const float n = (float)(CH * NPIX);
  const float mean = t1 / n;
  const float var = fmaxf(t2 / n - mean * mean, 0.f);
  const float adj = fmaxf(sqrtf(var), rsqrtf(n));  // tf.image.per_image_standardization
  const float inv = 1.f / adj;
  __nv_bfloat16* dst = out + (long long)b * NPIX * cpad;
  for (int i = threadIdx.x; i < NPIX * cpad; i += blockDim.x) {
    const int c = i % cpad, px = i / cpad;
    const float v = c < CH ? (s_img[c * NPIX + px] - mean) * inv : 0.f;
    dst[i] = __float2bfloat16(v);
  }
}

// one thread per (image, s2d pixel); writes 16 channels = (dy*2+dx)*3 + c, 12..15 = 0
__global__ void imagenet_prep_kernel(const uint8_t* __restrict__ src, long long image_stride,
                                     int src_h, int src_w, const int* __restrict__ crop_yx_flip,
                                     __nv_bfloat16* __restrict__ out, int B, int crop, float mean_r,
                                     float mean_g, float mean_b, int out_pitch, int out_xoff) {
  const int half = crop / 2;
  const long long total = (long long)B * half * half;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int j = (int)(i % half);
    const long long t = i / half;
    const int ii = (int)(t % half);
    const int b = (int)(t / half);
    int oy = 0, ox = 0, flip = 0;
    if (crop_yx_flip != nullptr) {
      oy = crop_yx_flip[b * 3 + 0];
      ox = crop_yx_flip[b * 3 + 1];
      flip = crop_yx_flip[b * 3 + 2];
    }
    const uint8_t* img = src + (long long)b * image_stride;
    const float means[3] = {mean_r, mean_g, mean_b};
    uint32_t packed[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) packed[k] = 0u;
    float vals[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) vals[k] = 0.f;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int y = 2 * ii + dy, x = 2 * j + dx;
        const int xs = flip ? (crop - 1 - x) : x;
        const int sy = y + oy, sx = xs + ox;
        const bool in = sy >= 0 && sy < src_h && sx >= 0 && sx < src_w;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const float p = in ? (float)img[((long long)sy * src_w + sx) * 3 + c] : 0.f;
          vals[(dy * 2 + dx) * 3 + c] = p * (1.f / 255.f) - means[c];
        }
      }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) packed[k] = pack_bf16x2(vals[2 * k], vals[2 * k + 1]);
    // rows of the output may be wider than crop/2 pixels (zero borders owned by the caller)
    uint4* dst = reinterpret_cast<uint4*>(out + (((long long)b * half + ii) * out_pitch + j + out_xoff) * 16);
    dst[0] = make_uint4(packed[0], packed[1], packed[2], packed[3]);
    dst[1] = make_uint4(packed[4], packed[5], packed[6], packed[7]);
  }
}

// TF1 resize_bilinear(align_corners=False): src = dst * (in/out), no half-pixel offset.
__global__ void resize_bilinear_kernel(const uint8_t* __restrict__ src, int sh, int sw,
                                       float* __restrict__ dst, int dh, int dw, float scale_y,
                                       float scale_x) {
  const int total = dh * dw;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int x = i % dw, y = i / dw;
    const float fy = y * scale_y, fx = x * scale_x;
    const int y0 = (int)floorf(fy), x0 = (int)floorf(fx);
    const int y1 = min(y0 + 1, sh - 1), x1 = min(x0 + 1, sw - 1);
    const float ly = fy - y0, lx = fx - x0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float tl = src[(y0 * sw + x0) * 3 + c], tr = src[(y0 * sw + x1) * 3 + c];
      const float bl = src[(y1 * sw + x0) * 3 + c], br = src[(y1 * sw + x1) * 3 + c];
      const float top = tl + (tr - tl) * lx, bot = bl + (br - bl) * lx;
      dst[i * 3 + c] = top + (bot - top) * ly;
    }
  }
}


// Batched "VGG" resize + crop straight from the planar RGB output of a GPU JPEG decoder
// (vgg_preprocessing.py:259-333): every output pixel of the 224x224 crop is sampled from the source
// with TF1 resize_bilinear(align_corners=False) semantics (src = dst * in/out, no half-pixel
// offset, edge clamp), rounded to the nearest grey level.  The resized image is never materialised.
// meta[b] = {src_h, src_w, new_h, new_w, off_y, off_x}; src[b] -> uint8 [3][src_h][src_w].
__global__ void resize_crop_batch_kernel(const uint8_t* const* __restrict__ src,
                                         const int* __restrict__ meta, uint8_t* __restrict__ out,
                                         int B, int crop) {
  const int b = blockIdx.y;
  const int sh = meta[b * 6 + 0], sw = meta[b * 6 + 1], nh = meta[b * 6 + 2], nw = meta[b * 6 + 3];
  const int oy = meta[b * 6 + 4], ox = meta[b * 6 + 5];
  const uint8_t* img = src[b];
  const float scale_y = (float)sh / (float)nh, scale_x = (float)sw / (float)nw;
  const long long plane = (long long)sh * sw;
  uint8_t* dst = out + (long long)b * crop * crop * 3;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < crop * crop; i += gridDim.x * blockDim.x) {
    const int x = i % crop, y = i / crop;
    const float fy = (y + oy) * scale_y, fx = (x + ox) * scale_x;
    const int y0 = min((int)floorf(fy), sh - 1), x0 = min((int)floorf(fx), sw - 1);
    const int y1 = min(y0 + 1, sh - 1), x1 = min(x0 + 1, sw - 1);
    const float ly = fy - (float)y0, lx = fx - (float)x0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const uint8_t* pl = img + c * plane;
      const float tl = pl[(long long)y0 * sw + x0], tr = pl[(long long)y0 * sw + x1];
      const float bl = pl[(long long)y1 * sw + x0], br = pl[(long long)y1 * sw + x1];
      const float top = tl + (tr - tl) * lx, bot = bl + (br - bl) * lx;
      const float v = top + (bot - top) * ly;
      dst[i * 3 + c] = (uint8_t)fminf(fmaxf(rintf(v), 0.f), 255.f);
    }
  }
}

}  // namespace
}  // namespace drn

using namespace drn;

extern "C" int drn_cifar_augment(const void* records, int record_stride, int image_offset, void* out,
                                 int B, int pad, int training, unsigned long long seed,
                                 unsigned int step, long long index_base, int cpad, void* stream) {
  DRN_REQUIRE(cpad >= 3 && cpad % 8 == 0, "cifar_augment: cpad must be a multiple of 8");
  DRN_REQUIRE(pad >= 0 && pad <= 16, "cifar_augment: pad out of range");
  cifar_augment_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(
      (const uint8_t*)records, record_stride, image_offset, (__nv_bfloat16*)out, B, pad, training,
      (uint32_t)seed, (uint32_t)(seed >> 32), step, index_base, cpad);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int drn_imagenet_prep(const void* src, long long image_stride, int src_h, int src_w,
                                 const int* crop_yx_flip, void* out, int B, int crop, float mean_r,
                                 float mean_g, float mean_b, int out_pitch, int out_xoff, void* stream) {
  DRN_REQUIRE(crop % 2 == 0, "imagenet_prep: crop size must be even");
  if (out_pitch <= 0) out_pitch = crop / 2;
  DRN_REQUIRE(out_xoff >= 0 && out_xoff + crop / 2 <= out_pitch, "imagenet_prep: bad output pitch/offset");
  const long long total = (long long)B * (crop / 2) * (crop / 2);
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  imagenet_prep_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(
      (const uint8_t*)src, image_stride, src_h, src_w, crop_yx_flip, (__nv_bfloat16*)out, B, crop,
      mean_r, mean_g, mean_b, out_pitch, out_xoff);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int drn_resize_bilinear(const void* src, int sh, int sw, float* dst, int dh, int dw,
                                   void* stream) {
  const int total = dh * dw;
  resize_bilinear_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
      (const uint8_t*)src, sh, sw, dst, dh, dw, (float)sh / (float)dh, (float)sw / (float)dw);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// src_ptrs: DEVICE array of B pointers to planar uint8 [3][h][w] images; meta: DEVICE int[B][6]
// {src_h, src_w, new_h, new_w, off_y, off_x}; out: uint8 [B][crop][crop][3] (what drn_imagenet_prep reads).
extern "C" int drn_resize_crop_batch(const void* src_ptrs, const int* meta, void* out, int B, int crop,
                                     void* stream) {
  DRN_REQUIRE(B > 0 && crop > 0, "resize_crop_batch: empty batch");
  dim3 grid((crop * crop + 255) / 256 > 64 ? 64 : (crop * crop + 255) / 256, B);
  resize_crop_batch_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const uint8_t* const*)src_ptrs, meta,
                                                                   (uint8_t*)out, B, crop);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}
