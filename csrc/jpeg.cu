// Batched JPEG decode on the GPU through nvJPEG (op K23 of SURVEY.md 2.6: the reference decodes with
// libjpeg inside tf.image.decode_jpeg on five CPU threads, resnet_imagenet_main.py:138-168).
//
// One decoder object per process/GPU.  It prefers the hardware JPEG engine
// (NVJPEG_BACKEND_HARDWARE: Huffman decoding + IDCT off the SMs and off the host) and falls back to
// the GPU-hybrid / default backends.  A batch is decoded into caller-provided planar RGB buffers
// ([3][h][w] uint8 per image - exactly what drn_resize_crop_batch consumes), images the hardware
// engine cannot take (progressive, CMYK, unusual subsampling) go through the default backend one by
// one, and whatever still fails is reported per image so that the caller can use a host decoder.
//
// nvJPEG allocates its own device / pinned work buffers: create the decoder at start-up.
#include <nvjpeg.h>

#include <vector>

#include "host.cuh"
#include "../include/drn_b200.h"

namespace {

struct Decoder {
  nvjpegHandle_t hw = nullptr;        // hardware-engine handle (may be null)
  nvjpegJpegState_t hw_state = nullptr;
  nvjpegJpegStream_t stream_probe = nullptr;
  nvjpegHandle_t sw = nullptr;        // GPU-hybrid / default handle (always present)
  nvjpegJpegState_t sw_state = nullptr;
  nvjpegJpegState_t sw_batch_state = nullptr;
  int hw_batch = 0, sw_batch = 0;
  int backend = 0;                    // 2 = hardware engine, 1 = GPU hybrid, 0 = default
};

#define DRN_NVJPEG(expr)                                                                   \
  do {                                                                                     \
    nvjpegStatus_t _s = (expr);                                                            \
    if (_s != NVJPEG_STATUS_SUCCESS)                                                       \
      return ::drn::fail("%s failed with nvjpegStatus %d (%s:%d)", #expr, (int)_s, __FILE__, __LINE__); \
  } while (0)

}  // namespace

extern "C" int drn_jpeg_create(int prefer_hardware, void** out) {
  Decoder* d = new Decoder();
  if (prefer_hardware &&
      nvjpegCreateEx(NVJPEG_BACKEND_HARDWARE, nullptr, nullptr, NVJPEG_FLAGS_DEFAULT, &d->hw) == NVJPEG_STATUS_SUCCESS) {
    if (nvjpegJpegStateCreate(d->hw, &d->hw_state) != NVJPEG_STATUS_SUCCESS ||
        nvjpegJpegStreamCreate(d->hw, &d->stream_probe) != NVJPEG_STATUS_SUCCESS) {
      nvjpegDestroy(d->hw);
      d->hw = nullptr;
    }
  } else {
    d->hw = nullptr;
  }
  nvjpegStatus_t s = nvjpegCreateEx(NVJPEG_BACKEND_GPU_HYBRID, nullptr, nullptr, NVJPEG_FLAGS_DEFAULT, &d->sw);
  d->backend = 1;
  if (s != NVJPEG_STATUS_SUCCESS) {
    s = nvjpegCreateEx(NVJPEG_BACKEND_DEFAULT, nullptr, nullptr, NVJPEG_FLAGS_DEFAULT, &d->sw);
    d->backend = 0;
  }
  if (s != NVJPEG_STATUS_SUCCESS) {
    delete d;
    return drn::fail("jpeg_create: nvjpegCreateEx failed with status %d", (int)s);
  }
  DRN_NVJPEG(nvjpegJpegStateCreate(d->sw, &d->sw_state));
  DRN_NVJPEG(nvjpegJpegStateCreate(d->sw, &d->sw_batch_state));
  if (d->hw != nullptr) d->backend = 2;
  *out = d;
  return 0;
}

// 2 = hardware JPEG engine, 1 = GPU hybrid, 0 = default backend
extern "C" int drn_jpeg_backend(void* handle) { return static_cast<Decoder*>(handle)->backend; }

extern "C" int drn_jpeg_destroy(void* handle) {
  Decoder* d = static_cast<Decoder*>(handle);
  if (d == nullptr) return 0;
  if (d->hw_state) nvjpegJpegStateDestroy(d->hw_state);
  if (d->stream_probe) nvjpegJpegStreamDestroy(d->stream_probe);
  if (d->hw) nvjpegDestroy(d->hw);
  if (d->sw_state) nvjpegJpegStateDestroy(d->sw_state);
  if (d->sw_batch_state) nvjpegJpegStateDestroy(d->sw_batch_state);
  if (d->sw) nvjpegDestroy(d->sw);
  delete d;
  return 0;
}

// width / height of each image from its header (host only); wh = int[B][2]; 0,0 = not a JPEG.
extern "C" int drn_jpeg_info(void* handle, const unsigned char* const* data, const size_t* lengths, int B,
                             int* wh) {
  Decoder* d = static_cast<Decoder*>(handle);
  for (int i = 0; i < B; ++i) {
    int comps = 0, w[NVJPEG_MAX_COMPONENT] = {0}, h[NVJPEG_MAX_COMPONENT] = {0};
    nvjpegChromaSubsampling_t ss;
    if (nvjpegGetImageInfo(d->sw, data[i], lengths[i], &comps, &ss, w, h) == NVJPEG_STATUS_SUCCESS) {
      wh[2 * i] = w[0];
      wh[2 * i + 1] = h[0];
    } else {
      wh[2 * i] = wh[2 * i + 1] = 0;
    }
  }
  return 0;
}

// Decode B images into planar RGB: out[i] = device uint8 [3][h_i][w_i] with (w_i, h_i) from
// drn_jpeg_info.  status[i] = 0 on success, non-zero if the image could not be decoded on the GPU
// (out[i] is then undefined).  The work is enqueued on `stream`; nvJPEG synchronises internally
// where its backends need it.
extern "C" int drn_jpeg_decode_batch(void* handle, const unsigned char* const* data, const size_t* lengths,
                                     int B, void* const* out, const int* wh, int* status, void* stream) {
  Decoder* d = static_cast<Decoder*>(handle);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  std::vector<nvjpegImage_t> img(B);
  for (int i = 0; i < B; ++i) {
    const int w = wh[2 * i], h = wh[2 * i + 1];
    memset(&img[i], 0, sizeof(nvjpegImage_t));
    unsigned char* base = static_cast<unsigned char*>(out[i]);
    for (int c = 0; c < 3; ++c) {
      img[i].channel[c] = base + (size_t)c * w * h;
      img[i].pitch[c] = (size_t)w;
    }
    status[i] = (w > 0 && h > 0) ? -1 : 1;     // -1 = still to do
  }
  // 1. hardware engine: everything it supports, in one batched call
  if (d->hw != nullptr) {
    std::vector<int> idx;
    for (int i = 0; i < B; ++i) {
      if (status[i] != -1) continue;
      int ok = -1;
      if (nvjpegJpegStreamParse(d->hw, data[i], lengths[i], 0, 0, d->stream_probe) == NVJPEG_STATUS_SUCCESS &&
          nvjpegDecodeBatchedSupported(d->hw, d->stream_probe, &ok) == NVJPEG_STATUS_SUCCESS && ok == 0)
        idx.push_back(i);
    }
    if (!idx.empty()) {
      const int n = (int)idx.size();
      std::vector<const unsigned char*> ptrs(n);
      std::vector<size_t> lens(n);
      std::vector<nvjpegImage_t> outs(n);
      for (int k = 0; k < n; ++k) {
        ptrs[k] = data[idx[k]];
        lens[k] = lengths[idx[k]];
        outs[k] = img[idx[k]];
      }
      bool good = true;
      if (d->hw_batch != n) {
        good = nvjpegDecodeBatchedInitialize(d->hw, d->hw_state, n, 1, NVJPEG_OUTPUT_RGB) == NVJPEG_STATUS_SUCCESS;
        d->hw_batch = good ? n : 0;
      }
      if (good && nvjpegDecodeBatched(d->hw, d->hw_state, ptrs.data(), lens.data(), outs.data(), st) ==
                      NVJPEG_STATUS_SUCCESS)
        for (int k = 0; k < n; ++k) status[idx[k]] = 0;
    }
  } else {
    // 1'. no hardware engine: one batched call on the GPU-hybrid backend
    bool good = true;
    bool all = true;
    for (int i = 0; i < B; ++i) all = all && status[i] == -1;
    if (all) {
      if (d->sw_batch != B) {
        good = nvjpegDecodeBatchedInitialize(d->sw, d->sw_batch_state, B, 1, NVJPEG_OUTPUT_RGB) == NVJPEG_STATUS_SUCCESS;
        d->sw_batch = good ? B : 0;
      }
      if (good && nvjpegDecodeBatched(d->sw, d->sw_batch_state, data, lengths, img.data(), st) == NVJPEG_STATUS_SUCCESS)
        for (int i = 0; i < B; ++i) status[i] = 0;
    }
  }
  // 2. the rest one by one on the default path
  for (int i = 0; i < B; ++i) {
    if (status[i] != -1) continue;
    const nvjpegStatus_t s = nvjpegDecode(d->sw, d->sw_state, data[i], lengths[i], NVJPEG_OUTPUT_RGB, &img[i], st);
    status[i] = s == NVJPEG_STATUS_SUCCESS ? 0 : 2;
  }
  return 0;
}
