// Softmax cross-entropy, fused forward + backward + batch precision.
//
// Reference graph: tf.nn.softmax (predictions), tf.nn.softmax_cross_entropy_with_logits +
// reduce_mean (resnet_model.py:77-82) and the train-batch precision
// mean(argmax(predictions) == argmax(labels)) (resnet_cifar_main.py:270-272).  Ops K15, K16, K18
// plus the dense layer's bias add and bias gradient (K14).  Labels are class indices (the
// reference's one-hot float labels carry the same information).
#include "common.cuh"
#include "host.cuh"
#include "../include/drn_b200.h"

namespace drn {
namespace {

constexpr int kThreads = 128;

__device__ __forceinline__ float block_reduce(float v, float* smem, bool is_max) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float u = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? fmaxf(v, u) : v + u;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  float r = smem[0];
  for (int w = 1; w < kThreads / 32; ++w) r = is_max ? fmaxf(r, smem[w]) : r + smem[w];
  return r;
}

__global__ void softmax_xent_kernel(const __nv_bfloat16* __restrict__ logits, int ld,
                                    const float* __restrict__ bias, const int* __restrict__ labels,
                                    int classes, float grad_scale, __nv_bfloat16* __restrict__ dlogits,
                                    float* __restrict__ dbias, float* __restrict__ probs,
                                    float* __restrict__ loss_sum, int* __restrict__ correct, unsigned int* det) {
  pdl_trigger();
  pdl_wait();
  __shared__ float s_red[kThreads / 32];
  __shared__ int s_arg[kThreads / 32];
  const int row = blockIdx.x;
  const __nv_bfloat16* z = logits + (long long)row * ld;
  const int label = labels[row];

  float m = -INFINITY;
  int arg = 0x7fffffff;
  for (int j = threadIdx.x; j < classes; j += kThreads) {
    const float v = __bfloat162float(z[j]) + (bias ? bias[j] : 0.f);
    if (v > m) {
      m = v;
      arg = j;
    }
  }
  const float gmax = block_reduce(m, s_red, true);
  // first index attaining the maximum
  int cand = (m == gmax) ? arg : 0x7fffffff;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cand = min(cand, __shfl_xor_sync(0xffffffffu, cand, o));
  __syncthreads();
  if ((threadIdx.x & 31) == 0) s_arg[threadIdx.x >> 5] = cand;
  __syncthreads();
  int garg = s_arg[0];
  for (int w = 1; w < kThreads / 32; ++w) garg = min(garg, s_arg[w]);

  float se = 0.f;
  for (int j = threadIdx.x; j < classes; j += kThreads) {
    const float v = __bfloat162float(z[j]) + (bias ? bias[j] : 0.f);
    se += __expf(v - gmax);
  }
  const float sum = block_reduce(se, s_red, false);
  const float inv = 1.f / sum;

  det_turn_begin(det, blockIdx.x);    // deterministic mode: the rows add their loss / dbias terms in order
  if (threadIdx.x == 0) {
    const float zl = __bfloat162float(z[label]) + (bias ? bias[label] : 0.f);
    atomicAdd(loss_sum, (__logf(sum) + gmax - zl));
    if (garg == label) atomicAdd(correct, 1);
  }
  for (int j = threadIdx.x; j < ld; j += kThreads) {
    float g = 0.f;
    if (j < classes) {
      const float v = __bfloat162float(z[j]) + (bias ? bias[j] : 0.f);
      const float pj = __expf(v - gmax) * inv;
      if (probs) probs[(long long)row * classes + j] = pj;
      g = (pj - (j == label ? 1.f : 0.f)) * grad_scale;
      if (dbias) atomicAdd(dbias + j, g);
    }
    if (dlogits) dlogits[(long long)row * ld + j] = __float2bfloat16(g);
  }
  det_turn_end(det, blockIdx.x + 1 == gridDim.x ? 0u : blockIdx.x + 1);
}

}  // namespace
}  // namespace drn

using namespace drn;

// logits [B][ld] bf16 (columns >= classes are padding), labels int32 [B].
// loss_sum += sum_i xent_i (caller divides by B); correct += #(argmax == label);
// dlogits (nullable) = (softmax - onehot) * grad_scale, padding columns zeroed;
// dbias (nullable, fp32 [classes]) += column sums of dlogits; probs (nullable) fp32 [B][classes].
extern "C" int drn_softmax_xent(const void* logits, int ld, const float* bias, const int* labels,
                                int B, int classes, float grad_scale, void* dlogits, float* dbias,
                                float* probs, float* loss_sum, int* correct, void* stream) {
  DRN_REQUIRE(classes <= ld, "softmax_xent: classes (%d) > ld (%d)", classes, ld);
  DRN_CHECK_CUDA(launch_pdl(softmax_xent_kernel, dim3(B), dim3(kThreads), 0, (cudaStream_t)stream,
                            (const __nv_bfloat16*)logits, ld, bias, labels, classes, grad_scale, (__nv_bfloat16*)dlogits,
      dbias, probs, loss_sum, correct, det_ticket(kDetLoss)));
  return 0;
}
