/* C ABI of libdrn_b200.so - the hand-written sm_100a kernels behind the
 * distributed_resnet_tensorflow_b200 Python package.
 *
 * Conventions
 *   - every pointer is a device pointer unless the name says `host`;
 *   - activations are NHWC bf16, filters are [Cout][R][S][Cin] bf16 ("KRSC");
 *   - `stream` is a cudaStream_t passed as void*;
 *   - every function returns 0 on success; on failure drn_last_error() describes why;
 *   - no function allocates device memory unless its comment says so.
 *
 * "reference" in the comments is hanalice/Distributed-ResNet-Tensorflow; the op numbers
 * (K1..K23, C1..C3) are the ones of SURVEY.md section 2.6 / 2.7.
 */
#ifndef DRN_B200_H_
#define DRN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* drn_last_error(void);
/* Non-zero if a bounded in-kernel barrier wait timed out since the last call (the kernel trapped). */
int drn_watchdog_status(void);
int drn_num_sms(void);

/* ---------------------------------------------------------------------------------------------
 * Convolution (reference: tf.layers.conv2d in resnet_model_official.py:80-91; ops K1-K4)
 * ------------------------------------------------------------------------------------------- */
typedef struct {
  int N, H, W, C; /* input NHWC; C is 16, 32 or a multiple of 64 (pad narrower inputs)       */
  int K;          /* output channels, multiple of 16                                          */
  int R, S;       /* filter height / width                                                    */
  int stride;     /* same in both directions                                                  */
  int pad_lo;     /* zero padding before (top/left)                                           */
  int pad_hi;     /* zero padding after (bottom/right)                                        */
  int P, Q;       /* output height / width = (H + pad_lo + pad_hi - R) / stride + 1           */
} drn_conv_desc;

enum {
  DRN_EPI_PLAIN = 0,  /* y = conv(x, w)                                                        */
  DRN_EPI_RESADD = 1, /* y = conv(x, w) + side                 (shortcut add, K10)             */
  DRN_EPI_BNBWD = 2   /* y = conv(x, w) * [side*scale+shift > 0]   (ReLU+BN backward mask, K6/K9);
                         stats then receive sum(y) and sum(y*side) per channel                  */
};

/* Implicit-GEMM convolution on tcgen05/TMEM with im2col TMA.
 *   x     [N,H,W,C] bf16      w [K][R*S*C] bf16      y [N,P,Q,K] bf16
 *   side  [N,P,Q,K] bf16 or NULL (see mode)
 *   stats float[2][stats_ld] or NULL: row 0 += per-channel sum of y, row 1 += sum of y*y
 *         (or sum of y*side in BNBWD mode). Accumulated with atomics: zero it first.
 *   bn_scale_shift float2[K] (BNBWD only)
 *   force_bn: 0 = pick the N tile automatically, else 16/32/64/128/256.
 * Also serves as the dgrad of a stride-1 convolution when given dy and the flipped filters. */
int drn_conv_igemm(const drn_conv_desc* desc, const void* x, const void* w, void* y,
                   const void* side, float* stats, int stats_ld, const void* bn_scale_shift,
                   int mode, int force_bn, void* stream);

/* Filter gradient (K3): dw[K][R][S][C] (fp32) += sum over pixels of dy[.,k] * x[. + (r,s), c].
 *   x  [N,H,W,C] bf16   dy [N,P,Q,K] bf16   dw fp32, ACCUMULATED with atomics (zero it first).
 *   split_k: 0 = choose automatically. */
int drn_conv_wgrad(const drn_conv_desc* desc, const void* x, const void* dy, float* dw,
                   int split_k, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DRN_B200_H_ */
