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
/* Give the watchdog a mapped pinned host word to report through (allocates 4 bytes of pinned host
 * memory once per device): lets drn_watchdog_status() name the stuck barrier even after the trap has
 * made the CUDA context unusable. (the reference relies on session re-creation.) */
int drn_watchdog_arm(void);
int drn_num_sms(void);
/* Programmatic dependent launch (griddepcontrol) for the library's launches: bit 0 = element-wise /
 * pooling / optimizer kernels, bit 1 = tcgen05 convolution kernels, -1 = default (off).  Pays for
 * launch-bound steps (CIFAR-sized layers: +4 % steps/s), costs throughput on ImageNet-sized ones. */
int drn_set_pdl_mask(int mask);
/* Deterministic mode (reference: none - TensorFlow's GPU reductions are not reproducible either; this
 * is for parity runs).  on = 1: every floating-point accumulation across thread blocks happens in a
 * fixed order, so a step is bit-identical from run to run on the same GPU type and grid: convolutions
 * take their BatchNorm statistics in a separate ordered pass instead of with epilogue atomics, the
 * filter gradient adds its split-K partial tiles in split order, the BatchNorm / loss / optimizer
 * reductions take turns block by block.  0 = off, -1 = follow the DRN_DETERMINISTIC environment
 * variable (default off).  Costs throughput. */
int drn_set_deterministic(int on);
int drn_deterministic(void);

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
  DRN_EPI_BNBWD = 2,  /* y = conv(x, w) * [side*scale+shift > 0]   (ReLU+BN backward mask, K6/K9);
                         stats then receive sum(y) and sum(y*side) per channel                  */
  /* two side inputs (drn_conv_igemm_epi2): */
  DRN_EPI_BNBWD_ADD = 3,   /* y = (conv + side2) * [side*scale+shift > 0]; stats as in BNBWD      */
  DRN_EPI_BNBWD_APPLY = 4  /* y = A*(conv [+ side2]) * [side*scale+shift > 0] + B*side + C [+ side2]:
                              the whole BatchNorm+ReLU backward of the layer in front (K6), with the
                              coefficients of drn_bn_bwd_finalize                                */
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

/* Epilogues with a second side tensor (same shape as y).  mode = DRN_EPI_BNBWD_ADD or _APPLY
 * (side2_first: _APPLY adds side2 before the mask - a second gradient branch - instead of after the
 * BatchNorm formula - the identity shortcut's gradient), or DRN_EPI_BNBWD / _ADD with y = NULL for a
 * statistics-only pass.  bn_coef = float4[K] from drn_bn_bwd_finalize.  Replaces "write the masked
 * gradient, then run drn_bn_bwd_apply over it" by recomputing the (cheap, 1x1) convolution. */
int drn_conv_igemm_epi2(const drn_conv_desc* desc, const void* x, const void* w, void* y,
                        const void* side, const void* side2, float* stats, int stats_ld,
                        const void* bn_scale_shift, const void* bn_coef, int side2_first, int mode,
                        int force_bn, void* stream);

/* The same two kernels over a strided (possibly overlapping) view of x: element strides between
 * pixels along W, rows and images, and zero padding along W given separately from desc->pad_*
 * (which then applies to H only). desc->W / desc->Q describe the view. */
typedef struct {
  long long stride_w, stride_h, stride_n; /* in elements, multiples of 8 */
  int pad_w_lo, pad_w_hi;
} drn_conv_view;
int drn_conv_igemm_view(const drn_conv_desc* desc, const drn_conv_view* view, const void* x,
                        const void* w, void* y, const void* side, float* stats, int stats_ld,
                        const void* bn_scale_shift, int mode, int force_bn, void* stream);

/* y = conv(relu(x * scale + shift), w): BatchNorm + ReLU of the consuming layer applied to the A tile
 * in shared memory (four extra warps between the TMA load and the MMA), so the activated tensor is
 * never materialised.  pre_scale_shift = float2[C], 16-byte aligned.  Any filter size / stride (taps
 * that fall into the zero padding stay zero); C = 16 (N tiles 16/32), 32 (32/64) or a multiple of 64
 * (64/128). */
int drn_conv_igemm_pre(const drn_conv_desc* desc, const void* x, const void* pre_scale_shift,
                       const void* w, void* y, const void* side, float* stats, int stats_ld,
                       const void* bn_scale_shift, int mode, int force_bn, void* stream);

/* Parameter-server mode: "this layer's filter bank still lives on the shard owners".  Given to
 * drn_conv_igemm_ex, the convolution kernel itself performs the pull (all-gather) before it uses the
 * weights: its CTAs wait for the owners' per-bucket epoch flags (drn_ps_update with pull = 1), copy the
 * bank from the owners' arenas over NVLink into the local arena and only then start their weight loads;
 * activation tiles stream in meanwhile.  (Reference: the ps->worker variable read of
 * replica_device_setter, resnet_cifar_main.py:392-396.) */
typedef struct {
  const void* const* shadows; /* HOST array of nranks device pointers: every rank's bf16 working arena */
  int nranks, rank;
  const void* my_pad;         /* this rank's signal pad                                           */
  void* local_arena;          /* this rank's working arena                                        */
  unsigned int* counters;     /* two zero-initialised device words reserved for this layer        */
  long long bucket_lo, per;   /* the bucket's first element / elements per owner (drn_shard_per)  */
  int num_ps, bucket;
  long long off, count;       /* the layer's weights inside the arena (elements, multiples of 8)  */
} drn_pull_desc;
/* drn_conv_igemm with any of: a strided view (NULL = dense), a BatchNorm+ReLU prologue (NULL = none),
 * a fused weight pull (NULL = weights are local). */
int drn_conv_igemm_ex(const drn_conv_desc* desc, const drn_conv_view* view, const void* pre_scale_shift,
                      const drn_pull_desc* pull, const void* x, const void* w, void* y, const void* side,
                      float* stats, int stats_ld, const void* bn_scale_shift, int mode, int force_bn,
                      void* stream);

/* BatchNorm finalize folded into the convolution that produces the statistics: the last CTA of the
 * grid to finish turns the completed sums into (scale, shift), (mean, 1/std) and the moving averages
 * (the arithmetic of drn_bn_finalize), so a prologue-fused consumer needs no BatchNorm launch in
 * between.  counter: one zero-initialised device word per BatchNorm (the kernel resets it).
 * (Reference: tf.layers.batch_normalization, resnet_model_official.py:41-50.) */
typedef struct {
  unsigned int* counter;
  long long count;           /* elements per channel (N*P*Q) */
  const float* gamma;
  const float* beta;
  float* run_mean;           /* NULL: no moving-average update */
  float* run_var;
  float decay, eps;
  void* scale_shift;         /* out float2[K] */
  void* mean_invstd;         /* out float2[K] */
} drn_bn_fin;
/* drn_conv_igemm_ex + finalize (fin may be NULL); PLAIN / RESADD epilogues with stats != NULL. */
int drn_conv_igemm_fin(const drn_conv_desc* desc, const drn_conv_view* view, const void* pre_scale_shift,
                       const drn_pull_desc* pull, const drn_bn_fin* fin, const void* x, const void* w, void* y,
                       const void* side, float* stats, int stats_ld, const void* bn_scale_shift, int mode,
                       int force_bn, void* stream);

/* Data gradient (K2) of the convolution described by `desc` (FORWARD geometry):
 *   dx[N,H,W,C] = sum_{k,r,s} dy[N,P,Q,K] (x) w,  given w_dgrad = drn_weight_prep_dgrad(w).
 * Runs the implicit-GEMM kernel above on the equivalent stride-1 problem; for stride > 1 dy is first
 * zero-inserted into `scratch` (bf16 [N,(P-1)s+1,(Q-1)s+1,K]; drn_conv_dgrad_scratch_bytes(desc),
 * 0 for stride 1). side / stats / bn_scale_shift / mode as in drn_conv_igemm, shaped like dx. */
long long drn_conv_dgrad_scratch_bytes(const drn_conv_desc* desc);
int drn_conv_dgrad(const drn_conv_desc* desc, const void* dy, const void* w_dgrad, void* dx,
                   const void* side, float* stats, int stats_ld, const void* bn_scale_shift, int mode,
                   void* scratch, long long scratch_bytes, void* stream);

/* Data gradient of a 3x3 / stride-2 / pad-(1,1) convolution as four parity-phase stride-1
 * convolutions over dy (no zero insertion: a quarter of the tensor-core work of drn_conv_dgrad).
 * banks = drn_weight_prep_dgrad_s2(w) (9*K*C bf16); out = [4][N][H/2][W/2][C] bf16 with phase
 * (a,b) at index a*2+b holding dx[:, a::2, b::2, :]; drn_bn_bwd_mask_reduce(addend_kind = 4) or any
 * other consumer interleaves them. */
int drn_weight_prep_dgrad_s2(const void* w_krsc, void* banks, int K, int C, void* stream);
int drn_conv_dgrad_s2(const drn_conv_desc* desc, const void* dy, const void* banks, void* out,
                      void* stream);

/* Same operation as drn_conv_dgrad for callers that cannot allocate device memory when they call -
 * e.g. a process that gave the whole GPU to its own memory pool, or a latency-critical loop where
 * cudaMalloc/cudaFree (which synchronise the device) are off limits.  The zero-insertion scratch
 * lives in a fixed DRN_DGRAD_STATIC_SCRATCH_BYTES array inside the library (reserved on every GPU
 * whose context loads the library's device code - with lazy module loading that is the first kernel
 * launch or this function's first call, so a pool that takes "all remaining memory" must be created
 * after it); larger problems are refused with an error.  The first call on a device resolves
 * the array's device address - make it at start-up, before any stream capture.  All calls share
 * the array: every call rewrites what it uses, so nothing has to be cleared, but calls that can
 * overlap on the GPU race with each other.  Allocates nothing.  No counterpart in the reference. */
#define DRN_DGRAD_STATIC_SCRATCH_BYTES (16u << 20)
int drn_conv_dgrad_static(const drn_conv_desc* desc, const void* dy, const void* w_dgrad, void* dx,
                          const void* side, float* stats, int stats_ld, const void* bn_scale_shift,
                          int mode, void* stream);

/* Filter gradient (K3): dw[K][R][S][C] (fp32) += sum over pixels of dy[.,k] * x[. + (r,s), c].
 *   x  [N,H,W,C] bf16   dy [N,P,Q,K] bf16   dw fp32, ACCUMULATED with atomics (zero it first).
 *   split_k: 0 = choose automatically. */
int drn_conv_wgrad(const drn_conv_desc* desc, const void* x, const void* dy, float* dw,
                   int split_k, void* stream);


/* dw += dy (x) relu(x * scale + shift): the activation is rebuilt from the RAW tensor in shared memory
 * (counterpart of drn_conv_igemm_pre).  C % 64 == 0; pre_scale_shift = float2[C], 16-byte aligned. */
int drn_conv_wgrad_pre(const drn_conv_desc* desc, const void* x, const void* pre_scale_shift,
                       const void* dy, float* dw, int split_k, void* stream);

int drn_conv_wgrad_view(const drn_conv_desc* desc, const drn_conv_view* view, const void* x,
                        const void* dy, float* dw, int split_k, void* stream);

/* ---------------------------------------------------------------------------------------------
 * BatchNorm + ReLU (reference: batch_norm_relu, resnet_model_official.py:41-50; ops K5-K9)
 * stats rows: [0] = per-channel sum, [1] = per-channel sum of squares (or sum dm*y in backward).
 * scale_shift / mean_invstd are float2[C]; coef is float4[C].
 * ------------------------------------------------------------------------------------------- */
int drn_bn_finalize(const float* stats, int stats_ld, long long count, const float* gamma,
                    const float* beta, float* run_mean /*nullable*/, float* run_var, float decay,
                    float eps, void* scale_shift, void* mean_invstd, int C, void* stream);
int drn_bn_inference_params(const float* gamma, const float* beta, const float* run_mean,
                            const float* run_var, float eps, void* scale_shift, int C, void* stream);
/* out = [relu](y*scale + shift [+ addend]) over an [M][C] bf16 tensor */
int drn_bn_apply(const void* y, const void* scale_shift, const void* addend, void* out, long long M,
                 int C, int relu, void* stream);
/* drn_bn_finalize + drn_bn_apply in ONE launch (every thread derives scale/shift for its own 8
 * channels; block 0 publishes scale_shift / mean_invstd / moving averages). count = M. */
int drn_bn_finalize_apply(const void* y, const float* stats, int stats_ld, long long count,
                          const float* gamma, const float* beta, float* run_mean, float* run_var,
                          float decay, float eps, void* scale_shift, void* mean_invstd, void* out,
                          int C, int relu, void* stream);
/* drn_bn_bwd_finalize + drn_bn_bwd_apply in ONE launch */
int drn_bn_bwd_finalize_apply(const void* dm, const void* y, const float* stats, int stats_ld,
                              long long count, const float* gamma, const void* mean_invstd,
                              float* dgamma, float* dbeta, const void* addend, void* dy, int C,
                              void* stream);
int drn_bn_stats(const void* y, float* stats, int stats_ld, long long M, int C, void* stream);
/* stats[0][c] += sum_m a[m][c];  stats[1][c] += sum_m a[m][c] * y[m][c]  (bf16 [M][C] tensors) */
int drn_bn_pair_reduce(const void* a, const void* y, float* stats, int stats_ld, long long M, int C,
                       void* stream);
/* dm = (da [+ addend]) * [y*scale+shift > 0]; stats += (sum dm, sum dm*y).
 * addend_kind: 0 none, 1 same shape, 2 compact stride-2 tensor [N,Ha,Wa,C] added at even pixels,
 * 3 per-image vector [N][C] times bcast_scale (average-pool gradient; da may then be NULL),
 * 4 no addend, but da is the output of drn_conv_dgrad_s2: four parity-phase tensors
 *   [4][N][H/2][W/2][C], phase = (h & 1) * 2 + (w & 1). */
int drn_bn_bwd_mask_reduce(const void* da, const void* addend, int addend_kind, int N, int H, int W,
                           int Ha, int Wa, float bcast_scale, const void* y, const void* scale_shift,
                           void* dm, float* stats, int stats_ld, int C, void* stream);
/* Post-activation (v1) residual blocks, out = relu(bn(y) + shortcut): dm = g * [out > 0],
 * stats += (sum dm, sum dm*y) and, for a projection shortcut with its own BatchNorm over y2,
 * stats2 += (sum dm, sum dm*y2) (y2 / stats2 nullable together).  The reference ships v2 only
 * (resnet_model_official.py:15-29); v1 is the north-star's extra model family. */
int drn_relu_bwd_reduce(const void* g, const void* out, const void* y, const void* y2, void* dm,
                        float* stats, float* stats2, int stats_ld, long long M, int C, void* stream);
/* dgamma += sum dm*xhat, dbeta += sum dm (both nullable), coef = (A,B,C) of dy = A*dm + B*y + C */
int drn_bn_bwd_finalize(const float* stats, int stats_ld, long long count, const float* gamma,
                        const void* mean_invstd, float* dgamma, float* dbeta, void* coef, int C,
                        void* stream);
int drn_bn_bwd_apply(const void* dm, const void* y, const void* coef, const void* addend, void* dy,
                     long long M, int C, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Pooling (reference resnet_model_official.py:314-316 max pool, :266-271/:334-339 average pool)
 * ------------------------------------------------------------------------------------------- */
int drn_maxpool_fwd(const void* x, void* out, void* argmax /*uint8 [N,P,Q,C]*/, int N, int H, int W,
                    int C, int ksize, int stride, int pad_lo, int P, int Q, void* stream);
int drn_maxpool_bwd(const void* dy, const void* argmax, void* dx, int N, int H, int W, int C,
                    int ksize, int stride, int pad_lo, int P, int Q, void* stream);
/* out[n][c] = mean_hw relu(y*scale + shift) */
int drn_bnrelu_avgpool(const void* y, const void* scale_shift, void* out, int N, int HW, int C,
                       void* stream);

/* ---------------------------------------------------------------------------------------------
 * Loss (reference resnet_model.py:77-82, resnet_cifar_main.py:270-272; ops K14-K16, K18)
 * ------------------------------------------------------------------------------------------- */
int drn_softmax_xent(const void* logits, int ld, const float* bias, const int* labels, int B,
                     int classes, float grad_scale, void* dlogits, float* dbias, float* probs,
                     float* loss_sum, int* correct, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Optimizer / weight layouts (reference resnet_model.py:85-99; ops K17, K19)
 * ------------------------------------------------------------------------------------------- */
/* hyper = device float[4] {lr, momentum, weight_decay, grad_scale}; n % 4 == 0 */
int drn_sgd_momentum(float* master, float* grad, float* mom, void* shadow_bf16, long long n,
                     const float* hyper, float* l2_sum, int zero_grad, void* stream);
/* w_dgrad[c][R-1-r][S-1-s][k] = w_krsc[k][r][s][c] (bf16) */
int drn_weight_prep_dgrad(const void* w_krsc, void* w_dgrad, int K, int R, int S, int C,
                          void* stream);
/* the same for many layers in one launch; table = device array of {int64 src_off, int64 dst_off,
 * int32 K, RS, C, tile_begin} (offsets in elements; tile_begin = running sum of
 * RS*ceil(K/64)*ceil(C/64)), total_tiles = grand total; K and C even */
int drn_weight_prep_dgrad_batched(const void* w_arena, void* dgrad_arena, const void* table,
                                  int n_layers, int total_tiles, void* stream);
/* out[n, p*stride, q*stride, :] = dy[n, p, q, :], zero elsewhere; out is [N,U,V,C] */
int drn_upsample_zero(const void* dy, void* out, int N, int P, int Q, int U, int V, int C,
                      int stride, void* stream);
int drn_f32_to_bf16(const float* src, void* dst, long long n, void* stream);
/* a[i] *= b[i] */
int drn_mul_inplace_f32(float* a, const float* b, long long n, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Input pipelines on device (reference resnet_cifar_main.py:157-200, cifar_input.py:66-75,
 * vgg_preprocessing.py; ops K13, K21, K22)
 * ------------------------------------------------------------------------------------------- */
int drn_cifar_augment(const void* records, int record_stride, int image_offset, void* out, int B,
                      int pad, int training, unsigned long long seed, unsigned int step,
                      long long index_base, int cpad, void* stream);
/* out is [B, crop/2, out_pitch, 16] bf16; pixels land at columns out_xoff .. out_xoff+crop/2-1
 * (out_pitch 0 = dense). Border columns are never written: keep them zero. */
int drn_imagenet_prep(const void* src, long long image_stride, int src_h, int src_w,
                      const int* crop_yx_flip, void* out, int B, int crop, float mean_r,
                      float mean_g, float mean_b, int out_pitch, int out_xoff, void* stream);
int drn_resize_bilinear(const void* src, int sh, int sw, float* dst, int dh, int dw, void* stream);
/* Batched aspect-preserving resize (TF1 bilinear, align_corners=False) + crop directly from the planar
 * RGB output of a GPU JPEG decoder: src_ptrs = DEVICE array of B pointers to uint8 [3][h][w];
 * meta = DEVICE int[B][6] {src_h, src_w, new_h, new_w, off_y, off_x}; out = uint8 [B][crop][crop][3]. */
int drn_resize_crop_batch(const void* src_ptrs, const int* meta, void* out, int B, int crop,
                          void* stream);

/* Batched JPEG decode on the GPU (nvJPEG; hardware engine when the device has one), op K23.  The
 * reference decodes on five CPU threads (tf.image.decode_jpeg, resnet_imagenet_main.py:138-168).
 * drn_jpeg_create allocates nvJPEG's internal device/pinned work buffers: call it at start-up.
 *   data/lengths: HOST arrays of B host pointers / sizes; wh: HOST int[B][2] (width, height);
 *   out: HOST array of B DEVICE pointers, each uint8 [3][h][w] (planar RGB, the input format of
 *   drn_resize_crop_batch); status: HOST int[B], 0 = decoded, otherwise use a host decoder. */
int drn_jpeg_create(int prefer_hardware, void** out_handle);
int drn_jpeg_backend(void* handle); /* 2 = hardware JPEG engine, 1 = GPU hybrid, 0 = default */
int drn_jpeg_destroy(void* handle);
int drn_jpeg_info(void* handle, const unsigned char* const* data, const size_t* lengths, int B, int* wh);
int drn_jpeg_decode_batch(void* handle, const unsigned char* const* data, const size_t* lengths, int B,
                          void* const* out, const int* wh, int* status, void* stream);

/* Measurement probe (csrc/probe_tma.cu, bench/probe_tma.py; not used by the framework): persistent TMA
 * copy of a [rows][cols] bf16 matrix in boxes of 128 rows x box_cols (32 or 64) columns. */
int drn_probe_tma_copy(const void* src, void* dst, long long rows, int cols, int box_cols, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Host utilities (no GPU): CRC-32C (Castagnoli) used by TensorBundle checkpoints, TFRecord files
 * and TensorBoard event files (the reference gets these from TensorFlow's runtime).
 * ------------------------------------------------------------------------------------------- */
uint32_t drn_crc32c(const void* data, size_t n, uint32_t crc);
long long drn_tfrecord_index(const uint8_t* buf, size_t n, uint64_t* offsets, uint64_t* lengths,
                             long long max_records, int verify);

/* ---------------------------------------------------------------------------------------------
 * Collectives over NVLink / NVSwitch peer memory (reference call sites C1-C3).
 * bufs / pads / grads / shadows are HOST arrays of `nranks` device pointers (index = rank) into
 * symmetric allocations; every rank launches the same call. nblocks must match on all ranks.
 * ------------------------------------------------------------------------------------------- */
int drn_comm_pad_bytes(void);
/* in-place: buf[offset:offset+count] = scale * sum over ranks (hvd.DistributedOptimizer, C2) */
int drn_allreduce_f32(const void* const* bufs, void* mc_ptr /*nullable*/, const void* const* pads,
                      int rank, int nranks, long long offset, long long count, float scale,
                      int nblocks, int channel, void* stream);
/* SyncReplicas parameter-server step (C1): owners (ranks < num_ps) average their gradient slice
 * over all workers, apply L2 + momentum-SGD to master/mom and hand the new bf16 weights to all workers.
 *   mc_grads  multicast address of the gradient arena or NULL: with it the N copies are added inside
 *             the switch (multimem.ld_reduce) and cleared with one multimem.st
 *   want_l2   every owner adds its share of sum(w^2) into the l2 slot of EVERY rank's pad
 *             (drn_comm_pad_offset(2); zero it at the start of the step)
 *   pull = 0  the owner pushes the weights into every working arena and the call ends with a barrier
 *   pull = 1  the weights stay in the owner's arena; the owner publishes epoch ready[bucket][owner]
 *             in every rank's pad instead - consumers fetch them (drn_conv_igemm_pull,
 *             drn_ps_sync_vectors).  bucket in [0, 64); offset % 8 == 0. */
int drn_ps_update(const void* const* grads, void* mc_grads /*nullable*/, const void* const* shadows,
                  float* master, float* mom, const float* hyper, int rank, int nranks, int num_ps,
                  long long offset, long long count, const void* const* pads, int want_l2, int pull,
                  int bucket, int nblocks, int channel, void* stream);
/* elements per shard owner of a bucket of `count` elements (multiple of 8) */
long long drn_shard_per(long long count, int num_ps);
/* One-shot allreduce fused with the optimizer, for small buckets (C2 + K19): every rank reduces the
 * WHOLE range from all ranks (multimem.ld_reduce when mc_grads is given), applies 1/N, L2 and
 * momentum-SGD to its own master/mom/shadow and clears its own gradients after a second barrier. */
int drn_allreduce_sgd_oneshot(const void* const* grads, void* mc_grads /*nullable*/, float* master,
                              float* mom, void* shadow_bf16, const float* hyper, int rank, int nranks,
                              long long offset, long long count, const void* const* pads,
                              float* l2_sum /*nullable*/, int nblocks, int channel, void* stream);
/* copy root's buffer into every rank's buffer (BroadcastGlobalVariablesHook, C3); with mc_ptr the
 * root issues one multimem.st per vector and the switch replicates it */
int drn_bcast_f32(const void* const* bufs, void* mc_ptr /*nullable*/, const void* const* pads, int rank,
                  int nranks, int root, long long count, int nblocks, int channel, void* stream);
/* fp32 copies of the vector variables (BatchNorm gamma/beta, dense bias) from the bf16 working arena.
 * table = device {int64 elem_off; int32 count; int32 map}[n_entries]; maps = device
 * {int64 lo; int64 per; int32 num_ps; int32 bucket}[]; pull = 1 fetches them from the shard owners
 * (after waiting for the shards' epochs) and refreshes the local working arena as well. */
int drn_ps_sync_vectors(const void* const* shadows, void* local_shadow, float* dst_f32,
                        const void* table, int n_entries, const void* maps, const void* my_pad,
                        int rank, int nranks, int pull, void* stream);
/* byte offsets inside a pad: 0 = ready[64][16] uint32, 1 = epoch[64] uint32, 2 = l2 float */
int drn_comm_pad_offset(int which);
int drn_comm_barrier(const void* const* pads, int rank, int nranks, int channel, void* stream);
/* start-up only; these two allocate / map device memory (CUDA IPC fallback for symmetric memory) */
int drn_ipc_alloc(size_t bytes, void** out_ptr, void* out_handle64);
int drn_ipc_open(const void* handle64, void** out_ptr);

#ifdef __cplusplus
}
#endif
#endif /* DRN_B200_H_ */
