// Implicit-GEMM convolution for sm_100a: persistent, warp-specialised, tcgen05 + TMEM + TMA.
//
//   D[m, n] = sum_{r,s,c} X[pixel(m) + (r,s), c] * Wt[n, (r,s,c)]        m = (img, p, q)  n = out ch
//
// * A operand (activations, NHWC bf16) is fetched with im2col-mode TMA: one instruction brings
//   128 consecutive output pixels x KC channels of one filter tap, zero-filled at the borders,
//   already in the 128/64/32-byte-swizzled K-major layout tcgen05.mma consumes.
// * B operand (filters, [Cout][R*S*C] bf16) is fetched with tiled TMA.
// * Accumulators live in TMEM, double-buffered, so the epilogue of tile i overlaps the main loop
//   of tile i+1.  One elected thread issues the MMAs.
// * The epilogue (two groups of 4 warps, alternating 32-column chunks, double-buffered) drains TMEM through swizzled
//   smem and TMA-stores bf16.  It optionally fuses
//     - a residual add (ResNet shortcut;  reference resnet_model_official.py:130,175),
//     - per-channel sum / sum-of-squares of the stored tensor (the batch statistics of the NEXT
//       BatchNorm; reference resnet_model_official.py:45-48),
//     - the ReLU/BatchNorm backward mask and the two per-channel reductions BatchNorm backward
//       needs (sum dm, sum dm*y) when the kernel is used as a dgrad.
//
// The same kernel is the dgrad for stride-1 convolutions: pass dy as X and the flipped,
// transposed filter bank (see optim.cu: drn_weight_prep) as Wt.
//
// Covers reference ops K1, K2, K4, K9 (as producer of statistics), K10.
#include <stdlib.h>

#include "common.cuh"
#include "comm_layout.cuh"
#include "host.cuh"
#include "../include/drn_b200.h"

namespace drn {

namespace {

constexpr int kBM = 128;            // output pixels per tile (UMMA M)
constexpr int kEpiBufs = 3;         // staging buffers per epilogue group (three: ONE barrier per chunk)
constexpr int kNumThreads = 384;    // warp0 TMA (A), warp1 MMA, warp2 TMEM alloc, warp3 TMA (B), warps 4-11 epilogue (2 groups)
constexpr int kEpiThreads = 128;
constexpr int kMaxStages = 8;

// BatchNorm "finalize" of the statistics this convolution accumulates, done by the LAST CTA of the grid
// to finish (threadfence + counter), so the consumer's scale/shift table exists when the kernel ends and
// no separate launch is needed between a convolution and the prologue-fused convolution that follows it.
struct BnFin {
  unsigned int* counter;   // zero before the launch, reset to zero by the last CTA; nullptr = no finalize
  float inv_count, unbias, one_minus_decay, eps;
  const float* gamma;
  const float* beta;
  float* run_mean;         // nullable (no moving-average update)
  float* run_var;
  float2* scale_shift;     // out [Cout]
  float2* mean_invstd;     // out [Cout]
  long long count;         // host side only (deterministic mode runs bn_finalize as its own launch)
  float decay;
};

struct ConvParams {
  int m_tiles, n_tiles, total_tiles;
  int k_iters;     // R*S*(C/KC)
  int c_blocks;    // C/KC
  int C;           // input channels (padded)
  int S;           // filter width
  int PQ, Q;       // output plane / width
  int stride_h, stride_w, pad_h, pad_w;
  int n_stages;    // smem pipeline depth (<= kMaxStages)
  int n_res;       // side-input ring depth per epilogue group (0 in PLAIN mode)
  int mode;        // DRN_EPI_*
  int want_stats;  // accumulate per-channel sums into stats[0][*], stats[1][*]
  int n_valid;     // Cout (columns >= n_valid are padding)
  int stats_ld;    // leading dimension of stats rows
  float* stats;    // [2][stats_ld]
  const float2* bn_scale_shift;  // [Cout] (scale, shift) of the BN whose backward mask is applied
  const float4* bn_coef;         // [Cout] (A, B, C, -) of dy = A*dm + B*y + C (BNBWD_APPLY)
  int no_store;                  // statistics only: the tile is reduced but not written (BNBWD / BNBWD_ADD)
  int side2_first;               // BNBWD_APPLY: side2 is added BEFORE the mask (gradient of a second branch)
                                 //              instead of after the BatchNorm-backward formula (shortcut gradient)
  // PRE variant: the A operand is relu(x * scale + shift) of the RAW tensor x (BatchNorm + ReLU of
  // the consuming layer applied in shared memory between the TMA load and the MMA)
  const float2* pre_ss;  // [C] (scale, shift)
  long long M;           // valid output rows (rows beyond are forced to zero after the transform)
  int H_in, W_in;        // input plane (PRE: filter taps that fall into the zero padding stay zero)
  unsigned spin_ns;      // sleep between polls of the non-critical mbarrier waits (0 = spin)
  // PULL: the filter bank is first fetched from the parameter-server shards (peer memory over
  // NVLink) into the local working arena by warp 2 of every CTA, each CTA taking 1/gridDim of it;
  // the weight producer (warp 3) issues its first TMA load only after all CTAs have delivered
  PullArgs pull;         // pull.shadows.p[0] == nullptr: weights are already local
  BnFin fin;
};
constexpr int kPreThreads = 128;    // warps 12-15 of the PRE variant

// second side input / coefficient table of the two-side epilogue modes (thread-local request set by
// drn_conv_igemm_epi2 around the ordinary launcher)
const void*& current_bnfin() {
  static thread_local const void* v = nullptr;
  return v;
}

struct Epi2 {
  const void* side2;
  const void* coef;
  int side2_first;
};

constexpr int kMaxRes = 8;          // side-input ring depth per epilogue group (upper bound)
constexpr int kSmemBudget = 226 * 1024;

template <int KC, int BN, bool TWO = false>
struct Cfg {
  static constexpr int kABytes = kBM * KC * 2;
  static constexpr int kBRows = TWO ? BN / 2 : BN;     // CTA-pair: each CTA stages half of the B tile
  static constexpr int kBBytesRaw = kBRows * KC * 2;
  static constexpr int kBBytes = (kBBytesRaw + 1023) / 1024 * 1024;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kCW = BN < 32 ? BN : 32;       // epilogue chunk width (columns)
  static constexpr int kChunks = BN / kCW;
  static constexpr int kRowBytes = kCW * 2;            // bytes per staged row (= swizzle span)
  static constexpr int kChunkBytes = kBM * kRowBytes;  // <= 8 KB
  static constexpr int kOutBytes = 2 * kEpiBufs * kChunkBytes;   // 2 groups x out[kEpiBufs]
  static constexpr int kAuxBytes = 4096 + 2048 + 1024 + 8192;  // 2 x scale/shift cache, 2 x stats scratch, barriers, 2 x coefficient cache
  static constexpr int kTmemColsRaw = 2 * BN;
  static constexpr int kTmemCols = kTmemColsRaw <= 32    ? 32
                                   : kTmemColsRaw <= 64  ? 64
                                   : kTmemColsRaw <= 128 ? 128
                                   : kTmemColsRaw <= 256 ? 256
                                                         : 512;
  // smem plan for a given side-input ring depth (per group)
  static constexpr int stages_for(int n_res) {
    int s = (kSmemBudget - 1024 - kAuxBytes - kOutBytes - 2 * n_res * kChunkBytes) / kStageBytes;
    return s > kMaxStages ? kMaxStages : s;
  }
  static constexpr int smem_for(int n_res, int stages) {
    return 1024 + kAuxBytes + kOutBytes + 2 * n_res * kChunkBytes + stages * kStageBytes;
  }
};

// EPI >= 0 fixes (epilogue mode, statistics) at compile time as mode * 2 + stats: the epilogue is
// unrolled over its chunks, and carrying all three modes in every copy made the kernel large enough
// to stall on instruction fetch (ncu: no_instruction).  EPI = -1 keeps them run-time parameters.
template <int KC, int BN, bool PRE = false, bool TWO = false, int EPI = -1>
__global__ void __launch_bounds__(PRE ? kNumThreads + kPreThreads : kNumThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapB,
                  const __grid_constant__ CUtensorMap tmapC, const __grid_constant__ CUtensorMap tmapR,
                  const __grid_constant__ CUtensorMap tmapR2, const ConvParams p) {
  using C_ = Cfg<KC, BN, TWO>;
  static_assert(!(PRE && TWO), "the BatchNorm prologue has no CTA-pair variant");
  const int kStages = p.n_stages;
  constexpr int kCW = C_::kCW;
  constexpr int kRowBytes = C_::kRowBytes;
  // CTA pair (TWO): the two CTAs of a (2,1,1) cluster own the row tiles 2*m_pair and 2*m_pair+1 of
  // the same column tile; rank 0 issues ONE 256-row tcgen05.mma.cta_group::2 per k-step that reads
  // both CTAs' A tiles and both halves of the B tile, so the weights cross the L2 fabric once per
  // pair instead of once per CTA.  p.m_tiles then counts row-tile PAIRS.
  const uint32_t cta_rank = TWO ? cluster_ctarank() : 0u;
  const bool leader_cta = cta_rank == 0;
  const int mode = EPI < 0 ? p.mode : (EPI >> 1);
  const bool want_stats = EPI < 0 ? (p.want_stats != 0) : ((EPI & 1) != 0);

  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment: required by the 128B swizzle atom (8 rows x 128 B).
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const uint32_t sStages = smem_base;
  const uint32_t sOut = sStages + kStages * C_::kStageBytes;
  const uint32_t sRes = sOut + C_::kOutBytes;
  const bool two_side = mode == DRN_EPI_BNBWD_ADD || mode == DRN_EPI_BNBWD_APPLY;
  const uint32_t sRes2 = sRes + 2 * p.n_res * C_::kChunkBytes;      // second side-input rings (two_side)
  const uint32_t sAux = sRes2 + (two_side ? 2 * p.n_res * C_::kChunkBytes : 0);
  float2* sScaleShift = reinterpret_cast<float2*>(smem_gen + (sAux - smem_base));  // 2 x 256 entries
  const uint32_t sBar = sAux + 4096 + 2048;
  // barrier map (8 bytes each)
  auto full_bar = [&](int s) { return sBar + 8u * s; };
  auto empty_bar = [&](int s) { return sBar + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int b) { return sBar + 8u * (2 * kMaxStages + b); };
  auto tempty_bar = [&](int b) { return sBar + 8u * (2 * kMaxStages + 2 + b); };
  auto res_bar = [&](int b) { return sBar + 8u * (2 * kMaxStages + 4 + b); };  // b in [0, 2*kMaxRes)
  float4* sCoef = reinterpret_cast<float4*>(smem_gen + (sBar + 1024 - smem_base));   // 2 x 256 entries
  const uint32_t sTmemPtr = sBar + 8u * (2 * kMaxStages + 4 + 2 * kMaxRes);
  auto ready_bar = [&](int s) { return sBar + 8u * (2 * kMaxStages + 4 + 2 * kMaxRes + 1 + s); };  // PRE
  volatile uint32_t* tmem_ptr_gen =
      reinterpret_cast<volatile uint32_t*>(smem_gen + (sTmemPtr - smem_base));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  pdl_trigger();

  // Each CTA owns a contiguous range of the n-major tile order, so it crosses a column-tile
  // boundary (statistics flush, scale/shift reload) at most once or twice.
  const int n_units = TWO ? (int)gridDim.x / 2 : (int)gridDim.x;   // CTAs, or CTA pairs
  const int unit = TWO ? (int)blockIdx.x / 2 : (int)blockIdx.x;
  const int t_q = p.total_tiles / n_units, t_r = p.total_tiles % n_units;
  const int t_begin = unit * t_q + (unit < t_r ? unit : t_r);
  const int t_end = t_begin + t_q + (unit < t_r ? 1 : 0);
  // row tile of this CTA for the scheduler's tile index
  auto row_tile = [&](int m_idx) { return TWO ? 2 * m_idx + (int)cta_rank : m_idx; };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmapA);
    tma_prefetch_desc(&tmapB);
    tma_prefetch_desc(&tmapC);
    if (mode != DRN_EPI_PLAIN) tma_prefetch_desc(&tmapR);
    if (two_side) tma_prefetch_desc(&tmapR2);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 2);   // the A producer and the B producer
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tfull_bar(b), 1);
      // one arrival per thread per chunk (from both CTAs of a pair: the leader's barrier gates the MMA)
      mbar_init(tempty_bar(b), TWO ? 2 * (kEpiThreads / 32) * C_::kChunks : kEpiThreads * C_::kChunks);
    }
    for (int b = 0; b < 2 * kMaxRes; ++b) mbar_init(res_bar(b), 1);
    if (PRE)
      for (int s = 0; s < kStages; ++s) mbar_init(ready_bar(s), kPreThreads / 32);
    fence_mbar_init();
  }
  if (warp == 2) {
    if (TWO) {
      tmem_alloc_2cta(sTmemPtr, C_::kTmemCols);
      tmem_relinquish_2cta();
    } else {
      tmem_alloc(sTmemPtr, C_::kTmemCols);
      tmem_relinquish();
    }
  }
  tc_fence_before();
  if (TWO) cluster_sync_all();   // the peer's barriers must be initialised before anything signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;
  // everything above (barriers, TMEM, descriptor prefetch) overlapped the previous kernel's tail
  pdl_wait();

  if (warp == 0 || warp == 3) {
    // ======================= TMA producers =======================
    // Two single-thread producers share every stage: warp 0 issues the im2col loads of the A tile,
    // warp 3 the loads of the weight tile.  One thread doing both (plus its index arithmetic and
    // two mbarrier operations per stage) was the slowest stage of the whole pipeline.
    const bool do_a = warp == 0;
    if (elect_one()) {
      if (!do_a && p.pull.shadows.p[0] != nullptr) {
        // all CTAs have delivered their part of the filter bank (generic-proxy stores, released with
        // __threadfence + atomicAdd); order them before this thread's TMA (async-proxy) reads
        const uint64_t t0 = globaltimer_ns();
        uint32_t it = 0;
        while (ld_acquire_gpu_u32(p.pull.counters) < gridDim.x) {
          if (((++it) & 0x3f) == 0 && globaltimer_ns() - t0 > DRN_WATCHDOG_NS) watchdog_fire(0x211);
        }
        fence_proxy_async_all();
        // the last CTA through resets the two words for the next launch of this layer
        if (atomicAdd(p.pull.counters + 1, 1u) == gridDim.x - 1) {
          p.pull.counters[1] = 0u;
          __threadfence();
          p.pull.counters[0] = 0u;
        }
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int t = t_begin; t < t_end; ++t) {
        const int n_tile = t / p.m_tiles;
        int img = 0, w0 = 0, h0 = 0;
        if (do_a) {
          const int m0 = row_tile(t - n_tile * p.m_tiles) * kBM;
          img = m0 / p.PQ;
          const int rem = m0 - img * p.PQ;
          const int op = rem / p.Q;
          const int oq = rem - op * p.Q;
          w0 = oq * p.stride_w - p.pad_w;
          h0 = op * p.stride_h - p.pad_h;
        }
        const int b_row = n_tile * BN + (TWO ? (int)cta_rank * C_::kBRows : 0);
        // (channel block, filter row, filter column, weight column) advance incrementally: integer
        // divisions per k-step would be a visible part of a single thread's loop
        int cb = 0, r = 0, s = 0, b_col = 0;
        for (int it = 0; it < p.k_iters; ++it) {
          mbar_wait_sleep(empty_bar(stage), phase ^ 1, 0x10 + stage, p.spin_ns);
          const uint32_t a_dst = sStages + stage * C_::kStageBytes;
          if (do_a) {
            if (TWO) {
              // both CTAs of a pair load their own rows; the bytes are counted on the leader's barrier
              if (leader_cta) mbar_arrive_expect_tx(full_bar(stage), 2 * C_::kABytes);
              tma_load_im2col_4d_2cta(a_dst, &tmapA, full_bar(stage), cb * KC, w0, h0, img,
                                      static_cast<uint16_t>(s), static_cast<uint16_t>(r));
            } else {
              mbar_arrive_expect_tx(full_bar(stage), C_::kABytes);
              tma_load_im2col_4d(a_dst, &tmapA, full_bar(stage), cb * KC, w0, h0, img,
                                 static_cast<uint16_t>(s), static_cast<uint16_t>(r));
            }
          } else {
            if (TWO) {
              if (leader_cta) mbar_arrive_expect_tx(full_bar(stage), 2 * C_::kBBytesRaw);
              tma_load_2d_2cta(a_dst + C_::kABytes, &tmapB, full_bar(stage), b_col, b_row);
            } else {
              mbar_arrive_expect_tx(full_bar(stage), C_::kBBytesRaw);
              tma_load_2d(a_dst + C_::kABytes, &tmapB, full_bar(stage), b_col, b_row);
            }
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
          b_col += KC;            // weight columns are ordered (tap, channel): tap * C + cb * KC
          if (++cb == p.c_blocks) {
            cb = 0;
            if (++s == p.S) {
              s = 0;
              ++r;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================= MMA issuer =======================
    if (leader_cta && elect_one()) {
      constexpr uint32_t kIdesc = make_idesc_bf16(TWO ? 2 * kBM : kBM, BN, 0, 0);
      constexpr uint32_t kLayout = umma_layout_for_swizzle_bytes(KC * 2);
      constexpr uint32_t kSBO = 8 * KC * 2;  // 8-row group stride (K-major, dense rows)
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = t_begin; t < t_end; ++t) {
        mbar_wait(tempty_bar(acc), acc_phase ^ 1, 0x30 + acc);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int it = 0; it < p.k_iters; ++it) {
          mbar_wait(PRE ? ready_bar(stage) : full_bar(stage), phase, 0x20 + stage);
          tc_fence_after();
          const uint32_t a_addr = sStages + stage * C_::kStageBytes;
          const uint32_t b_addr = a_addr + C_::kABytes;
          const uint64_t a_desc = make_smem_desc(a_addr, 0, kSBO, kLayout);
          const uint64_t b_desc = make_smem_desc(b_addr, 0, kSBO, kLayout);
#pragma unroll
          for (int kk = 0; kk < KC / 16; ++kk) {
            // advance 16 elements (32 B) along K inside the swizzle span: +2 in (addr >> 4) units
            if (TWO) umma_f16_2cta(d_tmem, a_desc + 2 * kk, b_desc + 2 * kk, kIdesc, (it | kk) != 0 ? 1u : 0u);
            else umma_f16(d_tmem, a_desc + 2 * kk, b_desc + 2 * kk, kIdesc, (it | kk) != 0 ? 1u : 0u);
          }
          if (TWO) umma_commit_2cta(empty_bar(stage), 3);   // frees the stage in both CTAs
          else umma_commit(empty_bar(stage));
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        if (TWO) umma_commit_2cta(tfull_bar(acc), 3);       // both epilogues may drain their 128 rows
        else umma_commit(tfull_bar(acc));
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (PRE && warp >= 12) {
    // ======================= A-operand transform (PRE): BatchNorm + ReLU in shared memory ===========
    // Thread = (16-byte channel group g of the 128-byte row, 8 rows): a warp touches 4 complete rows
    // per access (512 contiguous bytes, conflict free).  The 128B swizzle stores logical chunk g of
    // row r at physical chunk g ^ (r & 7).
    if constexpr (PRE) {
      // Thread = (16-byte channel group g of a row, kRows rows): a warp touches whole rows per access
      // (512 contiguous bytes, conflict free).  The swizzle of a KC*2-byte row (128/64/32 B) stores
      // logical chunk g of row r at physical chunk g ^ ((r >> kSwzShift) & (kCPR - 1)).
      constexpr int kCPR = KC / 8;                 // 16-byte chunks per row
      constexpr int kRows = kCPR;                  // rows per thread: 128 threads cover 128 rows x kCPR chunks
      constexpr int kRowStep = kPreThreads / kCPR;
      constexpr int kRB = KC * 2;
      constexpr int kSwzShift = KC == 64 ? 0 : (KC == 32 ? 1 : 2);
      const int tt = threadIdx.x - kNumThreads;
      const int g = tt % kCPR;
      const int r0 = tt / kCPR;  // rows r0 + kRowStep * i
      uint32_t off[kRows];
#pragma unroll
      for (int i = 0; i < kRows; ++i) {
        const int row = r0 + kRowStep * i;
        off[i] = row * kRB + ((g ^ ((row >> kSwzShift) & (kCPR - 1))) << 4);
      }
      // a filter tap that falls into the zero padding must stay zero after relu(0 * scale + shift):
      // (input row, input column) of tap (0,0) for each of my rows, per tile; 1x1 layers without
      // padding only have the rows beyond M to blank
      const bool taps = p.k_iters != p.c_blocks || p.pad_h != 0 || p.pad_w != 0;
      int stage = 0;
      uint32_t phase = 0;
      for (int t = t_begin; t < t_end; ++t) {
        const int n_tile = t / p.m_tiles;
        const int m_tile = t - n_tile * p.m_tiles;
        const long long m0 = (long long)m_tile * kBM;
        const bool ragged = m0 + kBM > p.M;   // only the last row tile has rows to blank
        int ih0[kRows], iw0[kRows];
#pragma unroll
        for (int i = 0; i < kRows; ++i) {
          const long long m = m0 + r0 + kRowStep * i;
          ih0[i] = iw0[i] = 0;
          if (m >= p.M) {
            ih0[i] = -(1 << 28);                 // never valid
          } else if (taps) {
            const int rem = (int)(m % p.PQ);
            const int op = rem / p.Q;
            ih0[i] = op * p.stride_h - p.pad_h;
            iw0[i] = (rem - op * p.Q) * p.stride_w - p.pad_w;
          }
        }
        int cb = 0, fr = 0, fs = 0;
        for (int it = 0; it < p.k_iters; ++it) {
          const float4* sp = reinterpret_cast<const float4*>(p.pre_ss + cb * KC + g * 8);
          float4 q[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) q[k] = __ldg(sp + k);  // (sc0, sh0, sc1, sh1) x 4
          mbar_wait_sleep(full_bar(stage), phase, 0x60 + stage, p.spin_ns);
          const uint32_t a_addr = sStages + stage * C_::kStageBytes;
          // all loads first, then all arithmetic, then all stores: one warp per scheduler has to
          // find its instruction-level parallelism inside its own stream
          uint4 v[kRows];
#pragma unroll
          for (int i = 0; i < kRows; ++i) v[i] = ld_shared_v4(a_addr + off[i]);
#pragma unroll
          for (int i = 0; i < kRows; ++i) {
            uint32_t w[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
            for (int k = 0; k < 4; ++k)   // fp32 fma, then ONE convert-with-ReLU packs both halves
              w[k] = pack_relu_bf16x2(fmaf(bf16lo(w[k]), q[k].x, q[k].y), fmaf(bf16hi(w[k]), q[k].z, q[k].w));
            const bool dead = (taps || ragged) &&
                              (static_cast<unsigned>(ih0[i] + fr) >= static_cast<unsigned>(p.H_in) ||
                               static_cast<unsigned>(iw0[i] + fs) >= static_cast<unsigned>(p.W_in));
            if (dead) w[0] = w[1] = w[2] = w[3] = 0u;
            v[i] = make_uint4(w[0], w[1], w[2], w[3]);
          }
#pragma unroll
          for (int i = 0; i < kRows; ++i) st_shared_v4(a_addr + off[i], v[i].x, v[i].y, v[i].z, v[i].w);
          fence_proxy_async_smem();   // generic-proxy writes -> visible to the tensor core's reads
          __syncwarp();
          if (lane == 0) mbar_arrive(ready_bar(stage));   // one arrival per warp
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
          if (++cb == p.c_blocks) {   // k-steps run over (filter row, filter column, channel block)
            cb = 0;
            if (++fs == p.S) {
              fs = 0;
              ++fr;
            }
          }
        }
      }
    }
  } else if (warp == 2) {
    // ======================= weight pull (parameter-server mode) =======================
    // The all-gather of the SyncReplicas "variable pull" (reference resnet_cifar_main.py:392-396:
    // variables live on the ps tasks, every worker reads them each step) happens HERE, inside the
    // first convolution that consumes the weights: the CTAs split the filter bank between them,
    // wait for the owning shards' epoch flags, and copy their part from the owners' memory over
    // NVLink into the local arena (which the data- and filter-gradient kernels re-use later),
    // while warp 0 already streams activation tiles and the pipeline fills.
    if (p.pull.shadows.p[0] != nullptr) {
      if (lane == 0) wait_shards_ready(p.pull.my_pad, p.pull.map, p.pull.off, p.pull.count);
      __syncwarp();
      const long long g0 = p.pull.off / 8, gn = p.pull.count / 8;
      const long long per_cta = (gn + gridDim.x - 1) / gridDim.x;
      long long gb = g0 + per_cta * blockIdx.x, ge = gb + per_cta;
      if (ge > g0 + gn) ge = g0 + gn;
      if (gb < ge) pull_groups(p.pull.shadows, p.pull.local_arena, p.pull.map, gb, ge, p.pull.self, lane, 32);
      __threadfence();      // my part of the bank is in local memory before I say so
      __syncwarp();
      if (lane == 0) atomicAdd(p.pull.counters, 1u);
    }
  } else if (warp >= 4 && warp < 12) {
    // ======================= epilogue: two groups of 4 warps =======================
    // A tile is drained in kChunks column chunks of kCW (<= 32) columns.  With kChunks >= 2 group
    // 0 takes the even chunks of every tile and group 1 the odd ones; with a single chunk per
    // tile the groups alternate tiles.  Each group owns two staging buffers, a ring of side-input
    // buffers and a named barrier, so one group's TMA waits overlap the other's arithmetic.
    const int grp = (warp - 4) >> 2;
    const int et = threadIdx.x - 128 - grp * kEpiThreads;  // 0..127 inside the group
    const int quad = warp & 3;               // TMEM lane quadrant this warp may read
    const int row = quad * 32 + lane;        // tile row == TMEM lane
    const bool leader = (et == 0);
    const bool has_res = (mode != DRN_EPI_PLAIN);
    const uint32_t bar_id = 1 + grp;
    const uint32_t out_base = sOut + grp * kEpiBufs * C_::kChunkBytes;
    const uint32_t res_base = sRes + grp * p.n_res * C_::kChunkBytes;
    const uint32_t res2_base = sRes2 + grp * p.n_res * C_::kChunkBytes;
    const uint32_t n_res = p.n_res;
    float2* my_ss = sScaleShift + grp * 256;
    float4* my_coef = sCoef + grp * 256;
    constexpr int kHalf = C_::kChunks >= 2 ? C_::kChunks / 2 : 1;  // chunks per tile per group
    // swizzled byte offsets of my row's 16-byte pieces inside a staging buffer (chunk independent)
    uint32_t row_off[kRowBytes / 16];
#pragma unroll
    for (int c16 = 0; c16 < kRowBytes / 16; ++c16)
      row_off[c16] = swz_off(row * kRowBytes + c16 * 16, kRowBytes);
    // statistics pass: thread = (8-column group scg, row slot srs), kNcg rows per thread
    constexpr int kNcg = kCW / 8;
    constexpr int kNrs = kEpiThreads / kNcg;
    const int scg = et % kNcg;
    const int srs = et / kNcg;
    uint32_t st_off[kNcg];
#pragma unroll
    for (int i = 0; i < kNcg; ++i)
      st_off[i] = swz_off((srs + kNrs * i) * kRowBytes + scg * 16, kRowBytes);
    float acc1[kHalf][8], acc2[kHalf][8];
#pragma unroll
    for (int a = 0; a < kHalf; ++a)
#pragma unroll
      for (int k = 0; k < 8; ++k) acc1[a][k] = acc2[a][k] = 0.f;
    int cur_n_tile = -1;

    // Reduce the per-thread partial sums over the group (shuffles, then shared-memory atomics)
    // and add one value per channel to the global statistics.
    // staging of the four warps' partial sums: [warp][a][scg][16] floats per group, in the coefficient
    // cache (only BNBWD_APPLY uses that, and it takes no statistics)
    float* my_red = reinterpret_cast<float*>(sCoef) + grp * 1024;
    auto flush_stats = [&](int n_tile) {
      if (!want_stats || n_tile < 0) return;
      // Every warp reduces its rows with shuffles and parks its per-channel sums in shared memory; after
      // ONE barrier 2*kHalf*kCW/4 threads add the four warps' values and issue one red.global.add.v4.f32
      // each.  (The first version zeroed a buffer, used shared-memory atomics and three barriers: 2 us of
      // a 7 us CIFAR-sized kernel.  Adding every warp's sums straight to global memory removes the last
      // barrier too but quadruples the global adds, which costs 2.7 us on a 128-tile, 64-channel layer.)
      const int wq = (et >> 5);
#pragma unroll
      for (int a = 0; a < kHalf; ++a) {
        float r1[8], r2[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          float v1 = acc1[a][k], v2 = acc2[a][k];
#pragma unroll
          for (int o = kNcg; o < 32; o <<= 1) {
            v1 += __shfl_xor_sync(0xffffffffu, v1, o);
            v2 += __shfl_xor_sync(0xffffffffu, v2, o);
          }
          r1[k] = v1;
          r2[k] = v2;
          acc1[a][k] = acc2[a][k] = 0.f;
        }
        if (lane < kNcg) {   // lane == scg
          float4* dst = reinterpret_cast<float4*>(my_red + ((wq * kHalf + a) * kNcg + scg) * 16);
          dst[0] = make_float4(r1[0], r1[1], r1[2], r1[3]);
          dst[1] = make_float4(r1[4], r1[5], r1[6], r1[7]);
          dst[2] = make_float4(r2[0], r2[1], r2[2], r2[3]);
          dst[3] = make_float4(r2[4], r2[5], r2[6], r2[7]);
        }
      }
      named_bar_sync(bar_id, kEpiThreads);
      // (the buffer is not rewritten before every thread has passed at least one chunk barrier of the
      //  next tile, which the adding threads below reach only after their reads)
      const bool vec_ok = ((reinterpret_cast<uintptr_t>(p.stats) | ((uintptr_t)p.stats_ld * 4)) & 15) == 0;
      if (et < kHalf * kNcg * 4) {
        const int q = et & 3;               // which float4 of the 16: (sum | sum2) x (channels 0-3 | 4-7)
        const int g8 = (et >> 2) % kNcg;    // 8-channel group
        const int a = (et >> 2) / kNcg;
        float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int w = 0; w < kEpiThreads / 32; ++w) {
          const float4 u = *reinterpret_cast<const float4*>(my_red + ((w * kHalf + a) * kNcg + g8) * 16 + q * 4);
          t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
        }
        const int j = C_::kChunks >= 2 ? a * 2 + grp : 0;
        const int col = n_tile * BN + j * kCW + g8 * 8 + (q & 1) * 4;
        float* dst = p.stats + (q >> 1) * p.stats_ld + col;
        if (vec_ok && col + 4 <= p.n_valid) {
          red_add_v4_f32(dst, t.x, t.y, t.z, t.w);
        } else {
          const float tv[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (col + k < p.n_valid) atomicAdd(dst + k, tv[k]);
        }
      }
    };
    // Side-input prefetch cursor (leader thread only): the group's chunks are requested strictly in
    // order, so (tile, chunk, column tile, row index, ring slot) advance incrementally - the leader
    // runs this once per chunk while 127 threads wait for it, no runtime divisions here.
    int pf_tile = t_begin + grp / C_::kChunks, pf_chunk = grp % C_::kChunks;
    int pf_n = pf_tile / p.m_tiles, pf_m = pf_tile - pf_n * p.m_tiles;
    uint32_t pf_slot = 0;
    auto issue_next_res_load = [&]() {
      if (pf_tile < t_end) {
        mbar_arrive_expect_tx(res_bar(grp * kMaxRes + pf_slot), (two_side ? 2 : 1) * C_::kChunkBytes);
        tma_load_2d(res_base + pf_slot * C_::kChunkBytes, &tmapR, res_bar(grp * kMaxRes + pf_slot),
                    pf_n * BN + pf_chunk * kCW, row_tile(pf_m) * kBM);
        if (two_side)   // both side tiles of a chunk land on the same barrier
          tma_load_2d(res2_base + pf_slot * C_::kChunkBytes, &tmapR2, res_bar(grp * kMaxRes + pf_slot),
                      pf_n * BN + pf_chunk * kCW, row_tile(pf_m) * kBM);
      }
      if (++pf_slot == n_res) pf_slot = 0;
      pf_chunk += 2;                       // the other group takes the chunks in between
      while (pf_chunk >= C_::kChunks) {
        pf_chunk -= C_::kChunks;
        ++pf_tile;
        if (++pf_m == p.m_tiles) {
          pf_m = 0;
          ++pf_n;
        }
      }
    };

    uint32_t mine = 0;  // chunks processed by this group so far
    uint32_t rb = 0, rb_phase = 0;  // ring slot / phase of the chunk being processed (= mine % n_res, ...)
    uint32_t ob = 0;                // staging buffer of the chunk being processed (= mine % kEpiBufs)
    const bool mask_mode = mode == DRN_EPI_BNBWD || two_side;   // the side tile is the BatchNorm input
    const bool late_refill = want_stats && (mode == DRN_EPI_BNBWD || mode == DRN_EPI_BNBWD_ADD);
    if (has_res && leader) {
      // fill the ring: chunks 0 .. n_res-2 (chunk i + n_res - 1 is requested when chunk i starts)
      for (uint32_t i = 0; i + 1 < n_res; ++i) issue_next_res_load();
    }

    int acc = 0;
    uint32_t acc_phase = 0;
    int tile_iter = 0;
    for (int t = t_begin; t < t_end; ++t, ++tile_iter) {
      if (t + 1 == t_end && threadIdx.x == 128) pdl_trigger_late();   // this CTA is on its last tile
      const bool my_tile = C_::kChunks >= 2 || (tile_iter & 1) == grp;
      if (my_tile) {
        const int n_tile = t / p.m_tiles;
        const int m_tile = row_tile(t - n_tile * p.m_tiles);
        if (n_tile != cur_n_tile) {
          flush_stats(cur_n_tile);
          cur_n_tile = n_tile;
          if (mask_mode) {
            named_bar_sync(bar_id, kEpiThreads);
            for (int i = et; i < BN; i += kEpiThreads) {
              const int col = n_tile * BN + i;
              my_ss[i] = col < p.n_valid ? p.bn_scale_shift[col] : make_float2(0.f, 0.f);
              if (mode == DRN_EPI_BNBWD_APPLY)
                my_coef[i] = col < p.n_valid ? p.bn_coef[col] : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            named_bar_sync(bar_id, kEpiThreads);
          }
        }
        mbar_wait_sleep(tfull_bar(acc), acc_phase, 0x40 + acc, p.spin_ns);
        tc_fence_after();

#pragma unroll
        for (int a = 0; a < kHalf; ++a) {
          const int j = C_::kChunks >= 2 ? 2 * a + grp : 0;
          const uint32_t out_buf = out_base + ob * C_::kChunkBytes;
          const uint32_t res_buf = res_base + rb * C_::kChunkBytes;
          const uint32_t res2_buf = res2_base + rb * C_::kChunkBytes;
          // request the side input n_res-1 chunks ahead; it lands in the ring slot of the previous
          // chunk, whose main-pass readers all passed that chunk's barrier.  When the statistics read
          // the side tile too (BNBWD), they may still be running: refill after THIS chunk's barrier.
          if (has_res && leader && !late_refill) issue_next_res_load();

          // ---- TMEM -> registers (kCW fp32 columns of my row)
          float v[kCW];
          {
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) +
                                   static_cast<uint32_t>(acc * BN + j * kCW);
            if constexpr (kCW == 16) {
              uint32_t r[16];
              tmem_ld_32x16(taddr, r);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
            } else {
              uint32_t r[32];
              tmem_ld_32x32(taddr, r);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
            }
          }
          // every chunk's 128 threads arrive once: the buffer is free after kChunks*128 arrivals
          tc_fence_before();
          if (TWO) {
            // one (possibly remote) arrival per warp: every lane has finished its tcgen05.ld
            __syncwarp();
            if (lane == 0) {
              if (leader_cta) mbar_arrive(tempty_bar(acc));
              else mbar_arrive_cluster(mapa_shared(tempty_bar(acc), 0));
            }
          } else {
            mbar_arrive(tempty_bar(acc));
          }

          if (has_res) {
            mbar_wait_sleep(res_bar(grp * kMaxRes + rb), rb_phase, 0x50 + grp * kMaxRes + rb, p.spin_ns);
            if (mode == DRN_EPI_RESADD) {
#pragma unroll
              for (int c16 = 0; c16 < kRowBytes / 16; ++c16) {
                const uint4 rv = ld_shared_v4(res_buf + row_off[c16]);
                const uint32_t rr[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  v[c16 * 8 + e * 2] += bf16lo(rr[e]);
                  v[c16 * 8 + e * 2 + 1] += bf16hi(rr[e]);
                }
              }
            } else if (mode == DRN_EPI_BNBWD) {  // keep the gradient where relu(bn(y)) was active
#pragma unroll
              for (int c16 = 0; c16 < kRowBytes / 16; ++c16) {
                const uint4 rv = ld_shared_v4(res_buf + row_off[c16]);
                const uint32_t rr[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const int c0 = c16 * 8 + e * 2;
                  const float2 ss0 = my_ss[j * kCW + c0];
                  const float2 ss1 = my_ss[j * kCW + c0 + 1];
                  v[c0] = (fmaf(bf16lo(rr[e]), ss0.x, ss0.y) > 0.f) ? v[c0] : 0.f;
                  v[c0 + 1] = (fmaf(bf16hi(rr[e]), ss1.x, ss1.y) > 0.f) ? v[c0 + 1] : 0.f;
                }
              }
            } else {
              // two side tiles: y (the BatchNorm input, mask source) and a second tensor that is either
              // added before the mask (gradient arriving through a second branch: BNBWD_ADD, or
              // BNBWD_APPLY with side2_first) or after the BatchNorm-backward formula (the identity
              // shortcut's gradient).  BNBWD_APPLY finishes the BatchNorm backward right here:
              //   out = A * (acc * [relu(bn(y)) active]) + B * y + C  (+ side2)
              const bool apply = mode == DRN_EPI_BNBWD_APPLY;
              const bool add_first = !apply || p.side2_first != 0;
#pragma unroll
              for (int c16 = 0; c16 < kRowBytes / 16; ++c16) {
                const uint4 rv = ld_shared_v4(res_buf + row_off[c16]);
                const uint4 sv = ld_shared_v4(res2_buf + row_off[c16]);
                const uint32_t rr[4] = {rv.x, rv.y, rv.z, rv.w};
                const uint32_t s2[4] = {sv.x, sv.y, sv.z, sv.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
#pragma unroll
                  for (int hh = 0; hh < 2; ++hh) {
                    const int c0 = c16 * 8 + e * 2 + hh;
                    const float yv = hh ? bf16hi(rr[e]) : bf16lo(rr[e]);
                    const float tv = hh ? bf16hi(s2[e]) : bf16lo(s2[e]);
                    const float2 ss0 = my_ss[j * kCW + c0];
                    float g = add_first ? v[c0] + tv : v[c0];
                    g = (fmaf(yv, ss0.x, ss0.y) > 0.f) ? g : 0.f;
                    if (apply) {
                      const float4 cf = my_coef[j * kCW + c0];
                      g = fmaf(cf.x, g, fmaf(cf.y, yv, cf.z));
                      if (!add_first) g += tv;
                    }
                    v[c0] = g;
                  }
                }
              }
            }
          }

          // ---- registers -> swizzled smem (bf16); out_buf was released by the previous chunk's barrier
#pragma unroll
          for (int c16 = 0; c16 < kRowBytes / 16; ++c16) {
            st_shared_v4(out_buf + row_off[c16], pack_bf16x2(v[c16 * 8 + 0], v[c16 * 8 + 1]),
                         pack_bf16x2(v[c16 * 8 + 2], v[c16 * 8 + 3]),
                         pack_bf16x2(v[c16 * 8 + 4], v[c16 * 8 + 5]),
                         pack_bf16x2(v[c16 * 8 + 6], v[c16 * 8 + 7]));
          }
          fence_proxy_async_smem();
          // The ONE barrier of the chunk.  Before it the leader makes sure that every store but the
          // newest has finished reading its staging buffer: that frees the buffer of chunk mine-2,
          // which is the one chunk mine+1 writes (three buffers in rotation).  Everything a thread
          // does after this barrier for chunk `mine` (the statistics pass) reads buffers that are not
          // rewritten before the thread itself has passed the next two barriers.
          if (leader) tma_store_wait_read<1>();
          named_bar_sync(bar_id, kEpiThreads);
          if (leader) {
            if (!p.no_store) {
              tma_store_2d(&tmapC, out_buf, n_tile * BN + j * kCW, m_tile * kBM);
              tma_store_commit();
            }
            if (has_res && late_refill) issue_next_res_load();
          }

          // ---- per-channel statistics of what was just stored (bf16-rounded values)
          if (want_stats) {
            if (mode == DRN_EPI_BNBWD || mode == DRN_EPI_BNBWD_ADD) {
#pragma unroll
              for (int i = 0; i < kNcg; ++i) {
                const uint4 ov = ld_shared_v4(out_buf + st_off[i]);
                const uint4 yv = ld_shared_v4(res_buf + st_off[i]);
                const uint32_t oo[4] = {ov.x, ov.y, ov.z, ov.w};
                const uint32_t yy[4] = {yv.x, yv.y, yv.z, yv.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float o0 = bf16lo(oo[e]), o1 = bf16hi(oo[e]);
                  acc1[a][2 * e] += o0;
                  acc1[a][2 * e + 1] += o1;
                  acc2[a][2 * e] = fmaf(o0, bf16lo(yy[e]), acc2[a][2 * e]);
                  acc2[a][2 * e + 1] = fmaf(o1, bf16hi(yy[e]), acc2[a][2 * e + 1]);
                }
              }
            } else {
#pragma unroll
              for (int i = 0; i < kNcg; ++i) {
                const uint4 ov = ld_shared_v4(out_buf + st_off[i]);
                const uint32_t oo[4] = {ov.x, ov.y, ov.z, ov.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float o0 = bf16lo(oo[e]), o1 = bf16hi(oo[e]);
                  acc1[a][2 * e] += o0;
                  acc1[a][2 * e + 1] += o1;
                  acc2[a][2 * e] = fmaf(o0, o0, acc2[a][2 * e]);
                  acc2[a][2 * e + 1] = fmaf(o1, o1, acc2[a][2 * e + 1]);
                }
              }
            }
          }
          ++mine;
          if (++ob == kEpiBufs) ob = 0;
          if (has_res && ++rb == n_res) {
            rb = 0;
            rb_phase ^= 1;
          }
        }
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    flush_stats(cur_n_tile);
    // the staging buffers must outlive the stores' READS only; completion of the writes is part of
    // grid completion (what the next kernel in the stream, or griddepcontrol.wait, orders behind)
    if (leader) tma_store_wait_read<0>();
  }

  if (p.fin.counter != nullptr) __threadfence();   // my statistics (red.global) before the CTA says "done"
  tc_fence_before();
  if (TWO) cluster_sync_all();   // the leader's MMAs read the peer's shared memory until the very end
  else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if (TWO) tmem_dealloc_2cta(tmem_base, C_::kTmemCols);
    else tmem_dealloc(tmem_base, C_::kTmemCols);
  }
  if (p.fin.counter != nullptr) {
    // the last CTA to get here sees every CTA's sums: it turns them into the consumer's tables
    // (same arithmetic as bn_finalize_kernel in bn.cu)
    __shared__ unsigned int s_last;
    if (threadIdx.x == 0) {
      __threadfence();
      s_last = atomicAdd(p.fin.counter, 1u) == gridDim.x - 1 ? 1u : 0u;
    }
    __syncthreads();
    if (s_last) {
      __threadfence();
      for (int c = threadIdx.x; c < p.n_valid; c += blockDim.x) {
        const float mean = __ldcg(p.stats + c) * p.fin.inv_count;
        float var = fmaf(-mean, mean, __ldcg(p.stats + p.stats_ld + c) * p.fin.inv_count);
        var = fmaxf(var, 0.f);
        const float invstd = rsqrtf(var + p.fin.eps);
        const float sc = p.fin.gamma[c] * invstd;
        p.fin.scale_shift[c] = make_float2(sc, fmaf(-mean, sc, p.fin.beta[c]));
        p.fin.mean_invstd[c] = make_float2(mean, invstd);
        if (p.fin.run_mean != nullptr) {
          p.fin.run_mean[c] -= (p.fin.run_mean[c] - mean) * p.fin.one_minus_decay;
          p.fin.run_var[c] -= (p.fin.run_var[c] - var * p.fin.unbias) * p.fin.one_minus_decay;
        }
      }
      if (threadIdx.x == 0) *p.fin.counter = 0u;
    }
  }
}

inline int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
}

template <int KC, int BN, bool PRE, bool TWO, int EPI>
int launch_instance(int grid, int threads, int smem_bytes, cudaStream_t stream, const CUtensorMap& tA,
                    const CUtensorMap& tB, const CUtensorMap& tC, const CUtensorMap& tR, const CUtensorMap& tR2, const ConvParams& p) {
  static PerDeviceOnce attr_set;
  if (attr_set.need()) {
    DRN_CHECK_CUDA(cudaFuncSetAttribute(conv_igemm_kernel<KC, BN, PRE, TWO, EPI>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBudget));
    attr_set.mark();
  }
  if constexpr (TWO) {
    DRN_CHECK_CUDA(launch_cluster2(conv_igemm_kernel<KC, BN, PRE, TWO, EPI>, dim3(grid), dim3(threads),
                                   smem_bytes, stream, tA, tB, tC, tR, tR2, p));
  } else {
    // size-based policy (experiment): programmatic dependent launch for layers of at most
    // DRN_PDL_SMALL_TILES output tiles - launch-bound layers gain from it, large ones lose
    static const int small_tiles = env_int("DRN_PDL_SMALL_TILES", 0);
    if (small_tiles > 0 && p.total_tiles <= small_tiles) pdl_override() = 1;
    const cudaError_t e = launch_pdl_conv(conv_igemm_kernel<KC, BN, PRE, TWO, EPI>, dim3(grid), dim3(threads),
                                          smem_bytes, stream, tA, tB, tC, tR, tR2, p);
    pdl_override() = -1;
    DRN_CHECK_CUDA(e);
  }
  return 0;
}

// (mode, statistics) combinations the networks use get their own instantiation for the 64-channel
// k-block kernels (ImageNet-sized layers); everything else runs the generic one.
template <int KC, int BN, bool PRE, bool TWO>
int launch_variant(int epi, int grid, int threads, int smem_bytes, cudaStream_t stream, const CUtensorMap& tA,
                   const CUtensorMap& tB, const CUtensorMap& tC, const CUtensorMap& tR, const CUtensorMap& tR2, const ConvParams& p) {
  if constexpr (KC == 64 && BN >= 64) {
    switch (epi) {
      case 0: return launch_instance<KC, BN, PRE, TWO, 0>(grid, threads, smem_bytes, stream, tA, tB, tC, tR, tR2, p);
      case 1: return launch_instance<KC, BN, PRE, TWO, 1>(grid, threads, smem_bytes, stream, tA, tB, tC, tR, tR2, p);
      default: break;
    }
    if (epi == 3) return launch_instance<KC, BN, PRE, TWO, 3>(grid, threads, smem_bytes, stream, tA, tB, tC, tR, tR2, p);
    if constexpr (!PRE) {
      if (epi == 5) return launch_instance<KC, BN, PRE, TWO, 5>(grid, threads, smem_bytes, stream, tA, tB, tC, tR, tR2, p);
    }
  }
  return launch_instance<KC, BN, PRE, TWO, -1>(grid, threads, smem_bytes, stream, tA, tB, tC, tR, tR2, p);
}

template <int KC, int BN, bool TWO = false>
int launch_conv(const drn_conv_desc& d, const void* x, const void* w, void* y, const void* side,
                float* stats, int stats_ld, const void* bn_scale_shift, int mode,
                cudaStream_t stream) {
  using C_ = Cfg<KC, BN, TWO>;
  const int P = d.P, Q = d.Q;
  const long long M = 1LL * d.N * P * Q;
  ConvParams p;
  p.m_tiles = (int)((M + kBM - 1) / kBM);
  if (TWO) p.m_tiles = (p.m_tiles + 1) / 2;   // row-tile pairs (an odd last tile pairs with an empty one)
  p.n_tiles = (d.K + BN - 1) / BN;
  p.total_tiles = p.m_tiles * p.n_tiles;
  p.c_blocks = d.C / KC;
  p.k_iters = d.R * d.S * p.c_blocks;
  p.C = d.C;
  p.S = d.S;
  p.PQ = P * Q;
  p.Q = Q;
  p.stride_h = d.stride;
  p.stride_w = d.stride;
  const ConvView* view = current_view();
  const int pad_w_lo = view ? view->pad_w_lo : d.pad_lo, pad_w_hi = view ? view->pad_w_hi : d.pad_hi;
  long long vstr[3] = {0, 0, 0};
  if (view) { vstr[0] = view->stride_w; vstr[1] = view->stride_h; vstr[2] = view->stride_n; }
  p.pad_h = d.pad_lo;
  p.pad_w = pad_w_lo;
  p.mode = mode;
  // side-input ring: short reductions leave shared memory for a deeper ring (3 chunks of lookahead),
  // long ones need the pipeline stages more (measured, bench/tune_igemm.py)
  const int k_iters_total = d.R * d.S * ((d.C + KC - 1) / KC);
  int n_res_dflt = k_iters_total <= 2 ? 4 : 2;
  // the BN-backward epilogue refills a ring slot half a chunk later (after the chunk's barrier) and
  // streams a wide side tensor: it wants the deeper ring up to longer reductions (measured)
  if (mode == DRN_EPI_BNBWD && (k_iters_total <= 4 || (k_iters_total <= 8 && d.K >= 4 * d.C))) n_res_dflt = 4;
  const bool two_side = mode == DRN_EPI_BNBWD_ADD || mode == DRN_EPI_BNBWD_APPLY;
  const Epi2* e2 = static_cast<const Epi2*>(current_epi2());
  if (two_side) {
    DRN_REQUIRE(e2 != nullptr && e2->side2 != nullptr, "conv_igemm: mode %d needs a second side input", mode);
    DRN_REQUIRE(mode != DRN_EPI_BNBWD_APPLY || e2->coef != nullptr, "conv_igemm: BNBWD_APPLY needs the coefficient table");
    n_res_dflt = env_int("DRN_NRES2", 2);          // two rings per group: keep them short
  }
  p.n_res = mode == DRN_EPI_PLAIN ? 0 : env_int("DRN_NRES", n_res_dflt);
  const int rings = p.n_res * (two_side ? 2 : 1);   // side-tile slots per epilogue group
  p.n_stages = C_::stages_for(rings);
  if (int cap = env_int("DRN_STAGES", 0)) p.n_stages = p.n_stages < cap ? p.n_stages : cap;   // experiments
  if (p.n_stages < 2) return fail("conv_igemm: shared memory plan does not fit (KC=%d BN=%d)", KC, BN);
  const int smem_bytes = C_::smem_for(rings, p.n_stages);
  p.no_store = y == nullptr ? 1 : 0;
  DRN_REQUIRE(!p.no_store || (stats != nullptr && (mode == DRN_EPI_BNBWD || mode == DRN_EPI_BNBWD_ADD)),
              "conv_igemm: y may only be NULL for a statistics-only BNBWD / BNBWD_ADD launch");
  p.bn_coef = e2 != nullptr ? reinterpret_cast<const float4*>(e2->coef) : nullptr;
  p.side2_first = e2 != nullptr ? e2->side2_first : 0;
  p.want_stats = stats != nullptr;
  p.n_valid = d.K;
  p.stats_ld = stats_ld;
  p.stats = stats;
  p.bn_scale_shift = reinterpret_cast<const float2*>(bn_scale_shift);

  CUtensorMap tA, tB, tC, tR, tR2;
  if (int e = tmap_im2col_nhwc(&tA, x, d.N, d.H, d.W, d.C, -d.pad_lo, -pad_w_lo,
                               d.pad_hi - (d.R - 1), pad_w_hi - (d.S - 1), d.stride, d.stride, KC,
                               kBM, KC * 2, view ? vstr : nullptr))
    return e;
  if (int e = tmap_tiled_2d(&tB, w, 2, d.K, (uint64_t)d.R * d.S * d.C, (uint64_t)d.R * d.S * d.C,
                            C_::kBRows, KC, KC * 2))
    return e;
  if (y != nullptr)
    if (int e = tmap_tiled_2d(&tC, y, 2, (uint64_t)M, d.K, d.K, kBM, C_::kCW, C_::kRowBytes)) return e;
  if (mode != DRN_EPI_PLAIN) {
    if (int e = tmap_tiled_2d(&tR, side, 2, (uint64_t)M, d.K, d.K, kBM, C_::kCW, C_::kRowBytes))
      return e;
    if (y == nullptr) tC = tR;
  } else {
    tR = tC;
  }
  if (two_side) {
    if (int e = tmap_tiled_2d(&tR2, e2->side2, 2, (uint64_t)M, d.K, d.K, kBM, C_::kCW, C_::kRowBytes))
      return e;
  } else {
    tR2 = tR;
  }

  // DRN_IGEMM_SMS (experiment): persistent CTAs of the dependent chain on fewer SMs, leaving the rest to
  // the filter gradients of the second stream.  Measured: monotonically slower (14 737 / 14 646 / 14 615 /
  // 14 409 / 14 214 img/s at 148 / 140 / 132 / 124 / 112 SMs)
  int sm_cap = env_int("DRN_IGEMM_SMS", 0);
  if (sm_cap <= 0 || sm_cap > num_sms()) sm_cap = num_sms();
  int grid = p.total_tiles < sm_cap ? p.total_tiles : sm_cap;
  p.pre_ss = reinterpret_cast<const float2*>(current_pre());
  p.H_in = d.H;
  p.W_in = d.W;
  if (current_bnfin() != nullptr && stats != nullptr) p.fin = *static_cast<const BnFin*>(current_bnfin());
  else memset(&p.fin, 0, sizeof(p.fin));
  if (current_pull() != nullptr) p.pull = *static_cast<const PullArgs*>(current_pull());
  else memset(&p.pull, 0, sizeof(p.pull));
  p.M = M;
  p.spin_ns = (unsigned)env_int("DRN_SPIN_NS", 0);
  const int epi = mode * 2 + (stats != nullptr ? 1 : 0);
  if constexpr (TWO) {
    DRN_REQUIRE(p.pre_ss == nullptr, "conv_igemm: no CTA-pair variant of the BatchNorm prologue");
    const int pairs = num_sms() / 2;
    grid = 2 * (p.total_tiles < pairs ? p.total_tiles : pairs);
    return launch_variant<KC, BN, false, true>(epi, grid, kNumThreads, smem_bytes, stream, tA, tB, tC, tR, tR2, p);
  }
  if (p.pre_ss != nullptr) {
    if constexpr ((KC == 64 && (BN == 64 || BN == 128)) || (KC == 32 && (BN == 32 || BN == 64)) ||
                  (KC == 16 && (BN == 16 || BN == 32))) {
      DRN_REQUIRE(view == nullptr, "conv_igemm: the BatchNorm prologue does not combine with an input view");
      return launch_variant<KC, BN, true, false>(epi, grid, kNumThreads + kPreThreads, smem_bytes, stream, tA,
                                                 tB, tC, tR, tR2, p);
    } else {
      return fail("conv_igemm: no BatchNorm-prologue variant for this tile (KC=%d BN=%d): C = 16 with N tiles 16/32, "
                  "C = 32 with 32/64, C %% 64 == 0 with 64/128", KC, BN);
    }
  }
  return launch_variant<KC, BN, false, false>(epi, grid, kNumThreads, smem_bytes, stream, tA, tB, tC, tR, tR2, p);
}

template <int KC>
int dispatch_bn(int bn, const drn_conv_desc& d, const void* x, const void* w, void* y,
                const void* side, float* stats, int stats_ld, const void* ss, int mode,
                cudaStream_t st, bool two = false) {
  if constexpr (KC == 64) {
    if (two && bn == 128) return launch_conv<KC, 128, true>(d, x, w, y, side, stats, stats_ld, ss, mode, st);
    if (two && bn == 256) return launch_conv<KC, 256, true>(d, x, w, y, side, stats, stats_ld, ss, mode, st);
  }
  switch (bn) {
    case 16: return launch_conv<KC, 16>(d, x, w, y, side, stats, stats_ld, ss, mode, st);
    case 32: return launch_conv<KC, 32>(d, x, w, y, side, stats, stats_ld, ss, mode, st);
    case 64: return launch_conv<KC, 64>(d, x, w, y, side, stats, stats_ld, ss, mode, st);
    case 128: return launch_conv<KC, 128>(d, x, w, y, side, stats, stats_ld, ss, mode, st);
    case 256: return launch_conv<KC, 256>(d, x, w, y, side, stats, stats_ld, ss, mode, st);
  }
  return fail("conv_igemm: unsupported N tile %d", bn);
}

}  // namespace

}  // namespace drn

extern "C" int drn_conv_igemm(const drn_conv_desc* desc, const void* x, const void* w, void* y,
                              const void* side, float* stats, int stats_ld,
                              const void* bn_scale_shift, int mode, int force_bn, void* stream) {
  using namespace drn;
  const drn_conv_desc& d = *desc;
  DRN_REQUIRE(d.C == 16 || d.C == 32 || d.C % 64 == 0,
              "conv_igemm: C must be 16, 32 or a multiple of 64 (got %d)", d.C);
  DRN_REQUIRE(d.K % 16 == 0, "conv_igemm: K must be a multiple of 16 (got %d)", d.K);
  DRN_REQUIRE(mode == DRN_EPI_PLAIN || side != nullptr, "conv_igemm: mode %d needs a side input",
              mode);
  DRN_REQUIRE(mode >= DRN_EPI_PLAIN && mode <= DRN_EPI_BNBWD_APPLY, "conv_igemm: unknown epilogue mode %d", mode);
  DRN_REQUIRE(mode < DRN_EPI_BNBWD || bn_scale_shift != nullptr,
              "conv_igemm: the BNBWD modes need bn_scale_shift");
  DRN_REQUIRE(mode != DRN_EPI_BNBWD_APPLY || stats == nullptr, "conv_igemm: BNBWD_APPLY takes no statistics");
  DRN_REQUIRE(stats == nullptr || stats_ld >= d.K, "conv_igemm: stats_ld < K");
  if (stats != nullptr && deterministic()) {
    // Deterministic mode: the epilogue adds its per-channel sums with float atomics in whatever order
    // the CTAs finish.  Run the convolution without them and take the sums of what it stored in an
    // ordered pass (thread-local prologue / view / pull / two-side settings stay in force for the call).
    DRN_REQUIRE(y != nullptr && mode != DRN_EPI_BNBWD_APPLY,
                "conv_igemm: statistics-only launches are not available in the deterministic mode");
    const BnFin* fin = static_cast<const BnFin*>(current_bnfin());
    current_bnfin() = nullptr;
    if (int rc = drn_conv_igemm(desc, x, w, y, side, nullptr, 0, bn_scale_shift, mode, force_bn, stream)) return rc;
    const long long M = 1LL * d.N * d.P * d.Q;
    if (mode == DRN_EPI_BNBWD || mode == DRN_EPI_BNBWD_ADD)   // (sum dm, sum dm * BatchNorm input)
      return drn_bn_pair_reduce(y, side, stats, stats_ld, M, d.K, stream);
    if (int rc = drn_bn_stats(y, stats, stats_ld, M, d.K, stream)) return rc;
    if (fin != nullptr)   // the finalize the last CTA would have done
      return drn_bn_finalize(stats, stats_ld, fin->count, fin->gamma, fin->beta, fin->run_mean, fin->run_var, fin->decay,
                             fin->eps, fin->scale_shift, fin->mean_invstd, d.K, stream);
    return 0;
  }
  int bn = force_bn;
  if (bn == 0) {
    // Measured on B200 over every ResNet-50 role (bench/tune_igemm.py): a 256-wide tile costs about
    // twice a 128-wide one (a little more when a side input is streamed), so it only pays when it
    // saves a partial round of the persistent grid, or when the reduction is a single k-block and
    // re-reading the A tile for a second N tile would double the input traffic.
    const long long m_tiles = (1LL * d.N * d.P * d.Q + 127) / 128;
    const int k_iters = d.R * d.S * ((d.C + 63) / 64);
    const int sms = num_sms();
    if (d.K % 256 == 0) {
      const long long r128 = (m_tiles * (d.K / 128) + sms - 1) / sms;
      const long long r256 = (m_tiles * (d.K / 256) + sms - 1) / sms;
      const double w256 = mode != DRN_EPI_PLAIN ? 2.2 : 1.9;  // side modes: 3 staging buffers + ring leave 256-wide tiles 2 stages
      bn = (k_iters == 1 || r256 * w256 < (double)r128) ? 256 : 128;
      if (current_pre() != nullptr) bn = 128;   // the prologue variant exists for N tiles <= 128
    } else if (d.K % 128 == 0) bn = 128;
    else if (d.K % 64 == 0) bn = 64;
    else if (d.K % 32 == 0) bn = 32;
    else bn = 16;
  }
  DRN_REQUIRE(d.K % bn == 0, "conv_igemm: K=%d not a multiple of the N tile %d", d.K, bn);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int kc = d.C >= 64 ? 64 : d.C;
  switch (kc) {
    case 16: return dispatch_bn<16>(bn, d, x, w, y, side, stats, stats_ld, bn_scale_shift, mode, st);
    case 32: return dispatch_bn<32>(bn, d, x, w, y, side, stats, stats_ld, bn_scale_shift, mode, st);
    default: {
      // CTA pairs (DRN_TWO_CTA: 0 never, 1 policy, 2 always where a variant exists).  Measured
      // (bench/check_two.py): halving the weight traffic pays only where the weights dominate the
      // L2->SM stream - 3x3 over 512 channels at 7x7, 49 -> 41 us; everywhere else it ties with the
      // single-CTA kernel (those layers are bound by their epilogue or by HBM).
      const int two_env = env_int("DRN_TWO_CTA", 1);
      const bool two = current_pre() == nullptr && current_view() == nullptr && (bn == 128 || bn == 256) &&
                       two_env >= 1 && (two_env >= 2 || d.R * d.S * d.C >= 4096);
      return dispatch_bn<64>(bn, d, x, w, y, side, stats, stats_ld, bn_scale_shift, mode, st, two);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Data gradient as a first-class entry point: callers describe the FORWARD convolution and get dx;
// the geometry of the equivalent stride-1 convolution over (zero-inserted) dy is derived here.
// ------------------------------------------------------------------------------------------------
namespace drn {
int dgrad_geometry(const drn_conv_desc& d, drn_conv_desc* g) {
  const int U = (d.P - 1) * d.stride + 1, V = (d.Q - 1) * d.stride + 1;
  const int pad_lo = d.R - 1 - d.pad_lo;
  const int pad_hi_h = d.H - U + d.pad_lo, pad_hi_w = d.W - V + d.pad_lo;
  DRN_REQUIRE(d.R == d.S && pad_hi_h == pad_hi_w, "conv_dgrad: square filters / symmetric geometry only");
  DRN_REQUIRE(pad_lo >= 0 && pad_hi_h >= 0, "conv_dgrad: padding larger than the filter is not supported");
  g->N = d.N; g->H = U; g->W = V; g->C = d.K; g->K = d.C; g->R = d.R; g->S = d.S; g->stride = 1;
  g->pad_lo = pad_lo; g->pad_hi = pad_hi_h; g->P = d.H; g->Q = d.W;
  return 0;
}
}  // namespace drn

extern "C" long long drn_conv_dgrad_scratch_bytes(const drn_conv_desc* desc) {
  if (desc->stride == 1) return 0;
  const long long U = (long long)(desc->P - 1) * desc->stride + 1, V = (long long)(desc->Q - 1) * desc->stride + 1;
  return 2LL * desc->N * U * V * desc->K;
}

extern "C" int drn_conv_dgrad(const drn_conv_desc* desc, const void* dy, const void* w_dgrad, void* dx,
                              const void* side, float* stats, int stats_ld, const void* bn_scale_shift,
                              int mode, void* scratch, long long scratch_bytes, void* stream) {
  using namespace drn;
  drn_conv_desc g;
  if (int e = dgrad_geometry(*desc, &g)) return e;
  const void* src = dy;
  if (desc->stride > 1) {
    const long long need = drn_conv_dgrad_scratch_bytes(desc);
    DRN_REQUIRE(scratch != nullptr && scratch_bytes >= need,
                "conv_dgrad: stride %d needs %lld bytes of scratch (got %lld)", desc->stride, need,
                scratch_bytes);
    if (int e = drn_upsample_zero(dy, scratch, desc->N, desc->P, desc->Q, g.H, g.W, desc->K,
                                  desc->stride, stream))
      return e;
    src = scratch;
  }
  return drn_conv_igemm(&g, src, w_dgrad, dx, side, stats, stats_ld, bn_scale_shift, mode, 0, stream);
}

// Convolution over a strided / overlapping VIEW of the input (strides in elements, separate W
// padding). Used for the space-to-depth ImageNet stem: viewing the zero-bordered [B,112,116,16]
// tensor as [B,112,112,64] with a W stride of 16 elements turns the four horizontal taps of the
// 4x4 filter into one 64-channel (128-byte) TMA row, i.e. a 4x1 convolution with K = 4 x 64.
extern "C" int drn_conv_igemm_view(const drn_conv_desc* desc, const drn_conv_view* view, const void* x,
                                   const void* w, void* y, const void* side, float* stats, int stats_ld,
                                   const void* bn_scale_shift, int mode, int force_bn, void* stream) {
  drn::ConvView v{view->stride_w, view->stride_h, view->stride_n, view->pad_w_lo, view->pad_w_hi};
  drn::current_view() = &v;
  const int rc = drn_conv_igemm(desc, x, w, y, side, stats, stats_ld, bn_scale_shift, mode, force_bn, stream);
  drn::current_view() = nullptr;
  return rc;
}

// y = conv(relu(x * scale + shift), w): the BatchNorm + ReLU that the reference applies in front of
// every convolution (batch_norm_relu, resnet_model_official.py:41-50,148-156) is executed on the A tile
// in shared memory, between the TMA load and the tensor-core read, by four extra warps - the
// activated tensor is never written to memory.  1x1 convolutions with C % 64 == 0 only.
extern "C" int drn_conv_igemm_pre(const drn_conv_desc* desc, const void* x, const void* pre_scale_shift,
                                  const void* w, void* y, const void* side, float* stats, int stats_ld,
                                  const void* bn_scale_shift, int mode, int force_bn, void* stream) {
  if (pre_scale_shift == nullptr) return drn::fail("conv_igemm_pre: pre_scale_shift is NULL");
  if ((reinterpret_cast<uintptr_t>(pre_scale_shift) & 15) != 0)
    return drn::fail("conv_igemm_pre: pre_scale_shift must be 16-byte aligned");
  if (force_bn != 0 && force_bn != 64 && force_bn != 128)
    return drn::fail("conv_igemm_pre: N tile must be 64 or 128");
  drn::current_pre() = pre_scale_shift;
  const int rc = drn_conv_igemm(desc, x, w, y, side, stats, stats_ld, bn_scale_shift, mode, force_bn, stream);
  drn::current_pre() = nullptr;
  return rc;
}

// Data gradient of a 3x3 / stride-2 / pad-(1,1) convolution WITHOUT zero insertion: the four parity
// phases of dx are four small stride-1 convolutions over dy ((1+a) x (1+b) taps, banks from
// drn_weight_prep_dgrad_s2), 9 tap-GEMMs over P*Q pixels in total instead of 9 over H*W.
//   out = [4][N][H/2][W/2][C] bf16, phase (a,b) at index a*2+b holds dx[:, a::2, b::2, :].
// drn_bn_bwd_mask_reduce(addend_kind = 4) interleaves the phases while applying the BN/ReLU mask.
extern "C" int drn_conv_dgrad_s2(const drn_conv_desc* desc, const void* dy, const void* banks, void* out,
                                 void* stream) {
  using namespace drn;
  const drn_conv_desc& d = *desc;
  DRN_REQUIRE(d.R == 3 && d.S == 3 && d.stride == 2 && d.pad_lo == 1 && d.pad_hi == 1 && d.H % 2 == 0 &&
                  d.W % 2 == 0 && d.P == d.H / 2 && d.Q == d.W / 2,
              "conv_dgrad_s2: needs a 3x3 / stride 2 / pad (1,1) convolution over even H and W");
  const long long phase_elems = 1LL * d.N * d.P * d.Q * d.C;
  long long bank_off = 0;
  for (int a = 0; a < 2; ++a) {
    for (int b = 0; b < 2; ++b) {
      drn_conv_desc g;
      g.N = d.N; g.H = d.P; g.W = d.Q; g.C = d.K; g.K = d.C;
      g.R = 1 + a; g.S = 1 + b; g.stride = 1;
      g.pad_lo = 0; g.pad_hi = a;           // rows: dy row i (+ row i+1 for odd output rows)
      g.P = d.P; g.Q = d.Q;
      drn_conv_view v;
      v.stride_w = d.K; v.stride_h = 1LL * d.Q * d.K; v.stride_n = 1LL * d.P * d.Q * d.K;
      v.pad_w_lo = 0; v.pad_w_hi = b;
      const char* bank = static_cast<const char*>(banks) + 2 * bank_off;
      char* dst = static_cast<char*>(out) + 2 * phase_elems * (a * 2 + b);
      if (int e = drn_conv_igemm_view(&g, &v, dy, bank, dst, nullptr, nullptr, 0, nullptr, DRN_EPI_PLAIN, 0, stream))
        return e;
      bank_off += 1LL * (1 + a) * (1 + b) * d.K * d.C;
    }
  }
  return 0;
}

// Superset entry point: optional strided view of x, optional BatchNorm+ReLU prologue, and optionally
// the parameter-server "variable pull" fused in - the kernel first fetches ITS OWN filter bank from
// the shard owners' memory over NVLink (see the PULL comment in the kernel) and then runs as usual.
// This is the all-gather of the SyncReplicas path (reference resnet_cifar_main.py:392-396,
// resnet_model.py:108-112; call site C1 "pull"), placed in the first consumer of the weights.
extern "C" int drn_conv_igemm_ex(const drn_conv_desc* desc, const drn_conv_view* view,
                                 const void* pre_scale_shift, const drn_pull_desc* pull, const void* x,
                                 const void* w, void* y, const void* side, float* stats, int stats_ld,
                                 const void* bn_scale_shift, int mode, int force_bn, void* stream) {
  using namespace drn;
  PullArgs pa;
  memset(&pa, 0, sizeof(pa));
  if (pull != nullptr) {
    DRN_REQUIRE(pull->nranks >= 1 && pull->nranks <= kMaxRanks, "conv_igemm_ex: nranks out of range");
    DRN_REQUIRE(pull->off % 8 == 0 && pull->count % 8 == 0 && pull->bucket_lo % 8 == 0 && pull->per % 8 == 0,
                "conv_igemm_ex: pull ranges must be multiples of 8 elements");
    DRN_REQUIRE(pull->bucket >= 0 && pull->bucket < kMaxBuckets && pull->num_ps >= 1, "conv_igemm_ex: bad shard map");
    DRN_REQUIRE(pull->counters != nullptr && pull->my_pad != nullptr && pull->local_arena != nullptr,
                "conv_igemm_ex: pull needs counters, pad and local arena");
    DRN_REQUIRE(static_cast<const char*>(w) == static_cast<const char*>(pull->local_arena) + 2 * pull->off,
                "conv_igemm_ex: w must be the layer's slot of the local arena");
    for (int i = 0; i < pull->nranks; ++i) pa.shadows.p[i] = const_cast<void*>(pull->shadows[i]);
    pa.my_pad = pull->my_pad;
    pa.local_arena = pull->local_arena;
    pa.counters = pull->counters;
    pa.map.lo = pull->bucket_lo;
    pa.map.per = pull->per;
    pa.map.num_ps = pull->num_ps;
    pa.map.bucket = pull->bucket;
    pa.off = pull->off;
    pa.count = pull->count;
    pa.self = pull->rank;
    current_pull() = &pa;
  }
  ConvView v;
  if (view != nullptr) {
    v = ConvView{view->stride_w, view->stride_h, view->stride_n, view->pad_w_lo, view->pad_w_hi};
    current_view() = &v;
  }
  if (pre_scale_shift != nullptr) current_pre() = pre_scale_shift;
  const int rc = drn_conv_igemm(desc, x, w, y, side, stats, stats_ld, bn_scale_shift, mode, force_bn, stream);
  current_pre() = nullptr;
  current_view() = nullptr;
  current_pull() = nullptr;
  return rc;
}

// drn_conv_igemm_ex + "finalize the BatchNorm whose statistics this convolution accumulates": the last
// CTA of the grid computes (scale, shift), (mean, 1/std) and the moving averages from the completed sums,
// so a prologue-fused consumer can follow without a BatchNorm launch in between.
extern "C" int drn_conv_igemm_fin(const drn_conv_desc* desc, const drn_conv_view* view,
                                  const void* pre_scale_shift, const drn_pull_desc* pull, const drn_bn_fin* fin,
                                  const void* x, const void* w, void* y, const void* side, float* stats,
                                  int stats_ld, const void* bn_scale_shift, int mode, int force_bn, void* stream) {
  using namespace drn;
  BnFin f;
  memset(&f, 0, sizeof(f));
  if (fin != nullptr) {
    DRN_REQUIRE(stats != nullptr && (mode == DRN_EPI_PLAIN || mode == DRN_EPI_RESADD),
                "conv_igemm_fin: the finalize needs the forward statistics epilogue (PLAIN / RESADD with stats)");
    DRN_REQUIRE(fin->counter != nullptr && fin->count > 0 && fin->gamma != nullptr && fin->beta != nullptr &&
                    fin->scale_shift != nullptr && fin->mean_invstd != nullptr,
                "conv_igemm_fin: counter, count, gamma, beta and both output tables are required");
    DRN_REQUIRE((fin->run_mean == nullptr) == (fin->run_var == nullptr), "conv_igemm_fin: run_mean and run_var go together");
    f.counter = fin->counter;
    f.inv_count = (float)(1.0 / (double)fin->count);
    f.unbias = fin->count > 1 ? (float)((double)fin->count / (double)(fin->count - 1)) : 1.f;
    f.one_minus_decay = 1.f - fin->decay;
    f.eps = fin->eps;
    f.gamma = fin->gamma;
    f.beta = fin->beta;
    f.run_mean = fin->run_mean;
    f.run_var = fin->run_var;
    f.scale_shift = static_cast<float2*>(fin->scale_shift);
    f.mean_invstd = static_cast<float2*>(fin->mean_invstd);
    f.count = fin->count;
    f.decay = fin->decay;
    current_bnfin() = &f;
  }
  const int rc = drn_conv_igemm_ex(desc, view, pre_scale_shift, pull, x, w, y, side, stats, stats_ld, bn_scale_shift,
                                   mode, force_bn, stream);
  current_bnfin() = nullptr;
  return rc;
}

// Two-side epilogues (BatchNorm backward finished inside the data-gradient kernel):
//   DRN_EPI_BNBWD_ADD    y = (conv + side2) * [relu(bn(side)) active]; stats += (sum y, sum y*side)
//   DRN_EPI_BNBWD_APPLY  y = A*(conv [+ side2 if side2_first]) * [active] + B*side + C [+ side2 otherwise]
// y may be NULL for a statistics-only launch of BNBWD / BNBWD_ADD (first of the two passes that
// replace "masked gradient to memory, then an element-wise BatchNorm-backward pass").
extern "C" int drn_conv_igemm_epi2(const drn_conv_desc* desc, const void* x, const void* w, void* y,
                                   const void* side, const void* side2, float* stats, int stats_ld,
                                   const void* bn_scale_shift, const void* bn_coef, int side2_first,
                                   int mode, int force_bn, void* stream) {
  drn::Epi2 e2{side2, bn_coef, side2_first};
  drn::current_epi2() = &e2;
  const int rc = drn_conv_igemm(desc, x, w, y, side, stats, stats_ld, bn_scale_shift, mode, force_bn, stream);
  drn::current_epi2() = nullptr;
  return rc;
}
