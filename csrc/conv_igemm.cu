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
// * The epilogue (4 warps) drains TMEM in 64-column chunks through swizzled smem and TMA-stores
//   bf16.  It optionally fuses
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
#include "common.cuh"
#include "host.cuh"
#include "../include/drn_b200.h"

namespace drn {

namespace {

constexpr int kBM = 128;            // output pixels per tile (UMMA M)
constexpr int kEpiBufs = 2;         // double-buffered epilogue chunk buffers
constexpr int kNumThreads = 256;    // warp0 TMA, warp1 MMA, warp2 TMEM alloc, warp3 idle, warps4-7 epilogue
constexpr int kEpiThreads = 128;
constexpr int kMaxStages = 8;

struct ConvParams {
  int m_tiles, n_tiles, total_tiles;
  int k_iters;     // R*S*(C/KC)
  int c_blocks;    // C/KC
  int C;           // input channels (padded)
  int S;           // filter width
  int PQ, Q;       // output plane / width
  int stride_h, stride_w, pad_h, pad_w;
  int mode;        // DRN_EPI_*
  int want_stats;  // accumulate per-channel sums into stats[0][*], stats[1][*]
  int n_valid;     // Cout (columns >= n_valid are padding)
  int stats_ld;    // leading dimension of stats rows
  float* stats;    // [2][stats_ld]
  const float2* bn_scale_shift;  // [Cout] (scale, shift) of the BN whose backward mask is applied
};

template <int KC, int BN>
struct Cfg {
  static constexpr int kABytes = kBM * KC * 2;
  static constexpr int kBBytesRaw = BN * KC * 2;
  static constexpr int kBBytes = (kBBytesRaw + 1023) / 1024 * 1024;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kCW = BN < 64 ? BN : 64;       // epilogue chunk width (columns)
  static constexpr int kChunks = BN / kCW;
  static constexpr int kRowBytes = kCW * 2;            // bytes per staged row (= swizzle span)
  static constexpr int kChunkBytes = kBM * kRowBytes;  // <= 16 KB
  static constexpr int kEpiBytes = 2 * kEpiBufs * kChunkBytes;  // out[2] + res[2]
  static constexpr int kAuxBytes = 2048 + 1024;        // scale/shift cache (256 * 8) + barriers
  static constexpr int kBudget = 222 * 1024;
  static constexpr int kStagesRaw = (kBudget - kEpiBytes - kAuxBytes) / kStageBytes;
  static constexpr int kStages = kStagesRaw > kMaxStages ? kMaxStages : kStagesRaw;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + kAuxBytes + 1024;
  static constexpr int kTmemColsRaw = 2 * BN;
  static constexpr int kTmemCols = kTmemColsRaw <= 32    ? 32
                                   : kTmemColsRaw <= 64  ? 64
                                   : kTmemColsRaw <= 128 ? 128
                                   : kTmemColsRaw <= 256 ? 256
                                                         : 512;
  static_assert(kStages >= 2, "pipeline too shallow");
};

template <int KC, int BN>
__global__ void __launch_bounds__(kNumThreads, 1)
conv_igemm_kernel(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapB,
                  const __grid_constant__ CUtensorMap tmapC, const __grid_constant__ CUtensorMap tmapR,
                  const ConvParams p) {
  using C_ = Cfg<KC, BN>;
  constexpr int kStages = C_::kStages;
  constexpr int kCW = C_::kCW;
  constexpr int kRowBytes = C_::kRowBytes;

  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment: required by the 128B swizzle atom (8 rows x 128 B).
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));

  const uint32_t sStages = smem_base;
  const uint32_t sOut = sStages + kStages * C_::kStageBytes;
  const uint32_t sRes = sOut + kEpiBufs * C_::kChunkBytes;
  const uint32_t sAux = sRes + kEpiBufs * C_::kChunkBytes;
  float2* sScaleShift = reinterpret_cast<float2*>(smem_gen + (sAux - smem_base));  // 256 entries
  const uint32_t sBar = sAux + 2048;
  // barrier map (8 bytes each)
  auto full_bar = [&](int s) { return sBar + 8u * s; };
  auto empty_bar = [&](int s) { return sBar + 8u * (kMaxStages + s); };
  auto tfull_bar = [&](int b) { return sBar + 8u * (2 * kMaxStages + b); };
  auto tempty_bar = [&](int b) { return sBar + 8u * (2 * kMaxStages + 2 + b); };
  auto res_bar = [&](int b) { return sBar + 8u * (2 * kMaxStages + 4 + b); };
  const uint32_t sTmemPtr = sBar + 8u * (2 * kMaxStages + 8);
  volatile uint32_t* tmem_ptr_gen =
      reinterpret_cast<volatile uint32_t*>(smem_gen + (sTmemPtr - smem_base));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmapA);
    tma_prefetch_desc(&tmapB);
    tma_prefetch_desc(&tmapC);
    if (p.mode != DRN_EPI_PLAIN) tma_prefetch_desc(&tmapR);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tfull_bar(b), 1);
      mbar_init(tempty_bar(b), kEpiThreads);
      mbar_init(res_bar(b), 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(sTmemPtr, C_::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_gen;

  if (warp == 0) {
    // ======================= TMA producer =======================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x) {
        const int n_tile = t / p.m_tiles;
        const int m_tile = t - n_tile * p.m_tiles;
        const int m0 = m_tile * kBM;
        const int img = m0 / p.PQ;
        const int rem = m0 - img * p.PQ;
        const int op = rem / p.Q;
        const int oq = rem - op * p.Q;
        const int w0 = oq * p.stride_w - p.pad_w;
        const int h0 = op * p.stride_h - p.pad_h;
        for (int it = 0; it < p.k_iters; ++it) {
          const int tap = it / p.c_blocks;
          const int cb = it - tap * p.c_blocks;
          const int r = tap / p.S;
          const int s = tap - r * p.S;
          mbar_wait(empty_bar(stage), phase ^ 1, 0x10 + stage);
          const uint32_t a_dst = sStages + stage * C_::kStageBytes;
          const uint32_t b_dst = a_dst + C_::kABytes;
          mbar_arrive_expect_tx(full_bar(stage), C_::kABytes + C_::kBBytesRaw);
          tma_load_im2col_4d(a_dst, &tmapA, full_bar(stage), cb * KC, w0, h0, img,
                             static_cast<uint16_t>(s), static_cast<uint16_t>(r));
          tma_load_2d(b_dst, &tmapB, full_bar(stage), tap * p.C + cb * KC, n_tile * BN);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ======================= MMA issuer =======================
    if (elect_one()) {
      constexpr uint32_t kIdesc = make_idesc_bf16(kBM, BN, 0, 0);
      constexpr uint32_t kLayout = umma_layout_for_swizzle_bytes(KC * 2);
      constexpr uint32_t kSBO = 8 * KC * 2;  // 8-row group stride (K-major, dense rows)
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x) {
        mbar_wait(tempty_bar(acc), acc_phase ^ 1, 0x30 + acc);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int it = 0; it < p.k_iters; ++it) {
          mbar_wait(full_bar(stage), phase, 0x20 + stage);
          tc_fence_after();
          const uint32_t a_addr = sStages + stage * C_::kStageBytes;
          const uint32_t b_addr = a_addr + C_::kABytes;
          const uint64_t a_desc = make_smem_desc(a_addr, 0, kSBO, kLayout);
          const uint64_t b_desc = make_smem_desc(b_addr, 0, kSBO, kLayout);
#pragma unroll
          for (int kk = 0; kk < KC / 16; ++kk) {
            // advance 16 elements (32 B) along K inside the swizzle span: +2 in (addr >> 4) units
            umma_f16(d_tmem, a_desc + 2 * kk, b_desc + 2 * kk, kIdesc, (it | kk) != 0 ? 1u : 0u);
          }
          umma_commit(empty_bar(stage));
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(tfull_bar(acc));
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (warp >= 4) {
    // ======================= epilogue =======================
    const int et = threadIdx.x - 128;       // 0..127
    const int quad = warp & 3;              // TMEM lane quadrant this warp may read
    const int row = quad * 32 + lane;       // tile row == TMEM lane
    const bool leader = (et == 0);
    const bool has_res = (p.mode != DRN_EPI_PLAIN);
    constexpr int kTPC = kEpiThreads / kCW;   // threads per column in the statistics pass
    constexpr int kRowsPerThread = kBM / kTPC;
    const int scol = et % kCW;                // my column inside a chunk
    const int srow0 = (et / kCW) * kRowsPerThread;

    float acc_s1[C_::kChunks], acc_s2[C_::kChunks];
#pragma unroll
    for (int j = 0; j < C_::kChunks; ++j) acc_s1[j] = acc_s2[j] = 0.f;
    int cur_n_tile = -1;

    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t g = 0;  // global chunk counter (selects out/res buffer and res barrier phase)

    auto flush_stats = [&](int n_tile) {
      if (!p.want_stats || n_tile < 0) return;
#pragma unroll
      for (int j = 0; j < C_::kChunks; ++j) {
        const int col = n_tile * BN + j * kCW + scol;
        if (col < p.n_valid) {
          atomicAdd(p.stats + col, acc_s1[j]);
          atomicAdd(p.stats + p.stats_ld + col, acc_s2[j]);
        }
        acc_s1[j] = acc_s2[j] = 0.f;
      }
    };
    auto issue_res_load = [&](int tile, int chunk, uint32_t gg) {
      const int n_tile = tile / p.m_tiles;
      const int m_tile = tile - n_tile * p.m_tiles;
      const uint32_t b = gg & 1;
      mbar_arrive_expect_tx(res_bar(b), C_::kChunkBytes);
      tma_load_2d(sRes + b * C_::kChunkBytes, &tmapR, res_bar(b), n_tile * BN + chunk * kCW,
                  m_tile * kBM);
    };

    const int first_tile = blockIdx.x;
    if (has_res && leader && first_tile < p.total_tiles) issue_res_load(first_tile, 0, 0);

    for (int t = first_tile; t < p.total_tiles; t += gridDim.x) {
      const int n_tile = t / p.m_tiles;
      const int m_tile = t - n_tile * p.m_tiles;
      if (n_tile != cur_n_tile) {
        flush_stats(cur_n_tile);
        cur_n_tile = n_tile;
        if (p.mode == DRN_EPI_BNBWD) {
          // refresh the (scale, shift) cache for this column tile
          named_bar_sync(1, kEpiThreads);
          for (int i = et; i < BN; i += kEpiThreads) {
            const int col = n_tile * BN + i;
            sScaleShift[i] = col < p.n_valid ? p.bn_scale_shift[col] : make_float2(0.f, 0.f);
          }
          named_bar_sync(1, kEpiThreads);
        }
      }
      mbar_wait(tfull_bar(acc), acc_phase, 0x40 + acc);
      tc_fence_after();

#pragma unroll 1
      for (int j = 0; j < C_::kChunks; ++j, ++g) {
        const uint32_t b = g & 1;
        const uint32_t out_buf = sOut + b * C_::kChunkBytes;
        const uint32_t res_buf = sRes + b * C_::kChunkBytes;

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
#pragma unroll
            for (int h = 0; h < kCW / 32; ++h) {
              uint32_t r[32];
              tmem_ld_32x32(taddr + h * 32, r);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) v[h * 32 + i] = __uint_as_float(r[i]);
            }
          }
        }
        if (j == C_::kChunks - 1) {
          // accumulator fully drained: hand the TMEM buffer back to the MMA warp
          tc_fence_before();
          mbar_arrive(tempty_bar(acc));
        }

        // ---- out[b] must no longer be read by the store issued two chunks ago; everybody has
        //      finished the statistics pass of the previous chunk after this barrier.
        if (leader) tma_store_wait_read<1>();
        named_bar_sync(1, kEpiThreads);

        // prefetch the side input of the next chunk into res[b^1] (its last reader was chunk g-1)
        if (has_res && leader) {
          int nt = t, nj = j + 1;
          if (nj == C_::kChunks) {
            nj = 0;
            nt = t + gridDim.x;
          }
          if (nt < p.total_tiles) issue_res_load(nt, nj, g + 1);
        }

        if (has_res) {
          mbar_wait(res_bar(b), (g >> 1) & 1, 0x50 + b);
#pragma unroll
          for (int c16 = 0; c16 < kRowBytes / 16; ++c16) {
            const uint32_t off = swz_off(row * kRowBytes + c16 * 16, kRowBytes);
            const uint4 rv = ld_shared_v4(res_buf + off);
            const uint32_t rr[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float y0 = bf16lo(rr[e]), y1 = bf16hi(rr[e]);
              const int c0 = c16 * 8 + e * 2;
              if (p.mode == DRN_EPI_RESADD) {
                v[c0] += y0;
                v[c0 + 1] += y1;
              } else {  // DRN_EPI_BNBWD: keep the gradient where relu(bn(y)) was active
                const float2 ss0 = sScaleShift[j * kCW + c0];
                const float2 ss1 = sScaleShift[j * kCW + c0 + 1];
                v[c0] = (fmaf(y0, ss0.x, ss0.y) > 0.f) ? v[c0] : 0.f;
                v[c0 + 1] = (fmaf(y1, ss1.x, ss1.y) > 0.f) ? v[c0 + 1] : 0.f;
              }
            }
          }
        }

        // ---- registers -> swizzled smem (bf16)
#pragma unroll
        for (int c16 = 0; c16 < kRowBytes / 16; ++c16) {
          const uint32_t off = swz_off(row * kRowBytes + c16 * 16, kRowBytes);
          st_shared_v4(out_buf + off, pack_bf16x2(v[c16 * 8 + 0], v[c16 * 8 + 1]),
                       pack_bf16x2(v[c16 * 8 + 2], v[c16 * 8 + 3]),
                       pack_bf16x2(v[c16 * 8 + 4], v[c16 * 8 + 5]),
                       pack_bf16x2(v[c16 * 8 + 6], v[c16 * 8 + 7]));
        }
        fence_proxy_async_smem();
        named_bar_sync(1, kEpiThreads);
        if (leader) {
          tma_store_2d(&tmapC, out_buf, n_tile * BN + j * kCW, m_tile * kBM);
          tma_store_commit();
        }

        // ---- per-channel statistics of what was just stored (bf16-rounded values)
        if (p.want_stats) {
          float s1 = 0.f, s2 = 0.f;
          const uint32_t cbyte = scol * 2;
#pragma unroll 8
          for (int rr = 0; rr < kRowsPerThread; ++rr) {
            const uint32_t off = swz_off((srow0 + rr) * kRowBytes + cbyte, kRowBytes);
            const float o = __uint_as_float(ld_shared_u16(out_buf + off) << 16);
            s1 += o;
            if (p.mode == DRN_EPI_BNBWD) {
              const float y = __uint_as_float(ld_shared_u16(res_buf + off) << 16);
              s2 = fmaf(o, y, s2);
            } else {
              s2 = fmaf(o, o, s2);
            }
          }
          acc_s1[j] += s1;
          acc_s2[j] += s2;
        }
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    flush_stats(cur_n_tile);
    if (leader) tma_store_wait_all<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, C_::kTmemCols);
  }
}

template <int KC, int BN>
int launch_conv(const drn_conv_desc& d, const void* x, const void* w, void* y, const void* side,
                float* stats, int stats_ld, const void* bn_scale_shift, int mode,
                cudaStream_t stream) {
  using C_ = Cfg<KC, BN>;
  const int P = d.P, Q = d.Q;
  const long long M = 1LL * d.N * P * Q;
  ConvParams p;
  p.m_tiles = (int)((M + kBM - 1) / kBM);
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
  p.pad_h = d.pad_lo;
  p.pad_w = d.pad_lo;
  p.mode = mode;
  p.want_stats = stats != nullptr;
  p.n_valid = d.K;
  p.stats_ld = stats_ld;
  p.stats = stats;
  p.bn_scale_shift = reinterpret_cast<const float2*>(bn_scale_shift);

  CUtensorMap tA, tB, tC, tR;
  if (int e = tmap_im2col_nhwc(&tA, x, d.N, d.H, d.W, d.C, -d.pad_lo, -d.pad_lo,
                               d.pad_hi - (d.R - 1), d.pad_hi - (d.S - 1), d.stride, d.stride, KC,
                               kBM, KC * 2))
    return e;
  if (int e = tmap_tiled_2d(&tB, w, 2, d.K, (uint64_t)d.R * d.S * d.C, (uint64_t)d.R * d.S * d.C,
                            BN, KC, KC * 2))
    return e;
  if (int e = tmap_tiled_2d(&tC, y, 2, (uint64_t)M, d.K, d.K, kBM, C_::kCW, C_::kRowBytes)) return e;
  if (mode != DRN_EPI_PLAIN) {
    if (int e = tmap_tiled_2d(&tR, side, 2, (uint64_t)M, d.K, d.K, kBM, C_::kCW, C_::kRowBytes))
      return e;
  } else {
    tR = tC;
  }

  static bool attr_set = false;
  if (!attr_set) {
    DRN_CHECK_CUDA(cudaFuncSetAttribute(conv_igemm_kernel<KC, BN>,
                                        cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        C_::kSmemBytes));
    attr_set = true;
  }
  int grid = p.total_tiles < num_sms() ? p.total_tiles : num_sms();
  conv_igemm_kernel<KC, BN><<<grid, kNumThreads, C_::kSmemBytes, stream>>>(tA, tB, tC, tR, p);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <int KC>
int dispatch_bn(int bn, const drn_conv_desc& d, const void* x, const void* w, void* y,
                const void* side, float* stats, int stats_ld, const void* ss, int mode,
                cudaStream_t st) {
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
  DRN_REQUIRE(mode != DRN_EPI_BNBWD || bn_scale_shift != nullptr,
              "conv_igemm: BNBWD mode needs bn_scale_shift");
  DRN_REQUIRE(stats == nullptr || stats_ld >= d.K, "conv_igemm: stats_ld < K");
  int bn = force_bn;
  if (bn == 0) {
    if (d.K % 256 == 0) bn = 256;
    else if (d.K % 128 == 0) bn = 128;
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
    default: return dispatch_bn<64>(bn, d, x, w, y, side, stats, stats_ld, bn_scale_shift, mode, st);
  }
}
