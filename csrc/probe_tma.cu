// Measurement probe (not used by the framework): a persistent TMA copy of a [rows][cols] bf16 matrix
// in boxes of 128 rows x box_cols columns - global -> shared (ring of 8 boxes) -> global - to find out
// how the achievable bandwidth of the epilogue's traffic pattern depends on the length of a box row
// (64-byte rows = 32-column chunks, 128-byte rows = 64-column chunks).
#include "common.cuh"
#include "host.cuh"
#include "../include/drn_b200.h"

namespace drn {
namespace {

constexpr int kRing = 8;

__global__ void __launch_bounds__(64, 1)
probe_tma_copy_kernel(const __grid_constant__ CUtensorMap tin, const __grid_constant__ CUtensorMap tout,
                      int row_tiles, int col_tiles, int box_bytes, int box_cols) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar0 = base + kRing * box_bytes;
  const int total = row_tiles * col_tiles;
  if (threadIdx.x == 0) {
    for (int i = 0; i < kRing; ++i) mbar_init(bar0 + 8 * i, 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  // one thread: keep kRing - 2 loads in flight, store each box as it lands
  int issued = 0, done = 0;
  uint32_t phase[kRing] = {0};
  auto tile_of = [&](int k) { return (int)blockIdx.x + k * (int)gridDim.x; };
  auto issue = [&](int k) {
    const int t = tile_of(k), slot = k % kRing;
    mbar_arrive_expect_tx(bar0 + 8 * slot, box_bytes);
    tma_load_2d(base + slot * box_bytes, &tin, bar0 + 8 * slot, (t % col_tiles) * box_cols, (t / col_tiles) * 128);
  };
  int mine = 0;
  for (int t = blockIdx.x; t < total; t += gridDim.x) ++mine;
  while (done < mine) {
    while (issued < mine && issued - done < kRing - 2) {
      // the slot's previous store must have finished reading it
      if (issued >= kRing) tma_store_wait_read<1>();
      issue(issued++);
    }
    const int slot = done % kRing, t = tile_of(done);
    mbar_wait(bar0 + 8 * slot, phase[slot], 0x300);
    phase[slot] ^= 1;
    fence_proxy_async_smem();
    tma_store_2d(&tout, base + slot * box_bytes, (t % col_tiles) * box_cols, (t / col_tiles) * 128);
    tma_store_commit();
    ++done;
  }
  tma_store_wait_read<0>();
}

}  // namespace
}  // namespace drn

extern "C" int drn_probe_tma_copy(const void* src, void* dst, long long rows, int cols, int box_cols, void* stream) {
  using namespace drn;
  DRN_REQUIRE(rows % 128 == 0 && cols % box_cols == 0 && (box_cols == 32 || box_cols == 64),
              "probe_tma_copy: rows %% 128, cols %% box_cols, box_cols 32 or 64");
  CUtensorMap tin, tout;
  if (int e = tmap_tiled_2d(&tin, src, 2, (uint64_t)rows, cols, cols, 128, box_cols, box_cols * 2)) return e;
  if (int e = tmap_tiled_2d(&tout, dst, 2, (uint64_t)rows, cols, cols, 128, box_cols, box_cols * 2)) return e;
  const int box_bytes = 128 * box_cols * 2;
  const int smem = kRing * box_bytes + 1024 + 256;
  static PerDeviceOnce once;
  if (once.need()) {
    DRN_CHECK_CUDA(cudaFuncSetAttribute(probe_tma_copy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    once.mark();
  }
  probe_tma_copy_kernel<<<num_sms(), 64, smem, (cudaStream_t)stream>>>(tin, tout, (int)(rows / 128), cols / box_cols,
                                                                       box_bytes, box_cols);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}
