// Layout of the symmetric "signal pad" shared by the collectives (comm.cu) and by the convolution
// kernel that pulls its own weights from the parameter-server shards (conv_igemm.cu, PULL variant),
// plus the device-side helpers both need.
//
//   [0, kBarrierBytes)        flag-word barriers: [channel][block][src rank]            (comm.cu)
//   kReadyOff  uint32[64][16] ready[bucket][owner]: epoch of the newest weights that shard
//                             `owner` has published for `bucket` (written by the owner into EVERY
//                             rank's pad with st.release.sys - the SyncReplicas token, per bucket)
//   kEpochOff  uint32[64]     local: number of parameter-server updates this rank has launched for
//                             the bucket = the epoch its consumers have to see in ready[][]
//   kDoneOff   uint32[64]     local: blocks of the running ps_update kernel that finished their slice
//   kL2Off     float[4]       sum of squared weights over ALL shards (every owner adds its part
//                             into every rank's pad), for the reported cost
#pragma once

#include <stdint.h>

#include "common.cuh"

namespace drn {

constexpr int kMaxRanks = 16;
constexpr int kMaxCommBlocks = 64;
constexpr int kCommChannels = 4;
constexpr int kMaxBuckets = 64;
constexpr int kBarrierBytes = kCommChannels * kMaxCommBlocks * kMaxRanks * 4;   // 16 KB
constexpr int kReadyOff = kBarrierBytes;
constexpr int kEpochOff = kReadyOff + kMaxBuckets * kMaxRanks * 4;
constexpr int kDoneOff = kEpochOff + kMaxBuckets * 4;
constexpr int kL2Off = kDoneOff + kMaxBuckets * 4;
constexpr int kPadBytes = kL2Off + 256;

struct PeerPtrs {
  void* p[kMaxRanks];
};

// How a range of the flat parameter arena is spread over the shard owners: the bucket
// [lo, lo + n) (elements) is cut into num_ps equal slices of `per` elements (a multiple of 8, so
// that a 16-byte group of bf16 weights never straddles two owners).
struct ShardMap {
  long long lo;    // first element of the bucket
  long long per;   // elements per owner
  int num_ps;
  int bucket;      // index into ready[][] / epoch[]
};

__host__ __device__ inline long long shard_per(long long n, int num_ps) {
  const long long per = (n + num_ps - 1) / num_ps;
  return (per + 7) / 8 * 8;
}

__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}

// Wait until every owner whose slice intersects [off, off + count) has published epoch >= the
// epoch this rank expects (one thread polls; bounded by the watchdog).
__device__ __forceinline__ void wait_shards_ready(const void* my_pad, const ShardMap& sm,
                                                  long long off, long long count) {
  const uint8_t* pad = static_cast<const uint8_t*>(my_pad);
  const uint32_t want = *reinterpret_cast<const volatile uint32_t*>(pad + kEpochOff + 4 * sm.bucket);
  int o0 = (int)((off - sm.lo) / sm.per), o1 = (int)((off + count - 1 - sm.lo) / sm.per);
  if (o0 < 0) o0 = 0;
  if (o1 >= sm.num_ps) o1 = sm.num_ps - 1;
  const uint32_t* ready =
      reinterpret_cast<const uint32_t*>(pad + kReadyOff) + sm.bucket * kMaxRanks;
  for (int o = o0; o <= o1; ++o) {
    const uint64_t t0 = globaltimer_ns();
    uint32_t it = 0;
    // epochs are compared modulo 2^32 ("not older than")
    while ((int32_t)(ld_acquire_sys_u32(ready + o) - want) < 0) {
      if (((++it) & 0x3f) == 0 && globaltimer_ns() - t0 > DRN_WATCHDOG_NS) watchdog_fire(0x210);
    }
  }
}

// Copy 16-byte groups [g_begin, g_end) (group index relative to the arena start; one group = 8 bf16
// weights) from the shards' owners into the local working arena.  `tid/nthreads` = cooperating
// threads; four loads are kept in flight per thread (a peer load takes ~2-3 us over NVLink).
__device__ __forceinline__ void pull_groups(const PeerPtrs& shadows, void* local_shadow,
                                            const ShardMap& sm, long long g_begin, long long g_end,
                                            int self, int tid, int nthreads) {
  const long long per_g = sm.per / 8, lo_g = sm.lo / 8;
  auto src_of = [&](long long g) -> const uint4* {
    int owner = (int)((g - lo_g) / per_g);
    if (owner >= sm.num_ps) owner = sm.num_ps - 1;
    return owner == self ? nullptr : static_cast<const uint4*>(shadows.p[owner]) + g;   // my own slice is in place
  };
  uint4* dst = static_cast<uint4*>(local_shadow);
  long long g = g_begin + tid;
  for (; g + 3LL * nthreads < g_end; g += 4LL * nthreads) {
    const uint4* s[4];
    uint4 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) s[k] = src_of(g + (long long)k * nthreads);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (s[k] != nullptr) v[k] = ld_relaxed_sys_v4(s[k]);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (s[k] != nullptr) dst[g + (long long)k * nthreads] = v[k];
  }
  for (; g < g_end; g += nthreads) {
    const uint4* s = src_of(g);
    if (s != nullptr) dst[g] = ld_relaxed_sys_v4(s);
  }
}

// What a convolution needs to fetch its own filter bank from the parameter-server shards.
struct PullArgs {
  PeerPtrs shadows;        // every rank's bf16 working arena (index = rank); p[0] == nullptr: no pull
  const void* my_pad;      // this rank's signal pad (ready[][] / epoch[])
  void* local_arena;       // this rank's working arena (destination)
  unsigned int* counters;  // two zero-initialised words owned by this layer: arrivals, departures
  ShardMap map;
  long long off, count;    // the layer's weights inside the arena (elements, multiples of 8)
  int self;
};

}  // namespace drn
