// In-kernel collectives over NVLink 5 / NVSwitch peer memory (one process per GPU; every rank
// launches the same kernel on its own GPU, buffers are symmetric allocations mapped into all
// peers: parallel/symm.py).
//
//  drn_allreduce_f32   "Horovod path" (reference hvd.DistributedOptimizer, resnet_model.py:114-116,
//                      call site C2): in-place average of a gradient range. One pass: rank r owns a
//                      1/N slice, reads that slice from all N arenas (peer loads or one
//                      multimem.ld_reduce through the switch), scales by 1/N and stores the result
//                      into all N arenas (peer stores or one multimem.st). Reduce-scatter and
//                      all-gather are fused, so there is no intermediate barrier and the same
//                      kernel is latency-optimal for small buckets and bandwidth-optimal for big.
//  drn_ps_update       "parameter-server path" (reference SyncReplicasOptimizer + ps-side
//                      ApplyMomentum, resnet_model.py:102-112; call site C1, ops K19/K20): shard p
//                      of the flat parameter arena lives on GPU p. The owner pulls its slice of all
//                      N workers' gradients over NVLink, averages, applies L2 + momentum-SGD to its
//                      fp32 master/momentum slice and pushes the new bf16 working weights straight
//                      into every worker's arena - push, accumulate, apply and pull in one kernel.
//  drn_bcast_f32       initial broadcast from the chief (BroadcastGlobalVariablesHook, C3).
//  drn_comm_barrier    token-queue equivalent (C1 barrier / W17).
//
//  drn_allreduce_sgd_oneshot  the latency-optimal variant for small buckets (CIFAR's whole 3 MB
//                      gradient, the 32-image ImageNet configuration): EVERY rank reduces the whole
//                      bucket (one multimem.ld_reduce per vector - the switch adds the N copies - or N
//                      peer loads), applies 1/N, L2 and momentum-SGD to its own replica in registers
//                      and writes nothing to its peers; one barrier in front, one behind.
//  drn_ps_sync_vectors gamma / beta / bias of every replica from the bf16 working arena (local after a
//                      push, the owners' shards in pull mode).
//
// Synchronisation: flag words in a symmetric "signal pad"; a flag is set by CAS 0->1 on the peer
// and consumed by CAS 1->0 locally, so no epoch has to be passed in (CUDA-graph friendly).  The
// per-bucket "weights ready" epochs of the pull mode count in device memory (comm_layout.cuh).  All
// spins are bounded by the %globaltimer watchdog.
#include "common.cuh"
#include "comm_layout.cuh"
#include "host.cuh"
#include "../include/drn_b200.h"

namespace drn {
namespace {

constexpr int kCommThreads = 512;

__device__ __forceinline__ void put_signal(uint32_t* addr) {
  uint64_t t0 = globaltimer_ns();
  uint32_t it = 0;
  while (atomicCAS_system(addr, 0u, 1u) != 0u) {
    if (((++it) & 0x3f) == 0 && globaltimer_ns() - t0 > DRN_WATCHDOG_NS) watchdog_fire(0x200);
  }
}
__device__ __forceinline__ void wait_signal(uint32_t* addr) {
  uint64_t t0 = globaltimer_ns();
  uint32_t it = 0;
  while (atomicCAS_system(addr, 1u, 0u) != 1u) {
    if (((++it) & 0x3f) == 0 && globaltimer_ns() - t0 > DRN_WATCHDOG_NS) watchdog_fire(0x201);
  }
}

// Barrier across ranks for one block: pad layout [channel][block][src_rank].
// Every thread's prior global writes are made visible system-wide before the signal is sent.
__device__ __forceinline__ void block_barrier(const PeerPtrs& pads, int rank, int nranks,
                                              int channel) {
  __syncthreads();
  if ((int)threadIdx.x < nranks) {
    __threadfence_system();
    const int peer = threadIdx.x;
    const size_t slot = ((size_t)channel * kMaxCommBlocks + blockIdx.x) * kMaxRanks;
    put_signal(reinterpret_cast<uint32_t*>(pads.p[peer]) + slot + rank);
    wait_signal(reinterpret_cast<uint32_t*>(pads.p[rank]) + slot + peer);
    __threadfence_system();
  }
  __syncthreads();
}

__device__ __forceinline__ float4 ld_f4(const float4* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_f4(float4* p, float4 v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ float4 mc_ld_reduce_f4(const float4* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f4(float4* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// [begin, end) of rank r's share of n4 vectors split into `parts` (shares are a multiple of two
// float4 vectors = 8 elements = one 16-byte group of bf16 weights; shard_per() in comm_layout.cuh)
__device__ __forceinline__ void slice_of(long long n4, int parts, int r, long long& b,
                                         long long& e) {
  const long long per = shard_per(n4 * 4, parts) / 4;
  b = per * r;
  e = b + per;
  if (b > n4) b = n4;
  if (e > n4) e = n4;
}

template <bool kMulticast>
__global__ void __launch_bounds__(kCommThreads)
allreduce_f32_kernel(PeerPtrs bufs, float4* mc, PeerPtrs pads, int rank, int nranks,
                     long long off4, long long n4, float scale, int channel) {
  block_barrier(pads, rank, nranks, channel);  // everybody's gradients are complete
  long long b, e;
  slice_of(n4, nranks, rank, b, e);
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += stride) {
    float4 acc;
    if constexpr (kMulticast) {
      acc = mc_ld_reduce_f4(mc + off4 + i);
    } else {
      acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
      for (int p = 0; p < nranks; ++p) {
        // start with my own arena so that every rank adds in a rotated but fixed order
        const int src = (rank + p) % nranks;
        const float4 v = ld_f4(reinterpret_cast<const float4*>(bufs.p[src]) + off4 + i);
        acc.x += v.x;
        acc.y += v.y;
        acc.z += v.z;
        acc.w += v.w;
      }
    }
    acc.x *= scale;
    acc.y *= scale;
    acc.z *= scale;
    acc.w *= scale;
    if constexpr (kMulticast) {
      mc_st_f4(mc + off4 + i, acc);
    } else {
#pragma unroll 8
      for (int p = 0; p < nranks; ++p) {
        const int dst = (rank + p) % nranks;
        st_f4(reinterpret_cast<float4*>(bufs.p[dst]) + off4 + i, acc);
      }
    }
  }
  block_barrier(pads, rank, nranks, channel);  // all results have landed everywhere
}

__device__ __forceinline__ void mc_st_zero_f4(float4* mc, uint32_t dep) {
  // `dep` is derived from the value the preceding multimem.ld_reduce returned: the clearing store
  // cannot be issued before that load has completed
  const float z = __uint_as_float(dep);
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %1, %1, %1};" ::"l"(mc), "f"(z) : "memory");
}
__device__ __forceinline__ uint32_t zero_after(float v) {
  uint32_t z;
  asm volatile("and.b32 %0, %1, 0;" : "=r"(z) : "r"(__float_as_uint(v)));
  return z;
}

__device__ __forceinline__ void sgd_apply(float4& w, float4& m, const float4& g, float gscale, float lr,
                                          float mu, float wd) {
  m.x = fmaf(mu, m.x, fmaf(wd, w.x, g.x * gscale));
  m.y = fmaf(mu, m.y, fmaf(wd, w.y, g.y * gscale));
  m.z = fmaf(mu, m.z, fmaf(wd, w.z, g.z * gscale));
  m.w = fmaf(mu, m.w, fmaf(wd, w.w, g.w * gscale));
  w.x = fmaf(-lr, m.x, w.x);
  w.y = fmaf(-lr, m.y, w.y);
  w.z = fmaf(-lr, m.z, w.z);
  w.w = fmaf(-lr, m.w, w.w);
}

// block-wide sum of `v`; the result is valid in thread 0
__device__ __forceinline__ float block_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __shared__ float s[kCommThreads / 32];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x == 0)
    for (int i = 0; i < kCommThreads / 32; ++i) t += s[i];
  return t;
}

// Sharded ("parameter server") update. Shards live on ranks [0, num_ps).
//   kMC    the N gradient copies are added inside the switch (multimem.ld_reduce) and cleared
//          with one multimem.st instead of N peer loads + N peer stores
//   kPull  the new bf16 weights stay in the OWNER's working arena; instead of pushing them and
//          closing with a barrier the owner publishes a per-bucket epoch to every rank
//          (ready[bucket][owner], st.release.sys) - the consumers (conv_igemm PULL variant,
//          drn_ps_sync_vectors) fetch the shard over NVLink when they first need it.
template <bool kMC, bool kPull>
__global__ void __launch_bounds__(kCommThreads)
ps_update_kernel(PeerPtrs grads, float4* __restrict__ mc_grads, PeerPtrs shadows,
                 float4* __restrict__ master, float4* __restrict__ mom,
                 const float* __restrict__ hyper, int rank, int nranks, int num_ps, long long off4,
                 long long n4, PeerPtrs pads, int channel, int bucket, int want_l2) {
  uint8_t* my_pad = static_cast<uint8_t*>(pads.p[rank]);
  uint32_t epoch = 0;
  if (kPull) {
    // every rank counts its launches per bucket: the epoch the consumers of this step's weights expect
    uint32_t* ep = reinterpret_cast<uint32_t*>(my_pad + kEpochOff) + bucket;
    epoch = *reinterpret_cast<volatile uint32_t*>(ep) + 1;
  }
  block_barrier(pads, rank, nranks, channel);  // every worker has pushed (finished) its gradients
  float l2 = 0.f;
  if (rank < num_ps) {
    const float lr = hyper[0], mu = hyper[1], wd = hyper[2];
    const float inv_n = 1.f / (float)nranks;
    long long b, e;
    slice_of(n4, num_ps, rank, b, e);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = b + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < e; i += stride) {
      const long long j = off4 + i;
      float4 g;
      if constexpr (kMC) {
        g = mc_ld_reduce_f4(mc_grads + j);
        // pull mode has no closing barrier behind which the replicas could clear their own
        // accumulators: the owner clears all N copies with one multicast store
        if constexpr (kPull) mc_st_zero_f4(mc_grads + j, zero_after(g.x));
      } else {
        g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
        for (int p = 0; p < nranks; ++p) {
          const int src = (rank + p) % nranks;
          float4* gp = reinterpret_cast<float4*>(grads.p[src]) + j;
          const float4 v = ld_f4(gp);
          g.x += v.x;
          g.y += v.y;
          g.z += v.z;
          g.w += v.w;
          if constexpr (kPull) st_f4(gp, make_float4(0.f, 0.f, 0.f, 0.f));  // accumulator taken: clear it
        }
      }
      float4 w = master[j], m = mom[j];
      l2 += w.x * w.x + w.y * w.y + w.z * w.z + w.w * w.w;
      sgd_apply(w, m, g, inv_n, lr, mu, wd);
      master[j] = w;
      mom[j] = m;
      const uint2 packed = make_uint2(pack_bf16x2(w.x, w.y), pack_bf16x2(w.z, w.w));
      if constexpr (kPull) {
        reinterpret_cast<uint2*>(shadows.p[rank])[j] = packed;
      } else {
#pragma unroll 8
        for (int p = 0; p < nranks; ++p) {
          const int dst = (rank + p) % nranks;
          uint2* sp = reinterpret_cast<uint2*>(shadows.p[dst]) + j;
          asm volatile("st.relaxed.sys.global.v2.b32 [%0], {%1, %2};" ::"l"(sp), "r"(packed.x),
                       "r"(packed.y)
                       : "memory");
        }
      }
    }
  }
  if (want_l2) {
    // every owner adds its share of sum(w^2) into EVERY rank's pad, so that each worker reports
    // the full L2 term of the cost (reference: model.cost, resnet_model.py:85-88)
    const float t = block_sum(l2);
    if (threadIdx.x == 0 && t != 0.f)
      for (int p = 0; p < nranks; ++p)
        atomicAdd_system(reinterpret_cast<float*>(static_cast<uint8_t*>(pads.p[p]) + kL2Off), t);
  }
  if constexpr (kPull) {
    // publish: the last block of this rank to finish stores the new epoch locally (every block read
    // the old one at its start, so nobody can see the new value early) and, on an owner, tells
    // every rank that the shard is at `epoch`
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t* done = reinterpret_cast<uint32_t*>(my_pad + kDoneOff) + bucket;
      if (atomicAdd(done, 1u) == gridDim.x - 1) {
        *reinterpret_cast<volatile uint32_t*>(done) = 0u;
        *reinterpret_cast<volatile uint32_t*>(my_pad + kEpochOff + 4 * bucket) = epoch;
        __threadfence_system();
        for (int p = 0; rank < num_ps && p < nranks; ++p) {
          const int dst = (rank + p) % nranks;
          st_release_sys_u32(reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(pads.p[dst]) + kReadyOff) +
                                 bucket * kMaxRanks + rank,
                             epoch);
        }
      }
    }
  } else {
    block_barrier(pads, rank, nranks, channel);  // new weights are visible on every worker
    // every owner has taken its slice of my accumulators: clear my whole range locally (HBM traffic
    // only - clearing from the owners' side would put another (N-1)/N * S bytes on the links)
    float4* mine = reinterpret_cast<float4*>(grads.p[rank]) + off4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride)
      mine[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

// One-shot allreduce fused with the optimizer (small buckets): see the file comment.
template <bool kMulticast>
__global__ void __launch_bounds__(kCommThreads)
allreduce_sgd_oneshot_kernel(PeerPtrs grads, const float4* __restrict__ mc_grads,
                             float4* __restrict__ master, float4* __restrict__ mom,
                             uint2* __restrict__ shadow, const float* __restrict__ hyper, int rank,
                             int nranks, long long off4, long long n4, PeerPtrs pads, int channel,
                             float* __restrict__ l2_sum) {
  block_barrier(pads, rank, nranks, channel);  // everybody's gradients are complete
  const float lr = hyper[0], mu = hyper[1], wd = hyper[2];
  const float inv_n = 1.f / (float)nranks;
  float l2 = 0.f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const long long j = off4 + i;
    float4 g;
    if constexpr (kMulticast) {
      g = mc_ld_reduce_f4(mc_grads + j);
    } else {
      g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
      for (int p = 0; p < nranks; ++p) {   // fixed order 0..N-1: every replica computes the same bits
        const float4 v = ld_f4(reinterpret_cast<const float4*>(grads.p[p]) + j);
        g.x += v.x;
        g.y += v.y;
        g.z += v.z;
        g.w += v.w;
      }
    }
    float4 w = master[j], m = mom[j];
    l2 += w.x * w.x + w.y * w.y + w.z * w.z + w.w * w.w;
    sgd_apply(w, m, g, inv_n, lr, mu, wd);
    master[j] = w;
    mom[j] = m;
    shadow[j] = make_uint2(pack_bf16x2(w.x, w.y), pack_bf16x2(w.z, w.w));
  }
  if (l2_sum != nullptr) {
    const float t = block_sum(l2);
    if (threadIdx.x == 0) atomicAdd(l2_sum, t);
  }
  block_barrier(pads, rank, nranks, channel);  // every peer has read my gradients: clear them
  float4* mine = reinterpret_cast<float4*>(grads.p[rank]) + off4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride)
    mine[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

template <bool kMulticast>
__global__ void __launch_bounds__(kCommThreads)
bcast_f32_kernel(PeerPtrs bufs, float4* __restrict__ mc, PeerPtrs pads, int rank, int nranks, int root,
                 long long n4, int channel) {
  block_barrier(pads, rank, nranks, channel);
  if (rank == root) {
    const float4* src = reinterpret_cast<const float4*>(bufs.p[root]);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
      const float4 v = src[i];
      if constexpr (kMulticast) {
        mc_st_f4(mc + i, v);          // one store, replicated to every replica by the switch
      } else {
        for (int p = 0; p < nranks; ++p)
          if (p != root) st_f4(reinterpret_cast<float4*>(bufs.p[p]) + i, v);
      }
    }
  }
  block_barrier(pads, rank, nranks, channel);
}

// gamma / beta / bias ("vector" variables) of this replica, fp32, from the bf16 working arena: in
// parameter-server mode only the shard owner holds the fp32 master, every replica (the owner
// included, so that all replicas compute with identical values) reads the working copy.
// One block per table entry {int64 elem_off, int32 count, int32 bucket_slot}; count % 8 == 0 is not
// required.  With `pull` the bf16 values come from their owners' arenas (after waiting for the
// shards' epochs) and are also stored into the local arena.
struct VecEntry {
  long long off;
  int count, map;   // map = index into `maps`
};
__global__ void __launch_bounds__(128)
ps_sync_vectors_kernel(PeerPtrs shadows, __nv_bfloat16* __restrict__ local_shadow,
                       float* __restrict__ dst, const VecEntry* __restrict__ table,
                       const ShardMap* __restrict__ maps, const void* my_pad, int self, int pull) {
  const VecEntry e = table[blockIdx.x];
  if (pull) {
    const ShardMap sm = maps[e.map];
    if (threadIdx.x == 0) wait_shards_ready(my_pad, sm, e.off, e.count);
    __syncthreads();
    for (int i = threadIdx.x; i < e.count; i += blockDim.x) {
      const long long idx = e.off + i;
      int owner = (int)((idx - sm.lo) / sm.per);
      if (owner >= sm.num_ps) owner = sm.num_ps - 1;
      const unsigned short raw =
          *reinterpret_cast<const volatile unsigned short*>(static_cast<const __nv_bfloat16*>(shadows.p[owner]) + idx);
      reinterpret_cast<unsigned short*>(local_shadow)[idx] = raw;
      dst[idx] = __uint_as_float((uint32_t)raw << 16);
    }
  } else {
    for (int i = threadIdx.x; i < e.count; i += blockDim.x)
      dst[e.off + i] = __bfloat162float(local_shadow[e.off + i]);
  }
}

__global__ void barrier_kernel(PeerPtrs pads, int rank, int nranks, int channel) {
  block_barrier(pads, rank, nranks, channel);
}

int fill_ptrs(PeerPtrs& out, const void* const* host_ptrs, int nranks) {
  if (nranks < 1 || nranks > kMaxRanks) return fail("comm: nranks %d out of range", nranks);
  for (int i = 0; i < kMaxRanks; ++i) out.p[i] = i < nranks ? const_cast<void*>(host_ptrs[i]) : nullptr;
  return 0;
}

}  // namespace
}  // namespace drn

using namespace drn;

extern "C" int drn_comm_pad_bytes(void) { return kPadBytes; }
// byte offsets inside the pad: which = 0 ready[64][16] (uint32), 1 epoch[64] (uint32), 2 l2 (float)
extern "C" int drn_comm_pad_offset(int which) {
  return which == 0 ? kReadyOff : which == 1 ? kEpochOff : which == 2 ? kL2Off : -1;
}

// bufs/pads: host arrays of nranks device pointers (index = rank). mc_ptr: multicast address of
// the same symmetric buffer or NULL. Range [offset, offset+count) floats, both multiples of 4.
extern "C" int drn_allreduce_f32(const void* const* bufs, void* mc_ptr, const void* const* pads,
                                 int rank, int nranks, long long offset, long long count,
                                 float scale, int nblocks, int channel, void* stream) {
  DRN_REQUIRE(offset % 4 == 0 && count % 4 == 0, "allreduce: offset/count must be multiples of 4");
  DRN_REQUIRE(channel >= 0 && channel < kCommChannels, "allreduce: channel out of range");
  PeerPtrs b, p;
  if (int e = fill_ptrs(b, bufs, nranks)) return e;
  if (int e = fill_ptrs(p, pads, nranks)) return e;
  if (nblocks <= 0) nblocks = 32;
  if (nblocks > kMaxCommBlocks) nblocks = kMaxCommBlocks;
  if (mc_ptr != nullptr)
    allreduce_f32_kernel<true><<<nblocks, kCommThreads, 0, (cudaStream_t)stream>>>(
        b, (float4*)mc_ptr, p, rank, nranks, offset / 4, count / 4, scale, channel);
  else
    allreduce_f32_kernel<false><<<nblocks, kCommThreads, 0, (cudaStream_t)stream>>>(
        b, nullptr, p, rank, nranks, offset / 4, count / 4, scale, channel);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int drn_ps_update(const void* const* grads, void* mc_grads, const void* const* shadows,
                             float* master, float* mom, const float* hyper, int rank, int nranks,
                             int num_ps, long long offset, long long count, const void* const* pads,
                             int want_l2, int pull, int bucket, int nblocks, int channel, void* stream) {
  DRN_REQUIRE(offset % 8 == 0 && count % 4 == 0, "ps_update: offset must be a multiple of 8, count of 4");
  DRN_REQUIRE(num_ps >= 1 && num_ps <= nranks, "ps_update: need 1 <= num_ps <= nranks");
  DRN_REQUIRE(channel >= 0 && channel < kCommChannels, "ps_update: channel out of range");
  DRN_REQUIRE(bucket >= 0 && bucket < kMaxBuckets, "ps_update: bucket index out of range");
  PeerPtrs g, s, p;
  if (int e = fill_ptrs(g, grads, nranks)) return e;
  if (int e = fill_ptrs(s, shadows, nranks)) return e;
  if (int e = fill_ptrs(p, pads, nranks)) return e;
  if (nblocks <= 0) nblocks = 32;
  if (nblocks > kMaxCommBlocks) nblocks = kMaxCommBlocks;
  cudaStream_t st = (cudaStream_t)stream;
  float4* mc = (float4*)mc_grads;
#define DRN_PS_LAUNCH(MC, PULL)                                                                       \
  ps_update_kernel<MC, PULL><<<nblocks, kCommThreads, 0, st>>>(g, mc, s, (float4*)master, (float4*)mom, \
                                                               hyper, rank, nranks, num_ps, offset / 4, \
                                                               count / 4, p, channel, bucket, want_l2)
  if (mc != nullptr && pull) DRN_PS_LAUNCH(true, true);
  else if (mc != nullptr) DRN_PS_LAUNCH(true, false);
  else if (pull) DRN_PS_LAUNCH(false, true);
  else DRN_PS_LAUNCH(false, false);
#undef DRN_PS_LAUNCH
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int drn_allreduce_sgd_oneshot(const void* const* grads, void* mc_grads, float* master,
                                         float* mom, void* shadow, const float* hyper, int rank,
                                         int nranks, long long offset, long long count,
                                         const void* const* pads, float* l2_sum, int nblocks,
                                         int channel, void* stream) {
  DRN_REQUIRE(offset % 4 == 0 && count % 4 == 0, "allreduce_sgd_oneshot: offset/count must be multiples of 4");
  DRN_REQUIRE(channel >= 0 && channel < kCommChannels, "allreduce_sgd_oneshot: channel out of range");
  PeerPtrs g, p;
  if (int e = fill_ptrs(g, grads, nranks)) return e;
  if (int e = fill_ptrs(p, pads, nranks)) return e;
  if (nblocks <= 0) nblocks = 32;
  if (nblocks > kMaxCommBlocks) nblocks = kMaxCommBlocks;
  cudaStream_t st = (cudaStream_t)stream;
  if (mc_grads != nullptr)
    allreduce_sgd_oneshot_kernel<true><<<nblocks, kCommThreads, 0, st>>>(
        g, (const float4*)mc_grads, (float4*)master, (float4*)mom, (uint2*)shadow, hyper, rank, nranks,
        offset / 4, count / 4, p, channel, l2_sum);
  else
    allreduce_sgd_oneshot_kernel<false><<<nblocks, kCommThreads, 0, st>>>(
        g, nullptr, (float4*)master, (float4*)mom, (uint2*)shadow, hyper, rank, nranks, offset / 4,
        count / 4, p, channel, l2_sum);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int drn_bcast_f32(const void* const* bufs, void* mc_ptr, const void* const* pads, int rank,
                             int nranks, int root, long long count, int nblocks, int channel,
                             void* stream) {
  DRN_REQUIRE(count % 4 == 0, "bcast: count must be a multiple of 4");
  PeerPtrs b, p;
  if (int e = fill_ptrs(b, bufs, nranks)) return e;
  if (int e = fill_ptrs(p, pads, nranks)) return e;
  if (nblocks <= 0) nblocks = 32;
  if (nblocks > kMaxCommBlocks) nblocks = kMaxCommBlocks;
  if (mc_ptr != nullptr)
    bcast_f32_kernel<true><<<nblocks, kCommThreads, 0, (cudaStream_t)stream>>>(
        b, (float4*)mc_ptr, p, rank, nranks, root, count / 4, channel);
  else
    bcast_f32_kernel<false><<<nblocks, kCommThreads, 0, (cudaStream_t)stream>>>(
        b, nullptr, p, rank, nranks, root, count / 4, channel);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// table: device array of {int64 elem_off, int32 count, int32 map}; maps: device array of
// {int64 lo, int64 per, int32 num_ps, int32 bucket} (drn_shard_map below builds one on the host)
extern "C" int drn_ps_sync_vectors(const void* const* shadows, void* local_shadow, float* dst_f32,
                                   const void* table, int n_entries, const void* maps,
                                   const void* my_pad, int rank, int nranks, int pull, void* stream) {
  PeerPtrs s;
  if (int e = fill_ptrs(s, shadows, nranks)) return e;
  if (n_entries <= 0) return 0;
  ps_sync_vectors_kernel<<<n_entries, 128, 0, (cudaStream_t)stream>>>(
      s, (__nv_bfloat16*)local_shadow, dst_f32, (const VecEntry*)table, (const ShardMap*)maps, my_pad, rank,
      pull);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// elements per owner for a bucket of `count` elements over num_ps owners (multiple of 8)
extern "C" long long drn_shard_per(long long count, int num_ps) { return shard_per(count, num_ps); }

extern "C" int drn_comm_barrier(const void* const* pads, int rank, int nranks, int channel,
                                void* stream) {
  PeerPtrs p;
  if (int e = fill_ptrs(p, pads, nranks)) return e;
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(p, rank, nranks, channel);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---- CUDA-IPC plumbing for the fallback symmetric allocator (parallel/symm.py) ----------------
// These two helpers DO allocate / map device memory; they run once at start-up.
extern "C" int drn_ipc_alloc(size_t bytes, void** out_ptr, void* out_handle64) {
  void* p = nullptr;
  DRN_CHECK_CUDA(cudaMalloc(&p, bytes));
  DRN_CHECK_CUDA(cudaMemset(p, 0, bytes));
  cudaIpcMemHandle_t h;
  DRN_CHECK_CUDA(cudaIpcGetMemHandle(&h, p));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is expected to be 64 bytes");
  memcpy(out_handle64, &h, 64);
  *out_ptr = p;
  return 0;
}
extern "C" int drn_ipc_open(const void* handle64, void** out_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  DRN_CHECK_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *out_ptr = p;
  return 0;
}
