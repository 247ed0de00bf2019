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
// Synchronisation: flag words in a symmetric "signal pad"; a flag is set by CAS 0->1 on the peer
// and consumed by CAS 1->0 locally, so no epoch has to be passed in (CUDA-graph friendly).  All
// spins are bounded by the %globaltimer watchdog.
#include "common.cuh"
#include "host.cuh"
#include "../include/drn_b200.h"

namespace drn {
namespace {

constexpr int kMaxRanks = 16;
constexpr int kCommThreads = 512;
constexpr int kMaxCommBlocks = 64;

struct PeerPtrs {
  void* p[kMaxRanks];
};

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

// [begin, end) of rank r's share of n4 vectors split into `parts`
__device__ __forceinline__ void slice_of(long long n4, int parts, int r, long long& b,
                                         long long& e) {
  const long long per = (n4 + parts - 1) / parts;
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

// Sharded ("parameter server") update. Shards live on ranks [0, num_ps).
__global__ void __launch_bounds__(kCommThreads)
ps_update_kernel(PeerPtrs grads, PeerPtrs shadows, float4* __restrict__ master,
                 float4* __restrict__ mom, const float* __restrict__ hyper, int rank, int nranks,
                 int num_ps, long long off4, long long n4, PeerPtrs pads, int channel,
                 float* __restrict__ l2_sum) {
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
      float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
      for (int p = 0; p < nranks; ++p) {
        const int src = (rank + p) % nranks;
        float4* gp = reinterpret_cast<float4*>(grads.p[src]) + j;
        const float4 v = ld_f4(gp);
        g.x += v.x;
        g.y += v.y;
        g.z += v.z;
        g.w += v.w;
        st_f4(gp, make_float4(0.f, 0.f, 0.f, 0.f));  // accumulator taken: clear it for the next step
      }
      float4 w = master[j], m = mom[j];
      l2 += w.x * w.x + w.y * w.y + w.z * w.z + w.w * w.w;
      m.x = fmaf(mu, m.x, fmaf(wd, w.x, g.x * inv_n));
      m.y = fmaf(mu, m.y, fmaf(wd, w.y, g.y * inv_n));
      m.z = fmaf(mu, m.z, fmaf(wd, w.z, g.z * inv_n));
      m.w = fmaf(mu, m.w, fmaf(wd, w.w, g.w * inv_n));
      w.x = fmaf(-lr, m.x, w.x);
      w.y = fmaf(-lr, m.y, w.y);
      w.z = fmaf(-lr, m.z, w.z);
      w.w = fmaf(-lr, m.w, w.w);
      master[j] = w;
      mom[j] = m;
      const uint2 packed = make_uint2(pack_bf16x2(w.x, w.y), pack_bf16x2(w.z, w.w));
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
  if (l2_sum != nullptr) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) l2 += __shfl_xor_sync(0xffffffffu, l2, o);
    __shared__ float s[kCommThreads / 32];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = l2;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int i = 0; i < kCommThreads / 32; ++i) t += s[i];
      if (t != 0.f) atomicAdd(l2_sum, t);
    }
  }
  block_barrier(pads, rank, nranks, channel);  // new weights are visible on every worker
}

__global__ void __launch_bounds__(kCommThreads)
bcast_f32_kernel(PeerPtrs bufs, PeerPtrs pads, int rank, int nranks, int root, long long n4,
                 int channel) {
  block_barrier(pads, rank, nranks, channel);
  if (rank == root) {
    const float4* src = reinterpret_cast<const float4*>(bufs.p[root]);
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
      const float4 v = src[i];
      for (int p = 0; p < nranks; ++p)
        if (p != root) st_f4(reinterpret_cast<float4*>(bufs.p[p]) + i, v);
    }
  }
  block_barrier(pads, rank, nranks, channel);
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

extern "C" int drn_comm_pad_bytes(void) {
  return 4 /*channels*/ * kMaxCommBlocks * kMaxRanks * (int)sizeof(uint32_t);
}

// bufs/pads: host arrays of nranks device pointers (index = rank). mc_ptr: multicast address of
// the same symmetric buffer or NULL. Range [offset, offset+count) floats, both multiples of 4.
extern "C" int drn_allreduce_f32(const void* const* bufs, void* mc_ptr, const void* const* pads,
                                 int rank, int nranks, long long offset, long long count,
                                 float scale, int nblocks, int channel, void* stream) {
  DRN_REQUIRE(offset % 4 == 0 && count % 4 == 0, "allreduce: offset/count must be multiples of 4");
  DRN_REQUIRE(channel >= 0 && channel < 4, "allreduce: channel out of range");
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

extern "C" int drn_ps_update(const void* const* grads, const void* const* shadows, float* master,
                             float* mom, const float* hyper, int rank, int nranks, int num_ps,
                             long long offset, long long count, const void* const* pads,
                             float* l2_sum, int nblocks, int channel, void* stream) {
  DRN_REQUIRE(offset % 4 == 0 && count % 4 == 0, "ps_update: offset/count must be multiples of 4");
  DRN_REQUIRE(num_ps >= 1 && num_ps <= nranks, "ps_update: need 1 <= num_ps <= nranks");
  DRN_REQUIRE(channel >= 0 && channel < 4, "ps_update: channel out of range");
  PeerPtrs g, s, p;
  if (int e = fill_ptrs(g, grads, nranks)) return e;
  if (int e = fill_ptrs(s, shadows, nranks)) return e;
  if (int e = fill_ptrs(p, pads, nranks)) return e;
  if (nblocks <= 0) nblocks = 32;
  if (nblocks > kMaxCommBlocks) nblocks = kMaxCommBlocks;
  ps_update_kernel<<<nblocks, kCommThreads, 0, (cudaStream_t)stream>>>(
      g, s, (float4*)master, (float4*)mom, hyper, rank, nranks, num_ps, offset / 4, count / 4, p,
      channel, l2_sum);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int drn_bcast_f32(const void* const* bufs, const void* const* pads, int rank, int nranks,
                             int root, long long count, int nblocks, int channel, void* stream) {
  DRN_REQUIRE(count % 4 == 0, "bcast: count must be a multiple of 4");
  PeerPtrs b, p;
  if (int e = fill_ptrs(b, bufs, nranks)) return e;
  if (int e = fill_ptrs(p, pads, nranks)) return e;
  if (nblocks <= 0) nblocks = 32;
  if (nblocks > kMaxCommBlocks) nblocks = kMaxCommBlocks;
  bcast_f32_kernel<<<nblocks, kCommThreads, 0, (cudaStream_t)stream>>>(b, p, rank, nranks, root,
                                                                       count / 4, channel);
  DRN_CHECK_CUDA(cudaGetLastError());
  return 0;
}

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
