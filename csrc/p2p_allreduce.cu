// p2p_allreduce.cu — gradient-bucket all-reduce over NVLink peer memory, written for the 8-GPU NVSwitch domain.
//
// Replaces DDP's Reducer + ncclAllReduce on the hot path (hetseq/controller.py:79-87).  Every rank maps every
// peer's gradient arena and flag array (CUDA IPC, set up by ipc_export/ipc_open below) and runs ONE kernel per
// bucket that
//   1. announces "my bucket is final" to every peer (st.release.sys into the peer's flag array),
//   2. waits until every peer has announced the same (ld.acquire.sys on its own, local flag array),
//   3. two-shot: sums its 1/W slice of the bucket over all ranks in fixed rank order (bit-identical result on
//      every rank) reading the peers' slices over NVLink, and writes the sum into every rank's arena
//      (reduce-scatter by peer loads + all-gather by peer stores: 2(W-1)/W * S bytes per direction per GPU);
//      one-shot for small buckets: every rank reads every peer's whole bucket and keeps the sum locally
//      ((W-1) * S inbound, one barrier less);
//   4. signals completion and waits for the peers' completion, so that when the kernel retires the local arena
//      holds the fully reduced bucket and no peer is still reading it.
// No host involvement after set-up; a dead peer turns into a recorded error after a bounded wait instead of a hang.
#include <cuda.h>
#include "common.h"
#include "optim_math.cuh"
#include "../include/b200seq.h"

#include <cstring>

namespace b200 {

constexpr int kMaxRanks = 8;
constexpr int kP2PThreads = 512;
constexpr int kOneShotMaxBlocks = 64;

struct P2PArgs {
  float* data[kMaxRanks];          // peers' arena base pointers (data[rank] is local)
  unsigned int* flags[kMaxRanks];  // peers' flag arrays
  unsigned int* counter;           // local: per-bucket block-arrival counters
  int rank, world;
  long long start, n;              // element range of the bucket inside the arena
  int bucket;                      // flag slot
  int max_buckets;
  unsigned int* epoch_dev;         // local: per-bucket epoch words (advanced by the kernel itself)
  int one_shot;
  int scatter_only;                // reduce-scatter mode: keep the sum of my slice locally, send nothing back
  float* sq_out;                   // local, optional: += sum of squares of the reduced values this rank produced
  float* mc;                       // multicast (NVLS) address of the arena, or null
  int tune;                        // bits 0-3: requests in flight per thread (U), bit 4: block-contiguous iteration
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer_f4(float* p, const float4& v) {
  asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// NVLS: the switch performs the sum over all ranks' copies / replicates a store to all ranks
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__device__ __forceinline__ float block_sum(float v) {
  __shared__ float red[kP2PThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x < 32) {
    t = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  __syncthreads();
  return t;   // valid in thread 0
}

// Watchdog.  A rank that never signals (dead process, stalled host) must not hang its peers for ever, and must not
// take their CUDA contexts down either: after g_wd_cycles SM cycles the waiting thread records
// (0x80000000 | slot << 8 | missing rank) in a word of mapped host memory (b200_p2p_set_watchdog), raises a
// device-wide "tripped" flag and gives up the wait.  The kernel then runs to completion on incomplete data, every
// later wait returns at once, and the host — which looks at the word after each step — stops the run with a
// message that names the missing rank.  No trap, so the context (and the device) stay usable for an orderly exit.
__device__ long long g_wd_cycles = 1200000000000ll;   // ~600 s at 1.965 GHz
__device__ unsigned int* g_wd_err = nullptr;
__device__ unsigned int g_wd_tripped = 0;

__device__ __forceinline__ bool wait_flag(const unsigned int* p, unsigned int epoch, int slot, int peer) {
  const long long t0 = clock64();
  // epochs only grow; (int) difference handles wrap-around
  while ((int)(ld_acquire_sys(p) - epoch) < 0) {
    const bool tripped = *reinterpret_cast<volatile unsigned int*>(&g_wd_tripped) != 0u;
    if (tripped || clock64() - t0 > g_wd_cycles) {
      if (!tripped && atomicExch(&g_wd_tripped, 1u) == 0u && g_wd_err != nullptr) {
        *reinterpret_cast<volatile unsigned int*>(g_wd_err) = 0x80000000u | ((unsigned)slot << 8) | (unsigned)peer;
        __threadfence_system();
      }
      return false;
    }
    __nanosleep(64);
  }
  return true;
}

__device__ __forceinline__ int flag_index(int phase, int bucket, int src, int max_buckets) {
  return (phase * max_buckets + bucket) * kMaxRanks + src;
}

// ---- slice loops of the two-shot / reduce-scatter paths (U independent 16-byte requests in flight per thread) ----
template <int U>
__device__ __forceinline__ float nvls_slice(const P2PArgs& a, long long base, long long step, long long limit) {
  float sq = 0.f;
  for (long long i0 = base; i0 < limit; i0 += U * step) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * step;
      if (i < limit) v[u] = multimem_ld_reduce_f4(a.mc + a.start + (i << 2));
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * step;
      if (i < limit) {
        if (a.scatter_only) *reinterpret_cast<float4*>(a.data[a.rank] + a.start + (i << 2)) = v[u];
        else multimem_st_f4(a.mc + a.start + (i << 2), v[u]);
        sq += v[u].x * v[u].x + v[u].y * v[u].y + v[u].z * v[u].z + v[u].w * v[u].w;
      }
    }
  }
  return sq;
}

template <int U>
__device__ __forceinline__ float peer_slice(const P2PArgs& a, long long base, long long step, long long limit) {
  float sq = 0.f;
  for (long long i0 = base; i0 < limit; i0 += U * step) {
    float4 v[U][kMaxRanks];                   // U x W float4 live: U = 4 would spill at W = 8
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * step;
      if (i < limit) {
        const long long e = a.start + (i << 2);
#pragma unroll
        for (int p = 0; p < kMaxRanks; ++p)
          if (p < a.world) v[u][p] = (p == a.rank) ? *reinterpret_cast<const float4*>(a.data[p] + e) : ld_peer_f4(a.data[p] + e);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * step;
      if (i < limit) {
        const long long e = a.start + (i << 2);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int p = 0; p < kMaxRanks; ++p)      // fixed rank order -> identical bits on every rank
          if (p < a.world) { acc.x += v[u][p].x; acc.y += v[u][p].y; acc.z += v[u][p].z; acc.w += v[u][p].w; }
        if (a.scatter_only) {
          *reinterpret_cast<float4*>(a.data[a.rank] + e) = acc;
        } else {
#pragma unroll
          for (int p = 0; p < kMaxRanks; ++p)
            if (p < a.world) st_peer_f4(a.data[p] + e, acc);
        }
        sq += acc.x * acc.x + acc.y * acc.y + acc.z * acc.z + acc.w * acc.w;
      }
    }
  }
  return sq;
}

__global__ void __launch_bounds__(kP2PThreads)
p2p_allreduce_kernel(const P2PArgs a) {
  const int t = threadIdx.x;
  // Every rank launches the same sequence of reductions, so a per-bucket counter kept on each device yields the
  // same epoch everywhere without a host-supplied value (lets a captured CUDA graph be replayed).  It is advanced
  // by the last block, after every block has read it.
  const unsigned int epoch = a.epoch_dev[a.bucket] + 1u;
  // ---- 1. announce, 2. wait for everyone's announcement --------------------------------------------------
  if (blockIdx.x == 0 && t < a.world && t != a.rank) {
    __threadfence_system();
    st_release_sys(a.flags[t] + flag_index(0, a.bucket, a.rank, a.max_buckets), epoch);
  }
  if (t < a.world && t != a.rank) wait_flag(a.flags[a.rank] + flag_index(0, a.bucket, t, a.max_buckets), epoch, a.bucket, t);
  __syncthreads();

  const long long n4 = a.n >> 2;   // bucket ranges are multiples of 64 elements (arena alignment)
  __shared__ bool last;

  if (a.one_shot) {
    // ---- one-shot (small buckets; host guarantees n4 <= gridDim.x * blockDim.x) ---------------------------
    // Every rank reads every peer's bucket, keeps the sum in registers, and may overwrite its own copy only
    // after ALL ranks have finished reading (second flag phase).
    const long long i = (long long)blockIdx.x * blockDim.x + t;
    const long long e = a.start + (i << 2);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < n4) {
#pragma unroll
      for (int p = 0; p < kMaxRanks; ++p) {
        if (p < a.world) {
          const float4 v = (p == a.rank) ? *reinterpret_cast<const float4*>(a.data[p] + e) : ld_peer_f4(a.data[p] + e);
          acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
      }
    }
    __syncthreads();
    if (t == 0) {
      const unsigned int prev = atomicAdd(a.counter + a.bucket, 1u);
      last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (last) {
      if (t == 0) { a.counter[a.bucket] = 0; a.epoch_dev[a.bucket] = epoch; }
      if (t < a.world && t != a.rank) st_release_sys(a.flags[t] + flag_index(1, a.bucket, a.rank, a.max_buckets), epoch);
    }
    if (t < a.world && t != a.rank) wait_flag(a.flags[a.rank] + flag_index(1, a.bucket, t, a.max_buckets), epoch, a.bucket, t);
    __syncthreads();
    if (i < n4) *reinterpret_cast<float4*>(a.data[a.rank] + e) = acc;
    if (a.sq_out != nullptr && (a.bucket % a.world) == a.rank) {       // one rank accounts for a small bucket's norm
      const float s = block_sum(i < n4 ? acc.x * acc.x + acc.y * acc.y + acc.z * acc.z + acc.w * acc.w : 0.f);
      if (t == 0) atomicAdd(a.sq_out, s);
    }
    return;
  }

  // ---- two-shot: reduce my slice over all ranks, push the result to every rank -----------------------------
  // Iteration space: interleaved (consecutive warps of the whole grid touch consecutive 512 B; a thread's U requests
  // are a grid apart) or block-contiguous (each block streams its own chunk; a thread's U requests are a block apart).
  float sq = 0.f;
  {
    const long long per = ((n4 + a.world - 1) / a.world);
    const long long lo = per * a.rank, hi = (lo + per < n4) ? lo + per : n4;
    long long base, step, limit;
    if (a.tune & 16) {
      const long long chunk = ((hi - lo + gridDim.x - 1) / gridDim.x + 3) & ~3ll;
      const long long c_lo = lo + (long long)blockIdx.x * chunk;
      base = c_lo + t; step = blockDim.x; limit = (c_lo + chunk < hi) ? c_lo + chunk : hi;
    } else {
      base = lo + (long long)blockIdx.x * blockDim.x + t; step = (long long)gridDim.x * blockDim.x; limit = hi;
    }
    const int u = a.tune & 15;
    if (a.mc != nullptr) {
      // NVLS: one multimem.ld_reduce pulls the W-way sum out of the switch, one multimem.st replicates it to all ranks
      sq = u >= 8 ? nvls_slice<8>(a, base, step, limit) : (u >= 4 ? nvls_slice<4>(a, base, step, limit) : nvls_slice<2>(a, base, step, limit));
    } else {
      sq = u >= 2 ? peer_slice<2>(a, base, step, limit) : peer_slice<1>(a, base, step, limit);
    }
  }
  if (a.sq_out != nullptr) {
    const float s = block_sum(sq);
    if (t == 0) atomicAdd(a.sq_out, s);
  }

  // ---- completion: when the kernel retires, every peer has finished reading from and writing to this rank ----
  __threadfence_system();
  __syncthreads();
  if (t == 0) {
    const unsigned int prev = atomicAdd(a.counter + a.bucket, 1u);
    last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (!last) return;
  if (t == 0) { a.counter[a.bucket] = 0; a.epoch_dev[a.bucket] = epoch; }
  // reduce-scatter mode has no second round: nothing was written into a peer, and the peers learn that this rank
  // is done reading THEIR gradients from the next collective it joins (the statistics tail), before any of them
  // touches its gradient arena again.
  if (a.scatter_only) return;
  if (t < a.world && t != a.rank) {
    __threadfence_system();
    st_release_sys(a.flags[t] + flag_index(1, a.bucket, a.rank, a.max_buckets), epoch);
    wait_flag(a.flags[a.rank] + flag_index(1, a.bucket, t, a.max_buckets), epoch, a.bucket, t);
  }
}


// ---- sharded optimizer step: Adam on my slice of every bucket, new parameters written into every rank -------------
// The gradient of my slice is already the sum over ranks (reduce-scatter kernels above) and the global norm is in
// device memory (statistics tail), so this is the third and last stage of the fused path
//   reduce-scatter (per bucket, under backward)  ->  {stats, sum g^2} one-shot  ->  scale / clip / Adam / broadcast.
// Optimizer work per rank is 1/W of the replicated kernel's; what crosses NVLink is the parameters (once, through
// multimem.st where the arena has a multicast mapping) instead of the reduced gradients.
struct ShardedAdamArgs {
  float* param[kMaxRanks];           // fp32 masters of every rank as mapped here (param[rank] local)
  float* param_mc;                   // multicast address of the masters, or null
  __nv_bfloat16* p16[kMaxRanks];     // bf16 working copies (null in fp32/tf32 mode)
  __nv_bfloat16* p16_mc;
  float* grad; float* m; float* v;   // local
  const long long* ranges;           // device: {start, n} per bucket
  int n_buckets;
  unsigned int* flags[kMaxRanks];
  unsigned int* counter;
  int rank, world, bucket, max_buckets;
  unsigned int* epoch_dev;
  const float* gate;                 // optional device flag: 0 -> nothing to apply (only the flag round runs)
  AdamArgsDev adam;
};

__device__ __forceinline__ void multimem_st_bf16x4(__nv_bfloat16* mc, uint32_t lo, uint32_t hi) {
  asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1, %2};" ::"l"(mc), "r"(lo), "r"(hi) : "memory");
}

__global__ void __launch_bounds__(256)
p2p_sharded_adam_kernel(const ShardedAdamArgs a) {
  const int t = threadIdx.x;
  const unsigned int epoch = a.epoch_dev[a.bucket] + 1u;
  AdamArgsDev ad = a.adam;
  float decay = 1.f - ad.weight_decay * ad.lr;
  if (ad.hyper_dev != nullptr) {
    ad.lr = ad.hyper_dev[0];
    ad.step_size = ad.hyper_dev[1];
    decay = ad.hyper_dev[2];
  }
  const StepScalars sc = derive_scalars(ad.mult, ad.sample_size, ad.sample_size_f32, ad.norm_sq, ad.max_norm);
  if (blockIdx.x == 0 && t == 0 && ad.stats_out != nullptr) {
    ad.stats_out[0] = sc.gnorm;
    ad.stats_out[1] = sc.scale;
    ad.stats_out[2] = sc.clipped;
  }
  const float one_m_b1 = 1.f - ad.beta1, one_m_b2 = 1.f - ad.beta2;
  const bool decay_on = ad.weight_decay != 0.f;
  const long long gtid = (long long)blockIdx.x * blockDim.x + t;
  const long long gstride = (long long)gridDim.x * blockDim.x;
  const bool bf16_mode = a.p16[a.rank] != nullptr;
  float* const p_local = a.param[a.rank];

  // a deferred update that was already flushed leaves the gate at 0: the launch still takes part in the flag round
  // (every rank launches the same sequence of kernels) but touches no data
  const int nb = (a.gate == nullptr || *a.gate != 0.f) ? a.n_buckets : 0;
  for (int b = 0; b < nb; ++b) {
    const long long start = a.ranges[2 * b], n4 = a.ranges[2 * b + 1] >> 2;
    const long long per = (n4 + a.world - 1) / a.world;
    const long long lo = per * a.rank, hi = (lo + per < n4) ? lo + per : n4;
    for (long long i = lo + gtid; i < hi; i += gstride) {
      const long long e = start + (i << 2);
      float4 pv = *reinterpret_cast<const float4*>(p_local + e);
      const float4 gv = *reinterpret_cast<const float4*>(a.grad + e);
      float4 mv = *reinterpret_cast<const float4*>(a.m + e);
      float4 vv = *reinterpret_cast<const float4*>(a.v + e);
      adam_element(pv.x, gv.x, mv.x, vv.x, sc.scale, ad.beta1, ad.beta2, one_m_b1, one_m_b2, ad.eps, ad.step_size, decay_on, decay);
      adam_element(pv.y, gv.y, mv.y, vv.y, sc.scale, ad.beta1, ad.beta2, one_m_b1, one_m_b2, ad.eps, ad.step_size, decay_on, decay);
      adam_element(pv.z, gv.z, mv.z, vv.z, sc.scale, ad.beta1, ad.beta2, one_m_b1, one_m_b2, ad.eps, ad.step_size, decay_on, decay);
      adam_element(pv.w, gv.w, mv.w, vv.w, sc.scale, ad.beta1, ad.beta2, one_m_b1, one_m_b2, ad.eps, ad.step_size, decay_on, decay);
      *reinterpret_cast<float4*>(a.m + e) = mv;
      *reinterpret_cast<float4*>(a.v + e) = vv;
      *reinterpret_cast<float4*>(a.grad + e) = make_float4(0.f, 0.f, 0.f, 0.f);
      if (bf16_mode) {
        // the fp32 master of a slice lives with its owner only; the bf16 working copy is what every rank computes with
        *reinterpret_cast<float4*>(p_local + e) = pv;
        __nv_bfloat162 q0 = __floats2bfloat162_rn(pv.x, pv.y), q1 = __floats2bfloat162_rn(pv.z, pv.w);
        const uint32_t w0 = *reinterpret_cast<uint32_t*>(&q0), w1 = *reinterpret_cast<uint32_t*>(&q1);
        if (a.p16_mc != nullptr) {
          multimem_st_bf16x4(a.p16_mc + e, w0, w1);
        } else {
#pragma unroll
          for (int p = 0; p < kMaxRanks; ++p)
            if (p < a.world) *reinterpret_cast<uint2*>(a.p16[p] + e) = make_uint2(w0, w1);
        }
      } else if (a.param_mc != nullptr) {
        multimem_st_f4(a.param_mc + e, pv);
      } else {
#pragma unroll
        for (int p = 0; p < kMaxRanks; ++p)
          if (p < a.world) st_peer_f4(a.param[p] + e, pv);
      }
    }
    // gradients outside my slice were only inputs to the peers' reductions: clear them for the next backward
    for (long long i = gtid; i < n4; i += gstride)
      if (i < lo || i >= hi) *reinterpret_cast<float4*>(a.grad + start + (i << 2)) = make_float4(0.f, 0.f, 0.f, 0.f);
  }

  // ---- completion round: when this kernel retires, every peer's slice has landed in this rank's parameters ----
  __shared__ bool last;
  __threadfence_system();
  __syncthreads();
  if (t == 0) {
    const unsigned int prev = atomicAdd(a.counter + a.bucket, 1u);
    last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (!last) return;
  if (t == 0) { a.counter[a.bucket] = 0; a.epoch_dev[a.bucket] = epoch; }
  if (t < a.world && t != a.rank) {
    __threadfence_system();
    st_release_sys(a.flags[t] + flag_index(1, a.bucket, a.rank, a.max_buckets), epoch);
    wait_flag(a.flags[a.rank] + flag_index(1, a.bucket, t, a.max_buckets), epoch, a.bucket, t);
  }
}

}  // namespace b200

using namespace b200;

// ---- CUDA IPC plumbing (host) -------------------------------------------------------------------------------
// export: 64-byte handle of the allocation that contains `ptr` + the byte offset of `ptr` inside it
extern "C" int b200_ipc_export(const void* ptr, unsigned char* handle_out /*64*/, long long* offset_out) {
  if (!ptr || !handle_out || !offset_out) return fail(B200_ERR_ARG, "ipc_export: null pointer");
  cudaIpcMemHandle_t h;
  B200_CUDA(cudaIpcGetMemHandle(&h, const_cast<void*>(ptr)));
  // the handle names the whole allocation; find its base to compute the offset
  typedef CUresult (*GetRangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);
  static GetRangeFn get_range = nullptr;
  if (!get_range) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &p, cudaEnableDefault, &q) != cudaSuccess || !p)
      return fail(B200_ERR_DRIVER, "ipc_export: cuMemGetAddressRange unavailable");
    get_range = reinterpret_cast<GetRangeFn>(p);
  }
  CUdeviceptr base = 0;
  size_t size = 0;
  if (get_range(&base, &size, reinterpret_cast<CUdeviceptr>(ptr)) != CUDA_SUCCESS)
    return fail(B200_ERR_DRIVER, "ipc_export: cuMemGetAddressRange failed");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  std::memcpy(handle_out, &h, 64);
  *offset_out = (long long)(reinterpret_cast<CUdeviceptr>(ptr) - base);
  return 0;
}

extern "C" int b200_ipc_open(const unsigned char* handle /*64*/, long long offset, void** ptr_out, void** base_out) {
  if (!handle || !ptr_out) return fail(B200_ERR_ARG, "ipc_open: null pointer");
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, 64);
  void* base = nullptr;
  B200_CUDA(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr_out = static_cast<char*>(base) + offset;
  if (base_out) *base_out = base;
  return 0;
}

extern "C" int b200_ipc_close(void* base) {
  if (!base) return 0;
  B200_CUDA(cudaIpcCloseMemHandle(base));
  return 0;
}

static int launch_reduce(void* const* data_ptrs, void* const* flag_ptrs, void* counter, int rank, int world,
                         long long start, long long n, int bucket, int max_buckets, void* epoch_dev, int one_shot,
                         int scatter_only, int blocks, float* sq_out, void* multicast_base, void* stream, const char* who,
                         int tune = 0) {
  if (!data_ptrs || !flag_ptrs || !counter || !epoch_dev) return fail(B200_ERR_ARG, "%s: null pointer", who);
  if (world < 2 || world > kMaxRanks || rank < 0 || rank >= world) return fail(B200_ERR_ARG, "%s: bad rank/world %d/%d", who, rank, world);
  if ((start & 3) || (n & 3)) return fail(B200_ERR_ARG, "%s: range must be a multiple of 4 elements", who);
  if (bucket < 0 || bucket >= max_buckets) return fail(B200_ERR_ARG, "%s: bucket index out of range", who);
  P2PArgs a;
  for (int i = 0; i < kMaxRanks; ++i) {
    a.data[i] = i < world ? static_cast<float*>(data_ptrs[i]) : nullptr;
    a.flags[i] = i < world ? static_cast<unsigned int*>(flag_ptrs[i]) : nullptr;
  }
  a.counter = static_cast<unsigned int*>(counter);
  a.rank = rank; a.world = world; a.start = start; a.n = n; a.bucket = bucket; a.max_buckets = max_buckets;
  a.epoch_dev = static_cast<unsigned int*>(epoch_dev); a.one_shot = one_shot; a.scatter_only = scatter_only;
  a.sq_out = sq_out;
  a.mc = one_shot ? nullptr : static_cast<float*>(multicast_base);
  if ((tune & 15) == 0) tune |= a.mc ? 4 : 2;
  a.tune = tune;
  if (blocks <= 0) blocks = 32;
  // blocks >= 1000 encodes a narrower CTA: value = threads * 1000 + blocks (threads a multiple of 32, <= 512).
  // Many narrow CTAs spread the copy warps over every SM instead of loading a few SMs heavily, which matters
  // while a persistent GEMM (one CTA per SM, static tile schedule) is running next to the reduction.
  int threads = kP2PThreads;
  if (blocks >= 1000) {
    threads = blocks / 1000;
    blocks %= 1000;
    if (threads < 32 || threads > kP2PThreads || (threads & 31) || blocks < 1) return fail(B200_ERR_ARG, "%s: bad threads/blocks encoding", who);
  }
  if (one_shot) {
    // Every block of a one-shot launch spins on a flag that a peer raises only after ALL of its blocks have
    // arrived, so the whole grid has to be resident at once — also next to a persistent GEMM that leaves room for
    // at most a few extra 512-thread CTAs per SM.  64 blocks (128 KB per rank) stay far below that.
    threads = kP2PThreads;
    const long long need = ((n >> 2) + kP2PThreads - 1) / kP2PThreads;
    if (need > kOneShotMaxBlocks) return fail(B200_ERR_ARG, "%s: bucket too large for one-shot mode (%lld blocks > %d)", who, need, kOneShotMaxBlocks);
    blocks = (int)(need > 0 ? need : 1);
  }
  p2p_allreduce_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(a);
  B200_LAUNCHED("p2p_allreduce_kernel");
  return 0;
}

extern "C" int b200_p2p_allreduce(void* const* data_ptrs, void* const* flag_ptrs, void* counter, int rank, int world,
                                  long long start, long long n, int bucket, int max_buckets, void* epoch_dev,
                                  int one_shot, int blocks, float* sq_out, void* multicast_base, void* stream) {
  const char* e = getenv("B200_P2P_TUNE");
  return launch_reduce(data_ptrs, flag_ptrs, counter, rank, world, start, n, bucket, max_buckets, epoch_dev, one_shot, 0,
                       blocks, sq_out, multicast_base, stream, "p2p_allreduce", e ? atoi(e) : 0);
}

extern "C" int b200_p2p_reduce_scatter(void* const* data_ptrs, void* const* flag_ptrs, void* counter, int rank, int world,
                                       long long start, long long n, int bucket, int max_buckets, void* epoch_dev,
                                       int blocks, float* sq_out, void* multicast_base, void* stream) {
  // B200_P2P_TUNE = U + 16 * contiguous (see P2PArgs::tune), read per launch so that the probe tool can sweep it
  const char* e = getenv("B200_P2P_TUNE");
  return launch_reduce(data_ptrs, flag_ptrs, counter, rank, world, start, n, bucket, max_buckets, epoch_dev, 0, 1,
                       blocks, sq_out, multicast_base, stream, "p2p_reduce_scatter", e ? atoi(e) : 0);
}

extern "C" int b200_p2p_sharded_adam(void* const* param_ptrs, void* param_mc, void* const* bf16_ptrs, void* bf16_mc,
                                     float* grad, float* m, float* v, const long long* ranges, int n_buckets, long long total,
                                     void* const* flag_ptrs, void* counter, int rank, int world, int bucket, int max_buckets,
                                     void* epoch_dev, const b200_adam_args* args, int blocks, const float* gate, void* stream) {
  if (!param_ptrs || !grad || !m || !v || !ranges || !flag_ptrs || !counter || !epoch_dev || !args)
    return fail(B200_ERR_ARG, "p2p_sharded_adam: null pointer");
  if (world < 2 || world > kMaxRanks || rank < 0 || rank >= world) return fail(B200_ERR_ARG, "p2p_sharded_adam: bad rank/world %d/%d", rank, world);
  if (bucket < 0 || bucket >= max_buckets || n_buckets < 1) return fail(B200_ERR_ARG, "p2p_sharded_adam: bad bucket arguments");
  if (total & 3) return fail(B200_ERR_ARG, "p2p_sharded_adam: arena size must be a multiple of 4 elements");
  ShardedAdamArgs a;
  for (int i = 0; i < kMaxRanks; ++i) {
    a.param[i] = i < world ? static_cast<float*>(param_ptrs[i]) : nullptr;
    a.p16[i] = (bf16_ptrs && i < world) ? static_cast<__nv_bfloat16*>(bf16_ptrs[i]) : nullptr;
    a.flags[i] = i < world ? static_cast<unsigned int*>(flag_ptrs[i]) : nullptr;
  }
  if (!a.param[rank]) return fail(B200_ERR_ARG, "p2p_sharded_adam: local parameter pointer missing");
  if (((reinterpret_cast<uintptr_t>(a.param[rank]) | reinterpret_cast<uintptr_t>(grad) | reinterpret_cast<uintptr_t>(m) |
        reinterpret_cast<uintptr_t>(v)) & 15) != 0)
    return fail(B200_ERR_ARG, "p2p_sharded_adam: buffers must be 16-byte aligned");
  a.param_mc = static_cast<float*>(param_mc);
  a.p16_mc = bf16_ptrs ? static_cast<__nv_bfloat16*>(bf16_mc) : nullptr;
  a.grad = grad; a.m = m; a.v = v; a.ranges = ranges; a.n_buckets = n_buckets;
  a.counter = static_cast<unsigned int*>(counter);
  a.rank = rank; a.world = world; a.bucket = bucket; a.max_buckets = max_buckets;
  a.epoch_dev = static_cast<unsigned int*>(epoch_dev);
  a.gate = gate;
  fill_adam_args(args, a.adam);
  if (blocks <= 0) blocks = num_sms() * 8;
  p2p_sharded_adam_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(a);
  B200_LAUNCHED("p2p_sharded_adam_kernel");
  return 0;
}

extern "C" int b200_p2p_set_watchdog(double timeout_seconds, unsigned int* host_error_word_devptr) {
  if (!(timeout_seconds > 0)) return fail(B200_ERR_ARG, "p2p_set_watchdog: timeout must be positive");
  int dev = 0, khz = 0;
  B200_CUDA(cudaGetDevice(&dev));
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  if (khz <= 0) khz = 1965000;
  double cyc = timeout_seconds * 1e3 * (double)khz;
  if (cyc > 9e18) cyc = 9e18;
  const long long cycles = (long long)cyc;
  B200_CUDA(cudaMemcpyToSymbol(g_wd_cycles, &cycles, sizeof(cycles)));
  B200_CUDA(cudaMemcpyToSymbol(g_wd_err, &host_error_word_devptr, sizeof(host_error_word_devptr)));
  return 0;
}

extern "C" int b200_p2p_watchdog_reset(void) {
  const unsigned int zero = 0;
  B200_CUDA(cudaMemcpyToSymbol(g_wd_tripped, &zero, sizeof(zero)));
  return 0;
}
