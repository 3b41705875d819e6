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
// No host involvement after set-up; a dead peer turns into a trap after a bounded spin instead of a hang.
#include <cuda.h>
#include "common.h"
#include "../include/b200seq.h"

#include <cstring>

namespace b200 {

constexpr int kMaxRanks = 8;
constexpr int kP2PThreads = 512;

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
  float* sq_out;                   // local, optional: += sum of squares of the reduced values this rank produced
  float* mc;                       // multicast (NVLS) address of the arena, or null
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
  asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
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

__device__ __forceinline__ void wait_flag(const unsigned int* p, unsigned int epoch) {
  long long t0 = clock64();
  // epochs only grow; (int) difference handles wrap-around
  while ((int)(ld_acquire_sys(p) - epoch) < 0) {
    if (clock64() - t0 > 20000000000ll) {   // ~10 s
      printf("b200seq p2p: peer flag wait timed out (block %d, want epoch %u, have %u)\n", blockIdx.x, epoch, *p);
      __trap();
    }
    __nanosleep(64);
  }
}

__device__ __forceinline__ int flag_index(int phase, int bucket, int src, int max_buckets) {
  return (phase * max_buckets + bucket) * kMaxRanks + src;
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
  if (t < a.world && t != a.rank) wait_flag(a.flags[a.rank] + flag_index(0, a.bucket, t, a.max_buckets), epoch);
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
    if (t < a.world && t != a.rank) wait_flag(a.flags[a.rank] + flag_index(1, a.bucket, t, a.max_buckets), epoch);
    __syncthreads();
    if (i < n4) *reinterpret_cast<float4*>(a.data[a.rank] + e) = acc;
    if (a.sq_out != nullptr && (a.bucket % a.world) == a.rank) {       // one rank accounts for a small bucket's norm
      const float s = block_sum(i < n4 ? acc.x * acc.x + acc.y * acc.y + acc.z * acc.z + acc.w * acc.w : 0.f);
      if (t == 0) atomicAdd(a.sq_out, s);
    }
    return;
  }

  // ---- two-shot: reduce my slice over all ranks, push the result to every rank -----------------------------
  float sq = 0.f;
  if (a.mc != nullptr) {
    // NVLS: one multimem.ld_reduce pulls the 8-way sum out of the switch, one multimem.st replicates it to all
    // ranks: S/W in + S/W out per GPU instead of (W-1)/W * S each way.
    const long long per = ((n4 + a.world - 1) / a.world);
    const long long lo = per * a.rank, hi = (lo + per < n4) ? lo + per : n4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    constexpr int U = 4;
    for (long long i0 = lo + (long long)blockIdx.x * blockDim.x + t; i0 < hi; i0 += U * stride) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + u * stride;
        if (i < hi) v[u] = multimem_ld_reduce_f4(a.mc + a.start + (i << 2));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + u * stride;
        if (i < hi) {
          multimem_st_f4(a.mc + a.start + (i << 2), v[u]);
          sq += v[u].x * v[u].x + v[u].y * v[u].y + v[u].z * v[u].z + v[u].w * v[u].w;
        }
      }
    }
  } else {
    const long long per = ((n4 + a.world - 1) / a.world);
    const long long lo = per * a.rank, hi = (lo + per < n4) ? lo + per : n4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    constexpr int U = 4;                      // independent 16-byte requests per peer in flight per thread
    for (long long i0 = lo + (long long)blockIdx.x * blockDim.x + t; i0 < hi; i0 += U * stride) {
      float4 v[U][kMaxRanks];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + u * stride;
        if (i < hi) {
          const long long e = a.start + (i << 2);
#pragma unroll
          for (int p = 0; p < kMaxRanks; ++p)
            if (p < a.world) v[u][p] = (p == a.rank) ? *reinterpret_cast<const float4*>(a.data[p] + e) : ld_peer_f4(a.data[p] + e);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + u * stride;
        if (i < hi) {
          const long long e = a.start + (i << 2);
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int p = 0; p < kMaxRanks; ++p)      // fixed rank order -> identical bits on every rank
            if (p < a.world) { acc.x += v[u][p].x; acc.y += v[u][p].y; acc.z += v[u][p].z; acc.w += v[u][p].w; }
#pragma unroll
          for (int p = 0; p < kMaxRanks; ++p)
            if (p < a.world) st_peer_f4(a.data[p] + e, acc);
          sq += acc.x * acc.x + acc.y * acc.y + acc.z * acc.z + acc.w * acc.w;
        }
      }
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
  if (t < a.world && t != a.rank) {
    __threadfence_system();
    st_release_sys(a.flags[t] + flag_index(1, a.bucket, a.rank, a.max_buckets), epoch);
    wait_flag(a.flags[a.rank] + flag_index(1, a.bucket, t, a.max_buckets), epoch);
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

extern "C" int b200_p2p_allreduce(void* const* data_ptrs, void* const* flag_ptrs, void* counter, int rank, int world,
                                  long long start, long long n, int bucket, int max_buckets, void* epoch_dev,
                                  int one_shot, int blocks, float* sq_out, void* multicast_base, void* stream) {
  if (!data_ptrs || !flag_ptrs || !counter || !epoch_dev) return fail(B200_ERR_ARG, "p2p_allreduce: null pointer");
  if (world < 2 || world > kMaxRanks || rank < 0 || rank >= world) return fail(B200_ERR_ARG, "p2p_allreduce: bad rank/world %d/%d", rank, world);
  if ((start & 3) || (n & 3)) return fail(B200_ERR_ARG, "p2p_allreduce: range must be a multiple of 4 elements");
  if (bucket < 0 || bucket >= max_buckets) return fail(B200_ERR_ARG, "p2p_allreduce: bucket index out of range");
  P2PArgs a;
  for (int i = 0; i < kMaxRanks; ++i) {
    a.data[i] = i < world ? static_cast<float*>(data_ptrs[i]) : nullptr;
    a.flags[i] = i < world ? static_cast<unsigned int*>(flag_ptrs[i]) : nullptr;
  }
  a.counter = static_cast<unsigned int*>(counter);
  a.rank = rank; a.world = world; a.start = start; a.n = n; a.bucket = bucket; a.max_buckets = max_buckets;
  a.epoch_dev = static_cast<unsigned int*>(epoch_dev); a.one_shot = one_shot;
  a.sq_out = sq_out;
  a.mc = one_shot ? nullptr : static_cast<float*>(multicast_base);
  if (blocks <= 0) blocks = 32;
  // blocks >= 1000 encodes a narrower CTA: value = threads * 1000 + blocks (threads a multiple of 32, <= 512).
  // Many narrow CTAs spread the copy warps over every SM instead of loading a few SMs heavily, which matters
  // while a persistent GEMM (one CTA per SM, static tile schedule) is running next to the reduction.
  int threads = kP2PThreads;
  if (blocks >= 1000) {
    threads = blocks / 1000;
    blocks %= 1000;
    if (threads < 32 || threads > kP2PThreads || (threads & 31) || blocks < 1) return fail(B200_ERR_ARG, "p2p_allreduce: bad threads/blocks encoding");
  }
  if (one_shot) {
    threads = kP2PThreads;
    const long long need = ((n >> 2) + kP2PThreads - 1) / kP2PThreads;
    if (need > 4 * num_sms()) return fail(B200_ERR_ARG, "p2p_allreduce: bucket too large for one-shot mode");
    blocks = (int)(need > 0 ? need : 1);
  }
  p2p_allreduce_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(a);
  B200_LAUNCHED("p2p_allreduce_kernel");
  return 0;
}
