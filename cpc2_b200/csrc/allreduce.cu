// In-switch gradient all-reduce over NVLink 5 / NVSwitch multicast memory (the collective of the data-parallel CPC
// step: DistributedDataParallel's gradient all-reduce, cpc/train.py:523-527).  One launch per gradient bucket, in
// place on a SYMMETRIC buffer (same offset on every rank, mapped through one multicast address):
//
//   barrier (every rank's gradients are complete)            -- signal pads, one slot per (block, peer)
//   rank r owns 1/world of the bucket:  v = multimem.ld_reduce.add [mc + i]   (the switch sums the 8 copies)
//                                        multimem.st [mc + i] = v * scale       (the switch writes all 8 copies)
//   barrier (every slice has landed everywhere)
//
// Per GPU the NVLink traffic is n*4*(world-1)/world bytes out and in (a ring all-reduce moves twice that, in
// 2(world-1) latency-bound hops), and the kernel needs no shared memory and few registers, so its CTAs co-reside with
// the persistent GEMM CTAs of the backward it overlaps instead of taking whole SMs away from them.
// CUDA-graph safe: the signal slots are consumed (1 -> 0) by the waiter, nothing counts epochs.
#include "cpc_internal.h"
#include "ptx.cuh"

namespace cpc {

constexpr int AR_THREADS = 512;
constexpr int AR_MAX_BLOCKS = 144;
constexpr int AR_DEFAULT_BLOCKS = 48;
constexpr int AR_UNROLL = 8;  // independent switch round trips in flight per thread
constexpr int AR_PAD_FIRST = 512;   // first u32 slot of the signal pad we use (the slots below are left to the host library)

__device__ __forceinline__ void signal_put(uint32_t* slot) {  // 0 -> 1 on the peer's pad
  uint32_t old;
  do {
    asm volatile("atom.global.release.sys.cas.b32 %0, [%1], 0, 1;" : "=r"(old) : "l"(slot) : "memory");
  } while (old != 0u);
}

__device__ __forceinline__ void signal_wait(uint32_t* slot) {  // 1 -> 0 on my pad
  uint32_t old;
  do {
    asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], 1, 0;" : "=r"(old) : "l"(slot) : "memory");
  } while (old != 1u);
}

// All CTAs with the same blockIdx on all ranks meet.  Thread t < world signals peer t and waits for peer t.
__device__ __forceinline__ void block_barrier(uint32_t* const* pads, int rank, int world) {
  __syncthreads();
  if ((int)threadIdx.x < world) {
    const int peer = threadIdx.x;
    signal_put(pads[peer] + AR_PAD_FIRST + blockIdx.x * world + rank);
    signal_wait(pads[rank] + AR_PAD_FIRST + blockIdx.x * world + peer);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(AR_THREADS) allreduce_mc_kernel(float* __restrict__ mc, long long n4,
                                                                  uint32_t* const* __restrict__ pads, int rank, int world,
                                                                  float scale) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  block_barrier(pads, rank, world);
  const long long per = (n4 + world - 1) / world;  // 16-byte units per rank
  const long long begin = per * rank;
  const long long end = begin + per < n4 ? begin + per : n4;
  const long long stride = (long long)gridDim.x * AR_THREADS;
  long long i = begin + (long long)blockIdx.x * AR_THREADS + threadIdx.x;
  for (; i + (AR_UNROLL - 1) * stride < end; i += AR_UNROLL * stride) {
    float4 v[AR_UNROLL];
#pragma unroll
    for (int u = 0; u < AR_UNROLL; ++u) {
      const float* a = mc + (i + u * stride) * 4;
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                   : "=f"(v[u].x), "=f"(v[u].y), "=f"(v[u].z), "=f"(v[u].w) : "l"(a) : "memory");
    }
#pragma unroll
    for (int u = 0; u < AR_UNROLL; ++u) {
      float* a = mc + (i + u * stride) * 4;
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                   :: "l"(a), "f"(v[u].x * scale), "f"(v[u].y * scale), "f"(v[u].z * scale), "f"(v[u].w * scale) : "memory");
    }
  }
  for (; i < end; i += stride) {
    float4 v;
    float* a = mc + i * 4;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(a) : "memory");
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 :: "l"(a), "f"(v.x * scale), "f"(v.y * scale), "f"(v.z * scale), "f"(v.w * scale) : "memory");
  }
  // multimem.st is weak: order it before the release of the barrier's signal
  __threadfence_system();
  block_barrier(pads, rank, world);
}

}  // namespace cpc

using namespace cpc;

extern "C" int cpc_allreduce_mc(void* mc_ptr, int64_t n, void* const* signal_pads_dev, int rank, int world, float scale,
                                int max_blocks, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (mc_ptr == nullptr || signal_pads_dev == nullptr || n <= 0 || (n & 3) != 0 || world < 2 || world > 16 || rank < 0 ||
      rank >= world || (reinterpret_cast<uintptr_t>(mc_ptr) & 15) != 0)
    return CPC_ERR_ARG;
  const long long n4 = n / 4;
  const long long per = (n4 + world - 1) / world;
  long long blocks = (per + AR_THREADS * AR_UNROLL - 1) / (AR_THREADS * AR_UNROLL);
  // 48 CTAs x 512 threads x 8 x 16 B = 3 MB in flight saturate the switch (measured at 8 GPUs: 66 MB in 191 us with 48
  // CTAs, 215 us with 144) and leave 100 SMs untouched for the backward kernels the reduction overlaps
  const int cap = max_blocks > 0 ? (max_blocks < AR_MAX_BLOCKS ? max_blocks : AR_MAX_BLOCKS) : AR_DEFAULT_BLOCKS;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  launch_k(allreduce_mc_kernel, dim3((unsigned)blocks), dim3(AR_THREADS), 0, stream, static_cast<float*>(mc_ptr), n4,
                                                                 reinterpret_cast<uint32_t* const*>(signal_pads_dev), rank,
                                                                 world, scale);
  return launch_status();
}

extern "C" int cpc_allreduce_pad_bytes(int world) { return 4 * (AR_PAD_FIRST + AR_MAX_BLOCKS * world); }
