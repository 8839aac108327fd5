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
// 2(world-1) latency-bound hops).  The kernel overlaps the backward, whose persistent tcgen05 GEMM CTAs hold one SM each
// with 320 threads x up to 168 registers = 53.8 K of the SM's 64 K registers: a reduction CTA co-resides with them only
// if it fits the remaining 11.7 K registers (and uses no shared memory).  The first form of this kernel ran 48 CTAs of
// 512 threads x 64 registers = 32 K registers each, which EVICTED the GEMM from 48 of the 148 SMs for the 150 us of the
// heads' bucket -- every full-grid GEMM of the backward then needed a second round (N = 2: +150 us per step).  Now:
// 128 threads x 64 registers = 8 K registers per CTA, up to 144 CTAs (the same ~2.4 MB of switch round trips in flight).
// CUDA-graph safe: the signal slots are consumed (1 -> 0) by the waiter, nothing counts epochs.
#include <cstdlib>

#include "cpc_internal.h"
#include "ptx.cuh"

namespace cpc {

constexpr int AR_MAX_BLOCKS = 144;
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

template <int AR_THREADS>
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
  // CPC_AR_THREADS=512 restores the first form (48 CTAs x 512 threads: 3 MB in flight saturate the switch -- measured at
  // 8 GPUs, kernel alone: 66 MB in 191 us with 48 such CTAs, 215 us with 144 -- but see the register note above)
  static const int threads = [] {
    const char* e = getenv("CPC_AR_THREADS");
    const int v = e != nullptr ? atoi(e) : 0;
    return v == 128 || v == 256 ? v : 512;
  }();
  static const int env_blocks = [] {  // CPC_AR_BLOCKS=<n>: CTA cap (A/B runs)
    const char* e = getenv("CPC_AR_BLOCKS");
    return e != nullptr ? atoi(e) : 0;
  }();
  long long blocks = (per + threads * AR_UNROLL - 1) / (threads * AR_UNROLL);
  int cap = max_blocks > 0 ? max_blocks : (env_blocks > 0 ? env_blocks : 48);
  if (cap > AR_MAX_BLOCKS) cap = AR_MAX_BLOCKS;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  float* mc = static_cast<float*>(mc_ptr);
  uint32_t* const* pads = reinterpret_cast<uint32_t* const*>(signal_pads_dev);
  if (threads == 512) launch_k(allreduce_mc_kernel<512>, dim3((unsigned)blocks), dim3(512), 0, stream, mc, n4, pads, rank, world, scale);
  else if (threads == 256) launch_k(allreduce_mc_kernel<256>, dim3((unsigned)blocks), dim3(256), 0, stream, mc, n4, pads, rank, world, scale);
  else launch_k(allreduce_mc_kernel<128>, dim3((unsigned)blocks), dim3(128), 0, stream, mc, n4, pads, rank, world, scale);
  return launch_status();
}

extern "C" int cpc_allreduce_pad_bytes(int world) { return 4 * (AR_PAD_FIRST + AR_MAX_BLOCKS * world); }
