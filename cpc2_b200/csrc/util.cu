// Small streaming kernels: fp32 -> bf16 casts (parameter arenas, activations), column sums (bias
// gradients), group sums, the integer index arithmetic of the negative sampler.
#include "cpc_internal.h"
#include "ptx.cuh"

namespace cpc {

__global__ void cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long n4 = n / 4;
  for (; i < n4; i += stride) {
    const float4 v = reinterpret_cast<const float4*>(in)[i];
    uint2 u;
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    u.x = *reinterpret_cast<uint32_t*>(&a);
    u.y = *reinterpret_cast<uint32_t*>(&b);
    reinterpret_cast<uint2*>(out)[i] = u;
  }
  for (long long j = n4 * 4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += stride)
    out[j] = __float2bfloat16(in[j]);
}

__global__ void cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ in, float* __restrict__ out, long long n) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    out[i] = __bfloat162float(in[i]);
}

// Inverted dropout of a contiguous tensor, in place: x[i] = keep(i) ? x[i] / (1 - p) : 0.  keep(i) is a counter hash of
// (seed, step, site, i) -- rng = device uint64[2] {seed, step} as in the GEMM / attention kernels -- so the backward
// regenerates the mask from the same (rng, site) instead of storing it.  One hash per element pair.
template <bool BF16>
__global__ void dropout_kernel(void* __restrict__ x_, long long n, const unsigned long long* __restrict__ rng,
                               uint32_t site, float p) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  const unsigned long long seed = rng[0] + rng[1] * 0x632BE59BD9B4E019ull;
  const uint32_t seed_lo = static_cast<uint32_t>(seed);
  const uint32_t seed_hi = static_cast<uint32_t>(seed >> 32) + site * 0x85EBCA77u;
  const uint32_t thresh = static_cast<uint32_t>(fminf(p, 0.9999f) * 65536.0f);
  const float scale = 1.f / (1.f - p);
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i2 = (long long)blockIdx.x * blockDim.x + threadIdx.x; 2 * i2 < n; i2 += stride) {
    const uint32_t h = rng_pair_u16(seed_lo, seed_hi, static_cast<unsigned long long>(i2));
    const bool k0 = (h & 0xFFFFu) >= thresh, k1 = (h >> 16) >= thresh;
    const long long i = 2 * i2;
    if (BF16) {
      __nv_bfloat16* x = static_cast<__nv_bfloat16*>(x_);
      x[i] = k0 ? __float2bfloat16(__bfloat162float(x[i]) * scale) : __float2bfloat16(0.f);
      if (i + 1 < n) x[i + 1] = k1 ? __float2bfloat16(__bfloat162float(x[i + 1]) * scale) : __float2bfloat16(0.f);
    } else {
      float* x = static_cast<float*>(x_);
      x[i] = k0 ? x[i] * scale : 0.f;
      if (i + 1 < n) x[i + 1] = k1 ? x[i + 1] * scale : 0.f;
    }
  }
}

// out[g][c] (+)= sum_r in[g][r][c]   (in bf16 or fp32), one block column-slab of 32 channels x row chunk
template <bool IN_BF16>
__global__ void __launch_bounds__(256) colsum_kernel(const void* __restrict__ in_, float* __restrict__ out, long long rows,
                                                     int C, long long ld, long long gs_in, int row_chunks) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  __shared__ float red[8][33];
  const int g = blockIdx.z;
  const int c = blockIdx.x * 32 + (threadIdx.x & 31);
  const int ry = threadIdx.x >> 5;
  const long long per = (rows + row_chunks - 1) / row_chunks;
  const long long rb = (long long)blockIdx.y * per;
  const long long re = rb + per < rows ? rb + per : rows;
  float acc = 0.f;
  if (c < C) {
    for (long long r = rb + ry; r < re; r += 8) {
      if (IN_BF16) acc += __bfloat162float(static_cast<const __nv_bfloat16*>(in_)[g * gs_in + r * ld + c]);
      else acc += static_cast<const float*>(in_)[g * gs_in + r * ld + c];
    }
  }
  red[ry][threadIdx.x & 31] = acc;
  __syncthreads();
  if (ry == 0 && c < C) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x & 31];
    red_add_f32(out + (long long)g * C + c, t);
  }
}

// out[i] = sum_g in[g][i]  (reduction of per-head input gradients), fp32 or bf16 input, fp32 and/or bf16 out
template <bool IN_BF16>
__global__ void groupsum_kernel(const void* __restrict__ in_, float* __restrict__ out_f32,
                                __nv_bfloat16* __restrict__ out_bf16, long long n, int groups, int accumulate) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float acc = 0.f;
    for (int g = 0; g < groups; ++g) {
      if (IN_BF16) acc += __bfloat162float(static_cast<const __nv_bfloat16*>(in_)[g * n + i]);
      else acc += static_cast<const float*>(in_)[g * n + i];
    }
    if (out_f32 != nullptr) out_f32[i] = accumulate ? out_f32[i] + acc : acc;
    if (out_bf16 != nullptr) out_bf16[i] = __float2bfloat16(acc);
  }
}

// extIdx = ((seqIdx + t) mod S) + batchIdx * S, flat order (b, n, t)   -- criterion.py:256-266, int64, exact
__global__ void neg_ext_index_kernel(const long long* __restrict__ batch_idx, const long long* __restrict__ seq_idx,
                                     long long* __restrict__ ext, long long n, int W, int S) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const long long t = i % W;
    long long s = (seq_idx[i] + t) % S;
    if (s < 0) s += S;  // torch.remainder semantics
    ext[i] = s + batch_idx[i] * S;
  }
}

static int grid_for(long long n, int per_block) {
  long long b = (n + per_block - 1) / per_block;
  if (b > 148 * 8) b = 148 * 8;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace cpc

using namespace cpc;

extern "C" int cpc_cast_f32_to_bf16(const void* in, void* out, int64_t n, void* stream_) {
  if (in == nullptr || out == nullptr || n < 0) return CPC_ERR_ARG;
  if (n == 0) return CPC_OK;
  if ((reinterpret_cast<uintptr_t>(in) & 15) || (reinterpret_cast<uintptr_t>(out) & 7)) return CPC_ERR_ALIGN;
  launch_k(cast_f32_bf16_kernel, dim3(grid_for(n, 1024)), dim3(256), 0, static_cast<cudaStream_t>(stream_), 
      static_cast<const float*>(in), static_cast<__nv_bfloat16*>(out), n);
  return launch_status();
}

extern "C" int cpc_cast_bf16_to_f32(const void* in, void* out, int64_t n, void* stream_) {
  if (in == nullptr || out == nullptr || n < 0) return CPC_ERR_ARG;
  if (n == 0) return CPC_OK;
  launch_k(cast_bf16_f32_kernel, dim3(grid_for(n, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream_), 
      static_cast<const __nv_bfloat16*>(in), static_cast<float*>(out), n);
  return launch_status();
}

extern "C" int cpc_dropout(void* x, int is_bf16, int64_t n, const void* rng, uint32_t site, float p, void* stream_) {
  if (x == nullptr || rng == nullptr || n < 0 || !(p >= 0.f) || !(p < 1.f)) return CPC_ERR_ARG;
  if (n == 0 || p == 0.f) return CPC_OK;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (is_bf16)
    launch_k(dropout_kernel<true>, dim3(grid_for((n + 1) / 2, 256)), dim3(256), 0, stream, x, n, static_cast<const unsigned long long*>(rng), site, p);
  else
    launch_k(dropout_kernel<false>, dim3(grid_for((n + 1) / 2, 256)), dim3(256), 0, stream, x, n, static_cast<const unsigned long long*>(rng), site, p);
  return launch_status();
}

extern "C" int cpc_colsum(const void* in, int in_is_bf16, void* out, int64_t rows, int C, int64_t ld, int64_t gs_in,
                          int groups, void* stream_) {
  if (in == nullptr || out == nullptr || rows <= 0 || C <= 0 || groups <= 0) return CPC_ERR_ARG;
  int chunks = (int)((rows + 255) / 256);
  if (chunks < 1) chunks = 1;
  if (chunks > 64) chunks = 64;
  dim3 grid((C + 31) / 32, chunks, groups);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (in_is_bf16) launch_k(colsum_kernel<true>, dim3(grid), dim3(256), 0, stream, in, static_cast<float*>(out), rows, C, ld, gs_in, chunks);
  else launch_k(colsum_kernel<false>, dim3(grid), dim3(256), 0, stream, in, static_cast<float*>(out), rows, C, ld, gs_in, chunks);
  return launch_status();
}

extern "C" int cpc_groupsum(const void* in, int in_is_bf16, void* out_f32, void* out_bf16, int64_t n, int groups,
                            int accumulate, void* stream_) {
  if (in == nullptr || n <= 0 || groups <= 0 || (out_f32 == nullptr && out_bf16 == nullptr)) return CPC_ERR_ARG;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (in_is_bf16)
    launch_k(groupsum_kernel<true>, dim3(grid_for(n, 256)), dim3(256), 0, stream, in, static_cast<float*>(out_f32),
                                                                 static_cast<__nv_bfloat16*>(out_bf16), n, groups, accumulate);
  else
    launch_k(groupsum_kernel<false>, dim3(grid_for(n, 256)), dim3(256), 0, stream, in, static_cast<float*>(out_f32),
                                                                  static_cast<__nv_bfloat16*>(out_bf16), n, groups, accumulate);
  return launch_status();
}

extern "C" int cpc_neg_ext_index(const void* batch_idx, const void* seq_idx, void* ext, int B, int NEG, int W, int S,
                                 void* stream_) {
  if (batch_idx == nullptr || seq_idx == nullptr || ext == nullptr || B <= 0 || NEG <= 0 || W <= 0 || S <= 0)
    return CPC_ERR_ARG;
  const long long n = (long long)B * NEG * W;
  launch_k(neg_ext_index_kernel, dim3(grid_for(n, 256)), dim3(256), 0, static_cast<cudaStream_t>(stream_), 
      static_cast<const long long*>(batch_idx), static_cast<const long long*>(seq_idx), static_cast<long long*>(ext), n, W, S);
  return launch_status();
}
