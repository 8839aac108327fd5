// Row normalisation kernels: ChannelNorm(+ReLU) of the encoder (cpc/model.py:52-60,103-107; unbiased
// variance over C) and nn.LayerNorm of the transformer blocks (cpc/transformers.py:124,130; biased
// variance), forward and backward.  HBM-bound: one warp per row, each lane owns V = C/32 contiguous
// channels (16-byte vector loads for C = 256), two shuffle reductions per row, statistics in fp32.
// Rows are addressed through a (sequence, time) -> row map so the kernels read and write the zero-halo
// channels-last buffers the implicit-GEMM convolutions consume.
#include "cpc_internal.h"
#include "ptx.cuh"

namespace cpc {

struct RowMap {
  long long seq_len;     // rows per sequence
  long long seq_stride;  // rows between consecutive sequences in the buffer
  long long row_off;     // first valid row inside a sequence (halo)
  // Row counts are below 2^31 (checked by the launchers): 32-bit division.  The 64-bit one is a ~100-instruction
  // routine, and a backward thread evaluates a dozen maps before it can issue its first load.
  __device__ __forceinline__ long long operator()(long long r) const {
    if (seq_stride == seq_len && row_off == 0) return r;  // dense buffer
    const unsigned n = (unsigned)r / (unsigned)seq_len;
    return (long long)n * seq_stride + row_off + ((unsigned)r - n * (unsigned)seq_len);
  }
};

template <int V>
__device__ __forceinline__ void load_bf16(const __nv_bfloat16* p, float (&x)[V]) {
  if constexpr (V >= 8) {
#pragma unroll
    for (int i = 0; i < V; i += 8) {
      const uint4 u = *reinterpret_cast<const uint4*>(p + i);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        x[i + 2 * j] = __low2float(h[j]);
        x[i + 2 * j + 1] = __high2float(h[j]);
      }
    }
  } else if constexpr (V == 4) {
    const uint2 u = *reinterpret_cast<const uint2*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
    x[0] = __low2float(h[0]); x[1] = __high2float(h[0]); x[2] = __low2float(h[1]); x[3] = __high2float(h[1]);
  } else if constexpr (V == 2) {
    const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(p);
    x[0] = __low2float(h); x[1] = __high2float(h);
  } else {
    x[0] = __bfloat162float(p[0]);
  }
}
// Raw 32-bit words of a row chunk: lets a kernel issue every load of several rows before the first conversion
// (load + convert in one helper made each row's loads wait for the previous row's data).
template <int W>
__device__ __forceinline__ void load_words(const void* p, uint32_t (&w)[W]) {
  if constexpr (W >= 4) {
#pragma unroll
    for (int i = 0; i < W; i += 4) {
      const uint4 u = *reinterpret_cast<const uint4*>(static_cast<const uint32_t*>(p) + i);
      w[i] = u.x; w[i + 1] = u.y; w[i + 2] = u.z; w[i + 3] = u.w;
    }
  } else if constexpr (W == 2) {
    const uint2 u = *static_cast<const uint2*>(p);
    w[0] = u.x; w[1] = u.y;
  } else {
    w[0] = *static_cast<const uint32_t*>(p);
  }
}
template <int V>
__device__ __forceinline__ void bf16_words_to_f32(const uint32_t (&w)[V / 2], float (&x)[V]) {
#pragma unroll
  for (int j = 0; j < V / 2; ++j) {
    x[2 * j] = __uint_as_float(w[j] << 16);
    x[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
  }
}
template <int V>
__device__ __forceinline__ void store_bf16(__nv_bfloat16* p, const float (&x)[V]) {
  if constexpr (V >= 8) {
#pragma unroll
    for (int i = 0; i < V; i += 8) {
      uint4 u;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
      for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(x[i + 2 * j], x[i + 2 * j + 1]);
      *reinterpret_cast<uint4*>(p + i) = u;
    }
  } else if constexpr (V == 4) {
    uint2 u;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
    h[0] = __floats2bfloat162_rn(x[0], x[1]);
    h[1] = __floats2bfloat162_rn(x[2], x[3]);
    *reinterpret_cast<uint2*>(p) = u;
  } else if constexpr (V == 2) {
    *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(x[0], x[1]);
  } else {
    p[0] = __float2bfloat16(x[0]);
  }
}
template <int V>
__device__ __forceinline__ void load_f32(const float* p, float (&x)[V]) {
  if constexpr (V >= 4) {
#pragma unroll
    for (int i = 0; i < V; i += 4) {
      const float4 u = *reinterpret_cast<const float4*>(p + i);
      x[i] = u.x; x[i + 1] = u.y; x[i + 2] = u.z; x[i + 3] = u.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < V; ++i) x[i] = p[i];
  }
}
template <int V>
__device__ __forceinline__ void store_f32(float* p, const float (&x)[V]) {
  if constexpr (V >= 4) {
#pragma unroll
    for (int i = 0; i < V; i += 4) *reinterpret_cast<float4*>(p + i) = make_float4(x[i], x[i + 1], x[i + 2], x[i + 3]);
  } else {
#pragma unroll
    for (int i = 0; i < V; ++i) p[i] = x[i];
  }
}

// Outputs addressed through a RowMap with halo rows (row_off > 0 or seq_stride > row_off + seq_len) are the
// zero-halo buffers of the implicit-GEMM convolutions: the warp that writes the first / last row of a sequence
// also clears the halo rows before / after it, so no separate memset launches exist.
template <int V, typename T>
__device__ __forceinline__ void zero_halo(T* base, const RowMap& m, long long r, int lane) {
  constexpr int C = 32 * V;
  if (m.seq_stride == m.seq_len) return;  // no halo rows
  const long long n = (unsigned)r / (unsigned)m.seq_len, t = r - n * m.seq_len;
  if (t != 0 && t != m.seq_len - 1) return;
  float z[V];
#pragma unroll
  for (int i = 0; i < V; ++i) z[i] = 0.f;
  if (t == 0)
    for (long long h = 0; h < m.row_off; ++h) {
      if constexpr (sizeof(T) == 2) store_bf16<V>(base + (n * m.seq_stride + h) * C + lane * V, z);
      else store_f32<V>(base + (n * m.seq_stride + h) * C + lane * V, z);
    }
  if (t == m.seq_len - 1)
    for (long long h = m.row_off + m.seq_len; h < m.seq_stride; ++h) {
      if constexpr (sizeof(T) == 2) store_bf16<V>(base + (n * m.seq_stride + h) * C + lane * V, z);
      else store_f32<V>(base + (n * m.seq_stride + h) * C + lane * V, z);
    }
}

struct RowNormFwdArgs {
  const __nv_bfloat16* in;
  RowMap in_map;
  const float* gamma;  // [groups][C] or null
  const float* beta;
  __nv_bfloat16* out_bf16;  // optional
  float* out_f32;           // optional
  RowMap out_map;
  float* mean;  // [rows]
  float* rstd;
  long long rows, rows_per_group;
  float eps;
  int unbiased, relu;
};

template <int V>
__global__ void __launch_bounds__(256) rownorm_fwd_kernel(const RowNormFwdArgs a) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  constexpr int C = 32 * V;
  const int lane = threadIdx.x & 31;
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long r = warp; r < a.rows; r += nwarps) {
    float x[V];
    load_bf16<V>(a.in + a.in_map(r) * C + lane * V, x);
    float ga[V], be[V];  // requested with the row, not behind the mean / rstd stores below
    if (a.gamma != nullptr) {
      const long long g = (unsigned)r / (unsigned)a.rows_per_group;
      load_f32<V>(a.gamma + g * C + lane * V, ga);
      load_f32<V>(a.beta + g * C + lane * V, be);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) s += x[i];
    const float mean = warp_sum(s) * (1.f / C);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      x[i] -= mean;
      q += x[i] * x[i];
    }
    const float var = warp_sum(q) / (a.unbiased ? (C - 1) : C);
    const float rstd = rsqrtf(var + a.eps);
    if (lane == 0 && a.mean != nullptr) {
      a.mean[r] = mean;
      a.rstd[r] = rstd;
    }
    if (a.gamma != nullptr) {
#pragma unroll
      for (int i = 0; i < V; ++i) x[i] = x[i] * rstd * ga[i] + be[i];
    } else {
#pragma unroll
      for (int i = 0; i < V; ++i) x[i] *= rstd;
    }
    if (a.relu) {
#pragma unroll
      for (int i = 0; i < V; ++i) x[i] = fmaxf(x[i], 0.f);
    }
    const long long o = a.out_map(r) * C + lane * V;
    if (a.out_bf16 != nullptr) {
      store_bf16<V>(a.out_bf16 + o, x);
      zero_halo<V>(a.out_bf16, a.out_map, r, lane);
    }
    if (a.out_f32 != nullptr) {
      store_f32<V>(a.out_f32 + o, x);
      zero_halo<V>(a.out_f32, a.out_map, r, lane);
    }
  }
}

struct RowNormBwdArgs {
  const void* dout;  // bf16 or f32 (dout_f32 flag), rows addressed by dout_map
  RowMap dout_map;
  const float* dout2;  // optional second fp32 gradient added to dout (same map as dout2_map)
  RowMap dout2_map;
  const __nv_bfloat16* in;  // forward input (pre-norm)
  RowMap in_map;
  const float* mean;
  const float* rstd;
  const float* gamma;
  const float* beta;
  __nv_bfloat16* dx_bf16;  // optional
  float* dx_f32;           // optional
  RowMap dx_map;
  float* dgamma;  // [groups][C] accumulated atomically (optional)
  float* dbeta;
  float* dxsum;  // [groups][C] column sum of dx (bias gradient of the producing conv), optional
  long long rows_per_group;
  int unbiased, relu, dout_f32, dx_accumulate;
};

// Two CTAs per SM for C <= 256 (128 registers): at 161 registers one CTA per SM was resident and a 256- or 348-CTA
// launch ran as two or three full latency chains (load -> reduce -> store -> block reduction -> atomics) back to back.
template <int V>
__global__ void __launch_bounds__(256, (V <= 8 ? 2 : 1)) rownorm_bwd_kernel(const RowNormBwdArgs a) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  constexpr int C = 32 * V;
  __shared__ float red[3][8][C > 512 ? 1 : C];
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5;
  const int wpb = blockDim.x >> 5;
  const long long g = blockIdx.y;
  const long long r0 = g * a.rows_per_group;
  float ag[V], ab[V], as[V];
#pragma unroll
  for (int i = 0; i < V; ++i) ag[i] = ab[i] = as[i] = 0.f;
  float ga[V], be[V];
  if (a.gamma != nullptr) {
    load_f32<V>(a.gamma + g * C + lane * V, ga);
    load_f32<V>(a.beta + g * C + lane * V, be);
  } else {
#pragma unroll
    for (int i = 0; i < V; ++i) { ga[i] = 1.f; be[i] = 0.f; }
  }
  const float inv_n = a.unbiased ? 1.f / (C - 1) : 1.f / C;
  // R rows per warp and iteration with every load issued before the first use: the kernel is bound by the
  // load -> reduce -> store latency chain of a row, not by bandwidth (a 1024-row layer took as long as an 8192-row one)
  constexpr int R = V <= 8 ? 4 : 2;
  for (long long rr0 = ((long long)blockIdx.x * wpb + wib) * R; rr0 < a.rows_per_group; rr0 += (long long)gridDim.x * wpb * R) {
    float x[R][V], d[R][V], mean[R], rstd[R];
    bool ok[R];
    long long rq[R];
#pragma unroll
    for (int q = 0; q < R; ++q) {
      ok[q] = rr0 + q < a.rows_per_group;
      rq[q] = r0 + rr0 + (ok[q] ? q : 0);
    }
    if constexpr (V >= 2) {
      // phase 1: every load of the R rows in flight (raw words), phase 2: conversions
      uint32_t xw[R][V / 2];
#pragma unroll
      for (int q = 0; q < R; ++q) load_words<V / 2>(a.in + a.in_map(rq[q]) * C + lane * V, xw[q]);
      if (a.dout_f32) {
        uint32_t dw[R][V];
#pragma unroll
        for (int q = 0; q < R; ++q)
          load_words<V>(static_cast<const float*>(a.dout) + a.dout_map(rq[q]) * C + lane * V, dw[q]);
#pragma unroll
        for (int q = 0; q < R; ++q) {
          mean[q] = a.mean[rq[q]];
          rstd[q] = a.rstd[rq[q]];
        }
#pragma unroll
        for (int q = 0; q < R; ++q)
#pragma unroll
          for (int i = 0; i < V; ++i) d[q][i] = __uint_as_float(dw[q][i]);
      } else {
        uint32_t dw[R][V / 2];
#pragma unroll
        for (int q = 0; q < R; ++q)
          load_words<V / 2>(static_cast<const __nv_bfloat16*>(a.dout) + a.dout_map(rq[q]) * C + lane * V, dw[q]);
#pragma unroll
        for (int q = 0; q < R; ++q) {
          mean[q] = a.mean[rq[q]];
          rstd[q] = a.rstd[rq[q]];
        }
#pragma unroll
        for (int q = 0; q < R; ++q) bf16_words_to_f32<V>(dw[q], d[q]);
      }
#pragma unroll
      for (int q = 0; q < R; ++q) bf16_words_to_f32<V>(xw[q], x[q]);
    } else {
#pragma unroll
      for (int q = 0; q < R; ++q) {
        const long long r = rq[q];
        load_bf16<V>(a.in + a.in_map(r) * C + lane * V, x[q]);
        if (a.dout_f32) load_f32<V>(static_cast<const float*>(a.dout) + a.dout_map(r) * C + lane * V, d[q]);
        else load_bf16<V>(static_cast<const __nv_bfloat16*>(a.dout) + a.dout_map(r) * C + lane * V, d[q]);
        mean[q] = a.mean[r];
        rstd[q] = a.rstd[r];
      }
    }
    if (a.dout2 != nullptr) {
#pragma unroll
      for (int q = 0; q < R; ++q) {
        const long long r = r0 + rr0 + (ok[q] ? q : 0);
        float d2[V];
        load_f32<V>(a.dout2 + a.dout2_map(r) * C + lane * V, d2);
#pragma unroll
        for (int i = 0; i < V; ++i) d[q][i] += d2[i];
      }
    }
    float s1[R], s2[R];
#pragma unroll
    for (int q = 0; q < R; ++q) {
      s1[q] = s2[q] = 0.f;
#pragma unroll
      for (int i = 0; i < V; ++i) {
        x[q][i] = (x[q][i] - mean[q]) * rstd[q];  // xhat
        if (!ok[q] || (a.relu && (x[q][i] * ga[i] + be[i]) <= 0.f)) d[q][i] = 0.f;
        ag[i] += d[q][i] * x[q][i];
        ab[i] += d[q][i];
        d[q][i] *= ga[i];  // dxhat
        s1[q] += d[q][i];
        s2[q] += d[q][i] * x[q][i];
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int q = 0; q < R; ++q) {
        s1[q] += __shfl_xor_sync(0xffffffffu, s1[q], o);
        s2[q] += __shfl_xor_sync(0xffffffffu, s2[q], o);
      }
    }
#pragma unroll
    for (int q = 0; q < R; ++q) {
      if (!ok[q]) continue;
      const long long r = r0 + rr0 + q;
      const float m1 = s1[q] * (1.f / C), m2 = s2[q] * inv_n;
#pragma unroll
      for (int i = 0; i < V; ++i) {
        d[q][i] = rstd[q] * (d[q][i] - m1 - x[q][i] * m2);
        as[i] += d[q][i];
      }
      const long long o = a.dx_map(r) * C + lane * V;
      if (a.dx_bf16 != nullptr) {
        store_bf16<V>(a.dx_bf16 + o, d[q]);
        zero_halo<V>(a.dx_bf16, a.dx_map, r, lane);
      }
      if (a.dx_f32 != nullptr) {
        if (!a.dx_accumulate) zero_halo<V>(a.dx_f32, a.dx_map, r, lane);
        if (a.dx_accumulate) {
          float old[V];
          load_f32<V>(a.dx_f32 + o, old);
#pragma unroll
          for (int i = 0; i < V; ++i) old[i] += d[q][i];
          store_f32<V>(a.dx_f32 + o, old);
        } else {
          store_f32<V>(a.dx_f32 + o, d[q]);
        }
      }
    }
  }
  // block reduction of the per-channel sums, then one atomic per channel per block
  if constexpr (C <= 512) {
#pragma unroll
    for (int i = 0; i < V; ++i) {
      red[0][wib][lane * V + i] = ag[i];
      red[1][wib][lane * V + i] = ab[i];
      red[2][wib][lane * V + i] = as[i];
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      float t0 = 0.f, t1 = 0.f, t2 = 0.f;
      for (int w = 0; w < wpb; ++w) {
        t0 += red[0][w][c];
        t1 += red[1][w][c];
        t2 += red[2][w][c];
      }
      if (a.dgamma != nullptr) {
        red_add_f32(a.dgamma + g * C + c, t0);
        red_add_f32(a.dbeta + g * C + c, t1);
      }
      if (a.dxsum != nullptr) red_add_f32(a.dxsum + g * C + c, t2);
    }
  }
}

static RowMap make_map(long long seq_len, long long seq_stride, long long row_off, long long rows) {
  RowMap m;
  m.seq_len = seq_len > 0 ? seq_len : (rows > 0 ? rows : 1);
  m.seq_stride = seq_len > 0 ? seq_stride : m.seq_len;
  m.row_off = row_off;
  return m;
}

}  // namespace cpc

using namespace cpc;

static RowMap to_map(const CpcRowMap& m, long long rows) { return make_map(m.seq_len, m.seq_stride, m.row_off, rows); }

extern "C" int cpc_rownorm_fwd(const CpcRowNormFwd* d, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (d == nullptr || d->rows <= 0 || d->rows > 0x7fffffffLL || d->C % 32 != 0 || d->in == nullptr) return CPC_ERR_ARG;
  RowNormFwdArgs a;
  a.in = static_cast<const __nv_bfloat16*>(d->in);
  a.in_map = to_map(d->in_map, d->rows);
  a.gamma = static_cast<const float*>(d->gamma);
  a.beta = static_cast<const float*>(d->beta);
  a.out_bf16 = static_cast<__nv_bfloat16*>(d->out_bf16);
  a.out_f32 = static_cast<float*>(d->out_f32);
  a.out_map = to_map(d->out_map, d->rows);
  a.mean = static_cast<float*>(d->mean);
  a.rstd = static_cast<float*>(d->rstd);
  a.rows = d->rows;
  a.rows_per_group = d->rows_per_group > 0 ? d->rows_per_group : d->rows;
  a.eps = d->eps; a.unbiased = d->unbiased; a.relu = d->relu;
  const int wpb = 8;
  long long blocks = (d->rows + wpb - 1) / wpb;
  if (blocks > 148 * 16) blocks = 148 * 16;
  switch (d->C / 32) {
    case 1: launch_k(rownorm_fwd_kernel<1>, dim3((unsigned)blocks), dim3(256), 0, stream, a); break;
    case 2: launch_k(rownorm_fwd_kernel<2>, dim3((unsigned)blocks), dim3(256), 0, stream, a); break;
    case 4: launch_k(rownorm_fwd_kernel<4>, dim3((unsigned)blocks), dim3(256), 0, stream, a); break;
    case 8: launch_k(rownorm_fwd_kernel<8>, dim3((unsigned)blocks), dim3(256), 0, stream, a); break;
    case 16: launch_k(rownorm_fwd_kernel<16>, dim3((unsigned)blocks), dim3(256), 0, stream, a); break;
    default: return CPC_ERR_UNSUPPORTED;
  }
  return launch_status();
}

extern "C" int cpc_rownorm_bwd(const CpcRowNormBwd* d, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (d == nullptr || d->rows <= 0 || d->rows > 0x7fffffffLL || d->C % 32 != 0 || d->in == nullptr || d->dout == nullptr)
    return CPC_ERR_ARG;
  const long long rpg = d->rows_per_group > 0 ? d->rows_per_group : d->rows;
  if (d->rows % rpg != 0) return CPC_ERR_ARG;
  RowNormBwdArgs a;
  a.dout = d->dout; a.dout_map = to_map(d->dout_map, d->rows); a.dout_f32 = d->dout_f32;
  a.dout2 = static_cast<const float*>(d->dout2); a.dout2_map = to_map(d->dout2_map, d->rows);
  a.in = static_cast<const __nv_bfloat16*>(d->in); a.in_map = to_map(d->in_map, d->rows);
  a.mean = static_cast<const float*>(d->mean); a.rstd = static_cast<const float*>(d->rstd);
  a.gamma = static_cast<const float*>(d->gamma); a.beta = static_cast<const float*>(d->beta);
  a.dx_bf16 = static_cast<__nv_bfloat16*>(d->dx_bf16); a.dx_f32 = static_cast<float*>(d->dx_f32);
  a.dx_map = to_map(d->dx_map, d->rows); a.dx_accumulate = d->dx_accumulate;
  a.dgamma = static_cast<float*>(d->dgamma); a.dbeta = static_cast<float*>(d->dbeta);
  a.dxsum = static_cast<float*>(d->dxsum);
  a.rows_per_group = rpg; a.unbiased = d->unbiased; a.relu = d->relu;
  const int groups = (int)(d->rows / rpg);
  const int wpb = 8;
  long long bx = (rpg + wpb * 4 - 1) / (wpb * 4);  // one 4-row batch per warp where the row count allows
  // at most one resident wave (2 CTAs per SM, 1 at C = 512): further rows are further loop iterations of the same
  // CTAs, which cost a load -> store chain, not another block reduction and launch ramp
  long long cap = (148 * (d->C <= 256 ? 2 : 1)) / groups;
  if (cap < 1) cap = 1;
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  dim3 grid((unsigned)bx, groups);
  switch (d->C / 32) {
    case 1: launch_k(rownorm_bwd_kernel<1>, dim3(grid), dim3(256), 0, stream, a); break;
    case 2: launch_k(rownorm_bwd_kernel<2>, dim3(grid), dim3(256), 0, stream, a); break;
    case 4: launch_k(rownorm_bwd_kernel<4>, dim3(grid), dim3(256), 0, stream, a); break;
    case 8: launch_k(rownorm_bwd_kernel<8>, dim3(grid), dim3(256), 0, stream, a); break;
    case 16: launch_k(rownorm_bwd_kernel<16>, dim3(grid), dim3(256), 0, stream, a); break;
    default: return CPC_ERR_UNSUPPORTED;
  }
  return launch_status();
}
