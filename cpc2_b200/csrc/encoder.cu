// Encoder-specific kernels (cpc/model.py:63-108).
//   * conv0 (C_in = 1, k = 10, stride 5, pad 3) + ChannelNorm + ReLU fused, forward and backward.  With one
//     input channel the layer is HBM-bound (10 MACs per output), so it runs on CUDA cores straight from the
//     fp32 waveform: one warp per output frame, lane owns C/32 channels, the whole row normalised in
//     registers and written once as bf16 channels-last (with the zero halo conv1's implicit GEMM needs).
//     The backward recomputes the conv from the waveform, so layer 0 keeps NO pre-norm activation.
//   * packing of conv1..4 weights into the two bf16 GEMM operand layouts (forward / wgrad and dgrad).
#include "cpc_internal.h"
#include "ptx.cuh"

namespace cpc {

constexpr int K0 = 10, S0 = 5, P0 = 3;

template <int V>
__device__ __forceinline__ void st_bf16v(__nv_bfloat16* p, const float (&x)[V]) {
  if constexpr (V >= 8) {
#pragma unroll
    for (int i = 0; i < V; i += 8) {
      uint4 u;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
      for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(x[i + 2 * j], x[i + 2 * j + 1]);
      *reinterpret_cast<uint4*>(p + i) = u;
    }
  } else {
#pragma unroll
    for (int i = 0; i < V; ++i) p[i] = __float2bfloat16(x[i]);
  }
}
template <int V>
__device__ __forceinline__ void ld_bf16v(const __nv_bfloat16* p, float (&x)[V]) {
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
  } else {
#pragma unroll
    for (int i = 0; i < V; ++i) x[i] = __bfloat162float(p[i]);
  }
}

struct Conv0Args {
  const float* x;  // [N, W] waveform
  const float* w;  // [C, 1, 10]
  const float* b;  // [C]
  const float* gamma;
  const float* beta;
  __nv_bfloat16* out;  // [N, L0 + 2*halo, C]
  float* mean;         // [N*L0]
  float* rstd;
  int N, W, L0, halo;
  float eps;
  // backward only
  const __nv_bfloat16* dout;  // [N, L0 + 2*halo, C] gradient w.r.t. the (padded) layer output
  float* dw;                  // [C, 10]
  float* db;
  float* dgamma;
  float* dbeta;
  int pass;  // channel sub-range processed by this launch (V > VA)
};

template <int V>
__global__ void __launch_bounds__(256) conv0_cn_relu_fwd_kernel(const Conv0Args a) {
  constexpr int C = 32 * V;
  const int lane = threadIdx.x & 31;
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  float w[V][K0], bias[V], ga[V], be[V];
#pragma unroll
  for (int i = 0; i < V; ++i) {
    const int c = lane * V + i;
#pragma unroll
    for (int k = 0; k < K0; ++k) w[i][k] = a.w[c * K0 + k];
    bias[i] = a.b[c];
    ga[i] = a.gamma[c];
    be[i] = a.beta[c];
  }
  const long long rows = (long long)a.N * a.L0;
  for (long long r = warp; r < rows; r += nwarps) {
    const int n = (int)(r / a.L0), t = (int)(r % a.L0);
    const float* xs = a.x + (long long)n * a.W;
    float xv[K0];
#pragma unroll
    for (int k = 0; k < K0; ++k) {
      const int i = t * S0 + k - P0;
      xv[k] = (i >= 0 && i < a.W) ? __ldg(xs + i) : 0.f;
    }
    float y[V];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      float acc = bias[i];
#pragma unroll
      for (int k = 0; k < K0; ++k) acc = fmaf(w[i][k], xv[k], acc);
      y[i] = acc;
      s += acc;
    }
    const float mean = warp_sum(s) * (1.f / C);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      y[i] -= mean;
      q += y[i] * y[i];
    }
    const float rstd = rsqrtf(warp_sum(q) / (C - 1) + a.eps);
    if (lane == 0) {
      a.mean[r] = mean;
      a.rstd[r] = rstd;
    }
#pragma unroll
    for (int i = 0; i < V; ++i) y[i] = fmaxf(y[i] * rstd * ga[i] + be[i], 0.f);
    st_bf16v<V>(a.out + ((long long)n * (a.L0 + 2 * a.halo) + a.halo + t) * C + lane * V, y);
  }
}

// VA = channels per lane accumulated by one launch (register budget: 13*VA accumulators)
template <int V, int VA>
__global__ void __launch_bounds__(256) conv0_cn_relu_bwd_kernel(const Conv0Args a) {
  constexpr int C = 32 * V;
  __shared__ float red[8][32 * VA];
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const long long warp = (long long)blockIdx.x * wpb + wib;
  const long long nwarps = (long long)gridDim.x * wpb;
  const int off = a.pass * VA;  // first of this launch's channels inside the lane's V
  float w[V][K0], bias[V], ga[V], be[V];
#pragma unroll
  for (int i = 0; i < V; ++i) {
    const int c = lane * V + i;
#pragma unroll
    for (int k = 0; k < K0; ++k) w[i][k] = a.w[c * K0 + k];
    bias[i] = a.b[c];
    ga[i] = a.gamma[c];
    be[i] = a.beta[c];
  }
  float aw[VA][K0], abias[VA], ag[VA], ab[VA];
#pragma unroll
  for (int i = 0; i < VA; ++i) {
    abias[i] = ag[i] = ab[i] = 0.f;
#pragma unroll
    for (int k = 0; k < K0; ++k) aw[i][k] = 0.f;
  }
  const long long rows = (long long)a.N * a.L0;
  for (long long r = warp; r < rows; r += nwarps) {
    const int n = (int)(r / a.L0), t = (int)(r % a.L0);
    const float* xs = a.x + (long long)n * a.W;
    float xv[K0];
#pragma unroll
    for (int k = 0; k < K0; ++k) {
      const int i = t * S0 + k - P0;
      xv[k] = (i >= 0 && i < a.W) ? __ldg(xs + i) : 0.f;
    }
    float d[V], xh[V];
    ld_bf16v<V>(a.dout + ((long long)n * (a.L0 + 2 * a.halo) + a.halo + t) * C + lane * V, d);
    const float mean = a.mean[r], rstd = a.rstd[r];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      float acc = bias[i];
#pragma unroll
      for (int k = 0; k < K0; ++k) acc = fmaf(w[i][k], xv[k], acc);
      xh[i] = (acc - mean) * rstd;
      if (xh[i] * ga[i] + be[i] <= 0.f) d[i] = 0.f;
    }
#pragma unroll
    for (int i = 0; i < VA; ++i) {
      ag[i] += d[off + i] * xh[off + i];
      ab[i] += d[off + i];
    }
#pragma unroll
    for (int i = 0; i < V; ++i) {
      d[i] *= ga[i];
      s1 += d[i];
      s2 += d[i] * xh[i];
    }
    s1 = warp_sum(s1) * (1.f / C);
    s2 = warp_sum(s2) * (1.f / (C - 1));
#pragma unroll
    for (int i = 0; i < VA; ++i) {
      const float dy = rstd * (d[off + i] - s1 - xh[off + i] * s2);
      abias[i] += dy;
#pragma unroll
      for (int k = 0; k < K0; ++k) aw[i][k] = fmaf(dy, xv[k], aw[i][k]);
    }
  }
  // reduce the 13 per-channel sums over the block's warps, one quantity at a time
  auto reduce_to = [&](float (&v)[VA], float* dst, int stride, int sub) {
    __syncthreads();
#pragma unroll
    for (int i = 0; i < VA; ++i) red[wib][lane * VA + i] = v[i];
    __syncthreads();
    for (int j = threadIdx.x; j < 32 * VA; j += blockDim.x) {
      float t = 0.f;
      for (int ww = 0; ww < wpb; ++ww) t += red[ww][j];
      const int c = (j / VA) * V + off + (j % VA);
      red_add_f32(dst + (long long)c * stride + sub, t);
    }
  };
  reduce_to(abias, a.db, 1, 0);
  reduce_to(ag, a.dgamma, 1, 0);
  reduce_to(ab, a.dbeta, 1, 0);
#pragma unroll
  for (int k = 0; k < K0; ++k) {
    float col[VA];
#pragma unroll
    for (int i = 0; i < VA; ++i) col[i] = aw[i][k];
    reduce_to(col, a.dw, K0, k);
  }
}

// conv weight [Co, Ci, k] fp32 -> Wf bf16 [Co, k*Ci] (index kk*Ci + ci; forward B operand, K-major) and
// Wd bf16 [s*Ci, 2*Co] (row r*Ci + ci, column h*Co + co = W[co, ci, r + (1-h)*s]; dgrad B operand).
__global__ void conv_weight_pack_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wf,
                                        __nv_bfloat16* __restrict__ wd, int Co, int Ci, int k, int s) {
  const long long total = (long long)Co * Ci * k;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int kk = (int)(i % k);
    const int ci = (int)((i / k) % Ci);
    const int co = (int)(i / ((long long)k * Ci));
    const __nv_bfloat16 v = __float2bfloat16(w[i]);
    wf[(long long)co * k * Ci + (long long)kk * Ci + ci] = v;
    if (wd != nullptr) {
      const int h = kk >= s ? 0 : 1;
      const int r = kk >= s ? kk - s : kk;
      wd[((long long)r * Ci + ci) * (2 * Co) + (long long)h * Co + co] = v;
    }
  }
}
// inverse of the forward packing for the weight gradient: dWf fp32 [Co, k*Ci] -> dW fp32 [Co, Ci, k]
__global__ void conv_wgrad_unpack_kernel(const float* __restrict__ dwf, float* __restrict__ dw, int Co, int Ci, int k) {
  const long long total = (long long)Co * Ci * k;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int kk = (int)(i % k);
    const int ci = (int)((i / k) % Ci);
    const int co = (int)(i / ((long long)k * Ci));
    dw[i] = dwf[(long long)co * k * Ci + (long long)kk * Ci + ci];
  }
}

}  // namespace cpc

using namespace cpc;

static int fill_conv0(Conv0Args& a, const CpcConv0* d) {
  if (d == nullptr || d->N <= 0 || d->W <= 0 || d->C % 32 != 0) return CPC_ERR_ARG;
  a.x = static_cast<const float*>(d->x); a.w = static_cast<const float*>(d->w); a.b = static_cast<const float*>(d->b);
  a.gamma = static_cast<const float*>(d->gamma); a.beta = static_cast<const float*>(d->beta);
  a.out = static_cast<__nv_bfloat16*>(d->out); a.mean = static_cast<float*>(d->mean); a.rstd = static_cast<float*>(d->rstd);
  a.N = d->N; a.W = d->W; a.L0 = (d->W + 2 * P0 - K0) / S0 + 1; a.halo = d->halo; a.eps = d->eps;
  a.dout = static_cast<const __nv_bfloat16*>(d->dout);
  a.dw = static_cast<float*>(d->dw); a.db = static_cast<float*>(d->db);
  a.dgamma = static_cast<float*>(d->dgamma); a.dbeta = static_cast<float*>(d->dbeta);
  a.pass = 0;
  return CPC_OK;
}

extern "C" int cpc_conv0_cn_relu_fwd(const CpcConv0* d, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  Conv0Args a;
  int rc = fill_conv0(a, d);
  if (rc != CPC_OK) return rc;
  if (a.out == nullptr || a.mean == nullptr) return CPC_ERR_ARG;
  const long long rows = (long long)a.N * a.L0;
  long long blocks = (rows + 7) / 8;
  if (blocks > 148 * 8) blocks = 148 * 8;
  switch (d->C / 32) {
    case 1: conv0_cn_relu_fwd_kernel<1><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 2: conv0_cn_relu_fwd_kernel<2><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 4: conv0_cn_relu_fwd_kernel<4><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 8: conv0_cn_relu_fwd_kernel<8><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 16: conv0_cn_relu_fwd_kernel<16><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    default: return CPC_ERR_UNSUPPORTED;
  }
  return launch_status();
}

extern "C" int cpc_conv0_cn_relu_bwd(const CpcConv0* d, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  Conv0Args a;
  int rc = fill_conv0(a, d);
  if (rc != CPC_OK) return rc;
  if (a.dout == nullptr || a.dw == nullptr || a.db == nullptr || a.dgamma == nullptr || a.dbeta == nullptr)
    return CPC_ERR_ARG;
  const long long rows = (long long)a.N * a.L0;
  long long blocks = (rows + 8 * 16 - 1) / (8 * 16);
  if (blocks > 148 * 2) blocks = 148 * 2;
  if (blocks < 1) blocks = 1;
  switch (d->C / 32) {
    case 1: conv0_cn_relu_bwd_kernel<1, 1><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 2: conv0_cn_relu_bwd_kernel<2, 2><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 4: conv0_cn_relu_bwd_kernel<4, 4><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 8: conv0_cn_relu_bwd_kernel<8, 8><<<(unsigned)blocks, 256, 0, stream>>>(a); break;
    case 16:
      for (int p = 0; p < 2; ++p) {
        a.pass = p;
        conv0_cn_relu_bwd_kernel<16, 8><<<(unsigned)blocks, 256, 0, stream>>>(a);
      }
      break;
    default: return CPC_ERR_UNSUPPORTED;
  }
  return launch_status();
}

extern "C" int cpc_conv_weight_pack(const void* w, void* wf, void* wd, int Co, int Ci, int k, int s, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (w == nullptr || wf == nullptr || Co <= 0 || Ci <= 0 || k != 2 * s) return CPC_ERR_ARG;
  const long long total = (long long)Co * Ci * k;
  const int blocks = (int)((total + 255) / 256 > 1184 ? 1184 : (total + 255) / 256);
  conv_weight_pack_kernel<<<blocks, 256, 0, stream>>>(static_cast<const float*>(w), static_cast<__nv_bfloat16*>(wf),
                                                      static_cast<__nv_bfloat16*>(wd), Co, Ci, k, s);
  return launch_status();
}

extern "C" int cpc_conv_wgrad_unpack(const void* dwf, void* dw, int Co, int Ci, int k, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (dwf == nullptr || dw == nullptr) return CPC_ERR_ARG;
  const long long total = (long long)Co * Ci * k;
  const int blocks = (int)((total + 255) / 256 > 1184 ? 1184 : (total + 255) / 256);
  conv_wgrad_unpack_kernel<<<blocks, 256, 0, stream>>>(static_cast<const float*>(dwf), static_cast<float*>(dw), Co, Ci, k);
  return launch_status();
}
