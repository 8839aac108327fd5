// Encoder normalisation variants of the reference other than the default ChannelNorm (cpc/model.py:71-82):
//   normMode "batchNorm"    -> nn.BatchNorm1d      : per-channel statistics over (batch, time), running estimates
//   normMode "instanceNorm" -> nn.InstanceNorm1d(affine=True): per-(sequence, channel) statistics over time
//   normMode "ID"           -> IDModule            : no normalisation, ReLU only
// each followed by ReLU (cpc/model.py:102-107), forward and backward, on the channels-last bf16 buffers the implicit-GEMM
// convolutions read and write.  Statistics run DOWN the columns of a [rows, C] matrix (the row kernels of rownorm.cu run
// along it), so a (group, channel) pair is reduced across many CTAs:
//   1. partial : every CTA reduces one chunk of CPC_COLNORM_CHUNK_ROWS rows to per-channel partials
//   2. finalize: one thread per (group, channel) combines the chunk partials in a fixed order (Chan's pairwise
//                (count, mean, M2) update in the forward -- no E[x^2] - E[x]^2 cancellation -- plain sums in the backward)
//   3. apply   : normalise + affine + ReLU (forward) / the input gradient (backward), row maps as in rownorm.cu
// All three are HBM-bound streams over the activation (2 bytes per element read twice, written once); results do not
// depend on the launch shape (no floating-point atomics inside a group).  conv0 (1 -> C channels, k 10, stride 5) without
// the fused ChannelNorm is here too: these modes cannot use cpc_conv0_cn_relu_*.
#include "cpc_internal.h"

namespace cpc {

struct CRowMap {
  long long seq_len, seq_stride, row_off;
  __device__ __forceinline__ long long operator()(long long r) const {
    if (seq_len == 0 || (seq_stride == seq_len && row_off == 0)) return r;
    const unsigned n = (unsigned)r / (unsigned)seq_len;
    return (long long)n * seq_stride + row_off + ((unsigned)r - n * (unsigned)seq_len);
  }
};
static CRowMap to_map(const CpcRowMap& m) { return CRowMap{m.seq_len, m.seq_stride, m.row_off}; }

constexpr int CN_THREADS = 256;
constexpr int CN_CHUNK = CPC_COLNORM_CHUNK_ROWS;

struct ColNormArgs {
  long long rows, rows_per_group;
  int C, groups, chunks;
  const __nv_bfloat16* in;
  const float* gamma;
  const float* beta;
  float* mean;
  float* rstd;
  float* partial;  // [groups][chunks][C] float2
  float* running_mean;
  float* running_var;
  float momentum, eps;
  int has_stats;  // 0: identity ("ID"), 1: statistics
  int frozen;     // statistics were given (eval-mode batchNorm): the backward treats them as constants
  int relu;
  __nv_bfloat16* out_bf16;
  CRowMap out_map;
  long long out_rows;  // buffer rows of out_bf16 (sequences * seq_stride)
  float* out_f32;
  // backward
  const void* dout;
  int dout_f32;
  CRowMap dout_map;
  float* sums;  // [groups][C] float2 (S1 = sum dy, S2 = sum dy * xhat)
  __nv_bfloat16* dx_bf16;
  CRowMap dx_map;
  long long dx_rows;
  float* dx_f32;
  float* dgamma;
  float* dbeta;
  float* dxsum;
};

__device__ __forceinline__ float2 bf2(const __nv_bfloat16* p) {
  const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(p);
  return make_float2(__low2float(h), __high2float(h));
}

// thread -> (channel pair, row lane): C / 2 threads cover a row with 4-byte accesses (a warp reads 128 contiguous bytes)
struct Lanes {
  int cp, rl, rpi;  // channel pair, row lane, rows per iteration
  __device__ __forceinline__ explicit Lanes(int C) {
    const int tpr = C >> 1;
    cp = threadIdx.x % tpr;
    rl = threadIdx.x / tpr;
    rpi = CN_THREADS / tpr;
  }
};

// ---- forward statistics ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(CN_THREADS) colnorm_partial_fwd_kernel(const ColNormArgs p) {
  pdl_sync();
  __shared__ float4 red[CN_THREADS];  // (n, mean0, M2_0, -) and a second slot for channel 1 below
  __shared__ float4 red1[CN_THREADS];
  const Lanes L(p.C);
  const int g = blockIdx.y, chunk = blockIdx.x;
  const long long r0 = (long long)chunk * CN_CHUNK;
  const long long r1 = min(p.rows_per_group, r0 + CN_CHUNK);
  float n = 0.f, s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
  // sums about the first value this thread sees (keeps the partial's M2 free of cancellation inside the chunk)
  float sh0 = 0.f, sh1 = 0.f;
  bool first = true;
  for (long long r = r0 + L.rl; r < r1; r += L.rpi) {
    const float2 v = bf2(p.in + ((long long)g * p.rows_per_group + r) * p.C + 2 * L.cp);
    if (first) { sh0 = v.x; sh1 = v.y; first = false; }
    const float a = v.x - sh0, b = v.y - sh1;
    n += 1.f; s0 += a; q0 += a * a; s1 += b; q1 += b * b;
  }
  float m0 = 0.f, m1 = 0.f, M0 = 0.f, M1 = 0.f;
  if (n > 0.f) {
    const float d0 = s0 / n, d1 = s1 / n;
    m0 = sh0 + d0; m1 = sh1 + d1;
    M0 = fmaxf(q0 - s0 * d0, 0.f); M1 = fmaxf(q1 - s1 * d1, 0.f);
  }
  red[threadIdx.x] = make_float4(n, m0, M0, 0.f);
  red1[threadIdx.x] = make_float4(n, m1, M1, 0.f);
  __syncthreads();
  if (L.rl == 0) {
    float na = n, ma0 = m0, Ma0 = M0, ma1 = m1, Ma1 = M1;
    for (int j = 1; j < L.rpi; ++j) {  // fixed order: deterministic
      const float4 b0 = red[j * (p.C >> 1) + L.cp];
      const float4 b1 = red1[j * (p.C >> 1) + L.cp];
      const float nb = b0.x;
      if (nb == 0.f) continue;
      const float nt = na + nb;
      const float d0 = b0.y - ma0, d1 = b1.y - ma1;
      ma0 += d0 * (nb / nt); Ma0 += b0.z + d0 * d0 * (na * nb / nt);
      ma1 += d1 * (nb / nt); Ma1 += b1.z + d1 * d1 * (na * nb / nt);
      na = nt;
    }
    float2* out = reinterpret_cast<float2*>(p.partial) + ((long long)g * p.chunks + chunk) * p.C + 2 * L.cp;
    out[0] = make_float2(ma0, Ma0);
    out[1] = make_float2(ma1, Ma1);
  }
}

__global__ void __launch_bounds__(CN_THREADS) colnorm_finalize_fwd_kernel(const ColNormArgs p) {
  pdl_sync();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= p.groups * p.C) return;
  const int g = idx / p.C, c = idx % p.C;
  const float2* part = reinterpret_cast<const float2*>(p.partial) + (long long)g * p.chunks * p.C + c;
  float na = 0.f, ma = 0.f, Ma = 0.f;
  for (int k = 0; k < p.chunks; ++k) {
    const float2 b = part[(long long)k * p.C];
    const float nb = (float)min((long long)CN_CHUNK, p.rows_per_group - (long long)k * CN_CHUNK);
    const float nt = na + nb;
    const float d = b.x - ma;
    ma += d * (nb / nt);
    Ma += b.y + d * d * (na * nb / nt);
    na = nt;
  }
  const float var = Ma / na;  // biased: what BatchNorm1d / InstanceNorm1d normalise with
  p.mean[idx] = ma;
  p.rstd[idx] = rsqrtf(var + p.eps);
  if (p.running_mean != nullptr) {  // nn.BatchNorm1d in training mode: unbiased variance in the running estimate
    const float unb = na > 1.f ? Ma / (na - 1.f) : var;
    p.running_mean[c] = (1.f - p.momentum) * p.running_mean[c] + p.momentum * ma;
    p.running_var[c] = (1.f - p.momentum) * p.running_var[c] + p.momentum * unb;
  }
}

// one thread per (buffer row, channel pair): halo rows of the output map are cleared here
__global__ void __launch_bounds__(CN_THREADS) colnorm_apply_fwd_kernel(const ColNormArgs p) {
  pdl_sync();
  const Lanes L(p.C);
  const long long stride = (long long)gridDim.x * L.rpi;
  const bool mapped = p.out_map.seq_len != 0;
  const long long nrows = p.out_bf16 != nullptr && mapped ? p.out_rows : p.rows;
  const int c = 2 * L.cp;
  float g0 = 1.f, g1 = 1.f, b0 = 0.f, b1 = 0.f;
  if (p.gamma != nullptr) { g0 = p.gamma[c]; g1 = p.gamma[c + 1]; b0 = p.beta[c]; b1 = p.beta[c + 1]; }
  for (long long br = (long long)blockIdx.x * L.rpi + L.rl; br < nrows; br += stride) {
    long long r = br;
    if (p.out_bf16 != nullptr && mapped) {
      const long long n = br / p.out_map.seq_stride;
      const long long t = br - n * p.out_map.seq_stride - p.out_map.row_off;
      if (t < 0 || t >= p.out_map.seq_len) {
        *reinterpret_cast<uint32_t*>(p.out_bf16 + br * p.C + c) = 0u;
        continue;
      }
      r = n * p.out_map.seq_len + t;
    }
    float2 v = bf2(p.in + r * p.C + c);
    if (p.has_stats) {
      const long long gi = (r / p.rows_per_group) * p.C + c;
      v.x = (v.x - p.mean[gi]) * p.rstd[gi] * g0 + b0;
      v.y = (v.y - p.mean[gi + 1]) * p.rstd[gi + 1] * g1 + b1;
    }
    if (p.relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); }
    if (p.out_bf16 != nullptr) *reinterpret_cast<__nv_bfloat162*>(p.out_bf16 + br * p.C + c) = __floats2bfloat162_rn(v.x, v.y);
    if (p.out_f32 != nullptr) *reinterpret_cast<float2*>(p.out_f32 + r * p.C + c) = v;
  }
}

// ---- backward ---------------------------------------------------------------------------------------------------------
// dy = dout masked by the ReLU (the mask is recomputed from the saved pre-norm activation), xhat = (x - mean) * rstd
__device__ __forceinline__ void load_dy_xhat(const ColNormArgs& p, long long r, int c, float g0, float g1, float b0,
                                             float b1, float2& dy, float2& xh) {
  const long long dr = p.dout_map(r);
  if (p.dout_f32) {
    dy = *reinterpret_cast<const float2*>(static_cast<const float*>(p.dout) + dr * p.C + c);
  } else {
    dy = bf2(static_cast<const __nv_bfloat16*>(p.dout) + dr * p.C + c);
  }
  const float2 x = bf2(p.in + r * p.C + c);
  float o0 = x.x, o1 = x.y;
  xh = x;
  if (p.has_stats) {
    const long long gi = (r / p.rows_per_group) * p.C + c;
    xh.x = (x.x - p.mean[gi]) * p.rstd[gi];
    xh.y = (x.y - p.mean[gi + 1]) * p.rstd[gi + 1];
    o0 = xh.x * g0 + b0;
    o1 = xh.y * g1 + b1;
  }
  if (p.relu) {
    if (!(o0 > 0.f)) dy.x = 0.f;
    if (!(o1 > 0.f)) dy.y = 0.f;
  }
}

__global__ void __launch_bounds__(CN_THREADS) colnorm_partial_bwd_kernel(const ColNormArgs p) {
  pdl_sync();
  __shared__ float4 red[CN_THREADS];
  const Lanes L(p.C);
  const int g = blockIdx.y, chunk = blockIdx.x;
  const long long r0 = (long long)chunk * CN_CHUNK;
  const long long r1 = min(p.rows_per_group, r0 + CN_CHUNK);
  const int c = 2 * L.cp;
  float g0 = 1.f, g1 = 1.f, b0 = 0.f, b1 = 0.f;
  if (p.gamma != nullptr) { g0 = p.gamma[c]; g1 = p.gamma[c + 1]; b0 = p.beta[c]; b1 = p.beta[c + 1]; }
  float a0 = 0.f, a1 = 0.f, e0 = 0.f, e1 = 0.f;
  for (long long r = r0 + L.rl; r < r1; r += L.rpi) {
    float2 dy, xh;
    load_dy_xhat(p, (long long)g * p.rows_per_group + r, c, g0, g1, b0, b1, dy, xh);
    a0 += dy.x; a1 += dy.y; e0 += dy.x * xh.x; e1 += dy.y * xh.y;
  }
  red[threadIdx.x] = make_float4(a0, a1, e0, e1);
  __syncthreads();
  if (L.rl == 0) {
    for (int j = 1; j < L.rpi; ++j) {
      const float4 b = red[j * (p.C >> 1) + L.cp];
      a0 += b.x; a1 += b.y; e0 += b.z; e1 += b.w;
    }
    float2* out = reinterpret_cast<float2*>(p.partial) + ((long long)g * p.chunks + chunk) * p.C + c;
    out[0] = make_float2(a0, e0);
    out[1] = make_float2(a1, e1);
  }
}

__global__ void __launch_bounds__(CN_THREADS) colnorm_finalize_bwd_kernel(const ColNormArgs p) {
  pdl_sync();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= p.groups * p.C) return;
  const int g = idx / p.C, c = idx % p.C;
  const float2* part = reinterpret_cast<const float2*>(p.partial) + (long long)g * p.chunks * p.C + c;
  float S1 = 0.f, S2 = 0.f;
  for (int k = 0; k < p.chunks; ++k) {
    const float2 b = part[(long long)k * p.C];
    S1 += b.x; S2 += b.y;
  }
  reinterpret_cast<float2*>(p.sums)[idx] = make_float2(S1, S2);
  if (p.has_stats && p.dgamma != nullptr) {
    atomicAdd(p.dgamma + c, S2);
    atomicAdd(p.dbeta + c, S1);
  }
  if (p.dxsum != nullptr) {
    // column sum of dx = bias gradient of the convolution in front.  Batch / instance statistics remove the column mean,
    // so that sum is zero (the reference's autograd returns rounding noise there); with given statistics dx = g rstd dy
    float v = 0.f;
    if (!p.has_stats) v = S1;
    else if (p.frozen) v = (p.gamma != nullptr ? p.gamma[c] : 1.f) * p.rstd[idx] * S1;
    if (v != 0.f) atomicAdd(p.dxsum + c, v);
  }
}

__global__ void __launch_bounds__(CN_THREADS) colnorm_apply_bwd_kernel(const ColNormArgs p) {
  pdl_sync();
  const Lanes L(p.C);
  const long long stride = (long long)gridDim.x * L.rpi;
  const bool mapped = p.dx_bf16 != nullptr && p.dx_map.seq_len != 0;
  const long long nrows = mapped ? p.dx_rows : p.rows;
  const int c = 2 * L.cp;
  float g0 = 1.f, g1 = 1.f, b0 = 0.f, b1 = 0.f;
  if (p.gamma != nullptr) { g0 = p.gamma[c]; g1 = p.gamma[c + 1]; b0 = p.beta[c]; b1 = p.beta[c + 1]; }
  const float inv_n = 1.f / (float)p.rows_per_group;
  for (long long br = (long long)blockIdx.x * L.rpi + L.rl; br < nrows; br += stride) {
    long long r = br;
    if (mapped) {
      const long long n = br / p.dx_map.seq_stride;
      const long long t = br - n * p.dx_map.seq_stride - p.dx_map.row_off;
      if (t < 0 || t >= p.dx_map.seq_len) {
        *reinterpret_cast<uint32_t*>(p.dx_bf16 + br * p.C + c) = 0u;
        continue;
      }
      r = n * p.dx_map.seq_len + t;
    }
    float2 dy, xh;
    load_dy_xhat(p, r, c, g0, g1, b0, b1, dy, xh);
    float2 dx = dy;
    if (p.has_stats) {
      const long long gi = (r / p.rows_per_group) * p.C + c;
      if (p.frozen) {
        dx.x = g0 * p.rstd[gi] * dy.x;
        dx.y = g1 * p.rstd[gi + 1] * dy.y;
      } else {
        const float2 sa = reinterpret_cast<const float2*>(p.sums)[gi];
        const float2 sb = reinterpret_cast<const float2*>(p.sums)[gi + 1];
        dx.x = g0 * p.rstd[gi] * (dy.x - (sa.x + xh.x * sa.y) * inv_n);
        dx.y = g1 * p.rstd[gi + 1] * (dy.y - (sb.x + xh.y * sb.y) * inv_n);
      }
    }
    if (p.dx_bf16 != nullptr) *reinterpret_cast<__nv_bfloat162*>(p.dx_bf16 + br * p.C + c) = __floats2bfloat162_rn(dx.x, dx.y);
    if (p.dx_f32 != nullptr) *reinterpret_cast<float2*>(p.dx_f32 + r * p.C + c) = dx;
  }
}

static int fill_args(const CpcColNorm* d, ColNormArgs& a) {
  if (d == nullptr || d->in == nullptr || d->rows <= 0) return CPC_ERR_ARG;
  const int C = d->C;
  if (C < 2 || (C & 1) || C / 2 > CN_THREADS || CN_THREADS % (C / 2) != 0) return CPC_ERR_UNSUPPORTED;
  if (d->rows >= (1ll << 31)) return CPC_ERR_TOO_LARGE;
  a = ColNormArgs{};
  a.rows = d->rows;
  a.rows_per_group = d->rows_per_group > 0 ? d->rows_per_group : d->rows;
  if (a.rows % a.rows_per_group != 0) return CPC_ERR_ARG;
  a.C = C;
  a.groups = (int)(a.rows / a.rows_per_group);
  a.chunks = (int)((a.rows_per_group + CN_CHUNK - 1) / CN_CHUNK);
  a.in = static_cast<const __nv_bfloat16*>(d->in);
  a.gamma = static_cast<const float*>(d->gamma);
  a.beta = static_cast<const float*>(d->beta);
  if ((a.gamma == nullptr) != (a.beta == nullptr)) return CPC_ERR_ARG;
  a.mean = static_cast<float*>(d->mean);
  a.rstd = static_cast<float*>(d->rstd);
  a.partial = static_cast<float*>(d->partial);
  a.running_mean = static_cast<float*>(d->running_mean);
  a.running_var = static_cast<float*>(d->running_var);
  a.momentum = d->momentum;
  a.eps = d->eps;
  a.has_stats = d->mode != CPC_COLNORM_ID;
  a.frozen = d->mode == CPC_COLNORM_GIVEN;
  if (d->mode < CPC_COLNORM_ID || d->mode > CPC_COLNORM_GIVEN) return CPC_ERR_ARG;
  if (a.has_stats && (a.mean == nullptr || a.rstd == nullptr)) return CPC_ERR_ARG;
  if (d->mode == CPC_COLNORM_STATS && a.partial == nullptr) return CPC_ERR_ARG;
  if (a.running_mean != nullptr && (a.groups != 1 || a.running_var == nullptr)) return CPC_ERR_ARG;
  a.relu = d->relu;
  return CPC_OK;
}

static int apply_grid(long long rows, int C) {
  const int rpi = CN_THREADS / (C / 2);
  const long long want = (rows + rpi - 1) / rpi;
  const long long cap = 148 * 8;
  return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace cpc

using namespace cpc;

extern "C" int cpc_colnorm_fwd(const CpcColNorm* d, void* stream) {
  ColNormArgs a;
  const int rc = fill_args(d, a);
  if (rc != CPC_OK) return rc;
  if (d->out_bf16 == nullptr && d->out_f32 == nullptr) return CPC_ERR_ARG;
  a.out_bf16 = static_cast<__nv_bfloat16*>(d->out_bf16);
  a.out_f32 = static_cast<float*>(d->out_f32);
  a.out_map = to_map(d->out_map);
  long long out_rows = a.rows;
  if (a.out_map.seq_len != 0) {
    if (a.rows % a.out_map.seq_len != 0 || a.out_map.row_off + a.out_map.seq_len > a.out_map.seq_stride) return CPC_ERR_ARG;
    out_rows = (a.rows / a.out_map.seq_len) * a.out_map.seq_stride;
  }
  a.out_rows = out_rows;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (d->mode == CPC_COLNORM_STATS) {
    launch_k(colnorm_partial_fwd_kernel, dim3(a.chunks, a.groups), dim3(CN_THREADS), 0, s, a);
    launch_k(colnorm_finalize_fwd_kernel, dim3((a.groups * a.C + CN_THREADS - 1) / CN_THREADS), dim3(CN_THREADS), 0, s, a);
  }
  launch_k(colnorm_apply_fwd_kernel, dim3(apply_grid(out_rows, a.C)), dim3(CN_THREADS), 0, s, a);
  return launch_status();
}

extern "C" int cpc_colnorm_bwd(const CpcColNorm* d, void* stream) {
  ColNormArgs a;
  const int rc = fill_args(d, a);
  if (rc != CPC_OK) return rc;
  if (d->dout == nullptr || d->partial == nullptr || d->sums == nullptr) return CPC_ERR_ARG;
  if (d->dx_bf16 == nullptr && d->dx_f32 == nullptr) return CPC_ERR_ARG;
  a.dout = d->dout;
  a.dout_f32 = d->dout_f32;
  a.dout_map = to_map(d->dout_map);
  a.sums = static_cast<float*>(d->sums);
  a.dx_bf16 = static_cast<__nv_bfloat16*>(d->dx_bf16);
  a.dx_f32 = static_cast<float*>(d->dx_f32);
  a.dx_map = to_map(d->dx_map);
  long long dx_rows = a.rows;
  if (a.dx_map.seq_len != 0) {
    if (a.rows % a.dx_map.seq_len != 0 || a.dx_map.row_off + a.dx_map.seq_len > a.dx_map.seq_stride) return CPC_ERR_ARG;
    dx_rows = (a.rows / a.dx_map.seq_len) * a.dx_map.seq_stride;
  }
  a.dx_rows = dx_rows;
  a.dgamma = static_cast<float*>(d->dgamma);
  a.dbeta = static_cast<float*>(d->dbeta);
  if ((a.dgamma == nullptr) != (a.dbeta == nullptr)) return CPC_ERR_ARG;
  a.dxsum = static_cast<float*>(d->dxsum);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  launch_k(colnorm_partial_bwd_kernel, dim3(a.chunks, a.groups), dim3(CN_THREADS), 0, s, a);
  launch_k(colnorm_finalize_bwd_kernel, dim3((a.groups * a.C + CN_THREADS - 1) / CN_THREADS), dim3(CN_THREADS), 0, s, a);
  launch_k(colnorm_apply_bwd_kernel, dim3(apply_grid(dx_rows, a.C)), dim3(CN_THREADS), 0, s, a);
  return launch_status();
}

// ---- conv0 without a fused norm ---------------------------------------------------------------------------------------
namespace cpc {

constexpr int C0_ROWS = 64;       // output frames per CTA, forward
constexpr int C0_ROWS_BWD = 256;  // backward: fewer CTAs, fewer atomics per weight
constexpr int C0_K = 10, C0_S = 5, C0_P = 3;

struct Conv0PlainArgs {
  int N, W, C, L0;
  long long x_stride;
  const float* x;
  const float* w;
  const float* b;
  __nv_bfloat16* y;  // [N * L0, C]
  const float* dy;   // [N * L0, C] fp32
  float* dw;         // [C, 10]
};

template <int ROWS>
__device__ __forceinline__ void stage_x(const Conv0PlainArgs& p, int n, int t0, float* xs) {
  const int nx = ROWS * C0_S + C0_K - C0_S;
  const long long base = (long long)t0 * C0_S - C0_P;
  for (int i = threadIdx.x; i < nx; i += blockDim.x) {
    const long long j = base + i;
    xs[i] = (j >= 0 && j < p.W) ? p.x[(long long)n * p.x_stride + j] : 0.f;
  }
}

__global__ void __launch_bounds__(256) conv0_plain_fwd_kernel(const Conv0PlainArgs p) {
  pdl_sync();
  __shared__ float xs[C0_ROWS * C0_S + C0_K - C0_S];
  const int n = blockIdx.y, t0 = blockIdx.x * C0_ROWS;
  stage_x<C0_ROWS>(p, n, t0, xs);
  __syncthreads();
  const int t1 = min(C0_ROWS, p.L0 - t0);
  for (int c = threadIdx.x; c < p.C; c += blockDim.x) {
    float w[C0_K];
#pragma unroll
    for (int j = 0; j < C0_K; ++j) w[j] = p.w[c * C0_K + j];
    const float bias = p.b != nullptr ? p.b[c] : 0.f;
    for (int t = 0; t < t1; ++t) {
      float acc = bias;
#pragma unroll
      for (int j = 0; j < C0_K; ++j) acc = fmaf(w[j], xs[t * C0_S + j], acc);
      p.y[((long long)n * p.L0 + t0 + t) * p.C + c] = __float2bfloat16(acc);
    }
  }
}

// dw[c, j] += sum_{n, t} dy[n, t, c] * x[n, 5 t - 3 + j]: per-CTA register partials, one atomic per (c, j) and CTA
__global__ void __launch_bounds__(256) conv0_plain_bwd_kernel(const Conv0PlainArgs p) {
  pdl_sync();
  __shared__ float xs[C0_ROWS_BWD * C0_S + C0_K - C0_S];
  const int n = blockIdx.y, t0 = blockIdx.x * C0_ROWS_BWD;
  stage_x<C0_ROWS_BWD>(p, n, t0, xs);
  __syncthreads();
  const int t1 = min(C0_ROWS_BWD, p.L0 - t0);
  for (int c = threadIdx.x; c < p.C; c += blockDim.x) {
    float acc[C0_K];
#pragma unroll
    for (int j = 0; j < C0_K; ++j) acc[j] = 0.f;
    for (int t = 0; t < t1; ++t) {
      const float g = p.dy[((long long)n * p.L0 + t0 + t) * p.C + c];
#pragma unroll
      for (int j = 0; j < C0_K; ++j) acc[j] = fmaf(g, xs[t * C0_S + j], acc[j]);
    }
#pragma unroll
    for (int j = 0; j < C0_K; ++j) atomicAdd(p.dw + c * C0_K + j, acc[j]);
  }
}

static int fill_conv0(const CpcConv0* d, Conv0PlainArgs& a) {
  if (d == nullptr || d->x == nullptr || d->N <= 0 || d->C <= 0 || d->W < C0_K - 2 * C0_P) return CPC_ERR_ARG;
  a = Conv0PlainArgs{};
  a.N = d->N; a.W = d->W; a.C = d->C;
  a.L0 = (d->W + 2 * C0_P - C0_K) / C0_S + 1;
  a.x_stride = d->x_stride > 0 ? d->x_stride : d->W;
  a.x = static_cast<const float*>(d->x);
  if (d->N > 65535) return CPC_ERR_TOO_LARGE;
  return CPC_OK;
}

}  // namespace cpc

extern "C" int cpc_conv0_plain_fwd(const CpcConv0* d, void* stream) {
  Conv0PlainArgs a;
  const int rc = fill_conv0(d, a);
  if (rc != CPC_OK) return rc;
  if (d->w == nullptr || d->out == nullptr) return CPC_ERR_ARG;
  a.w = static_cast<const float*>(d->w);
  a.b = static_cast<const float*>(d->b);
  a.y = static_cast<__nv_bfloat16*>(d->out);
  launch_k(conv0_plain_fwd_kernel, dim3((a.L0 + C0_ROWS - 1) / C0_ROWS, a.N), dim3(256), 0,
           static_cast<cudaStream_t>(stream), a);
  return launch_status();
}

extern "C" int cpc_conv0_plain_bwd(const CpcConv0* d, void* stream) {
  Conv0PlainArgs a;
  const int rc = fill_conv0(d, a);
  if (rc != CPC_OK) return rc;
  if (d->dout == nullptr || d->dw == nullptr) return CPC_ERR_ARG;
  a.dy = static_cast<const float*>(d->dout);
  a.dw = static_cast<float*>(d->dw);
  launch_k(conv0_plain_bwd_kernel, dim3((a.L0 + C0_ROWS_BWD - 1) / C0_ROWS_BWD, a.N), dim3(256), 0,
           static_cast<cudaStream_t>(stream), a);
  return launch_status();
}
