// InfoNCE tail of CPCUnsupersivedCriterion (cpc/criterion/criterion.py:237-286, 329-363).
//
// The reference materialises, for each of the K prediction steps, the gathered candidates
// [B, NEG+1, W, D] and a broadcast product of the same size.  Here the K logit matrices
//   Lg[k][row, col] = <pred_k[row], z[col]> / D,   row = b*W + t,  col = b'*S + s   (all pairs, tensor cores)
// are produced by cpc_gemm_bf16 and these kernels only GATHER the NEG+1 scalars a row needs:
//   positive  col = b*S + t + k          (k = 1..K)
//   negative  col = extIdx[b, n, t]      (shared by all k)
// forward : per row log-sum-exp, cross entropy against class 0, first-max argmax == 0 (exact tie rule of
//           torch.max: the positive wins ties), weighted loss sum and hit count per k.
// backward: d loss / d Lg scattered into a dense bf16 matrix (duplicates among the sampled negatives are
//           accumulated in shared memory first) which the two backward GEMMs consume.
#include "cpc_internal.h"
#include "ptx.cuh"

namespace cpc {

struct NceArgs {
  const float* logits;  // [K][T][ld]
  long long ld;
  const long long* ext;  // [B][NEG][W]
  const float* weights;  // [T] or null
  float* lse;            // [K][T]
  float* loss_sum;       // [K]  sum_rows w*ce
  float* hits;           // [K]  number of rows whose argmax is the positive
  float out_scale;       // factor applied to what is added to loss_sum / hits
  const float* gscale;   // [K]  upstream gradient of each loss_k (backward)
  __nv_bfloat16* dlogits;  // [K][T][ld]
  __nv_bfloat16* dlogits_lo;  // optional residual: dlogits + dlogits_lo carries ~16 mantissa bits
  int K, B, W, S, NEG;
};

// One warp per (row = (b, t), chunk of KU prediction steps).  The NEG sampled columns of a row are the same for
// every step k (sampleClean draws them once, criterion.py:256-266): they are loaded once into registers (J = ceil(NEG
// / 32) per lane; the index tensor is [B, NEG, W], so these loads are strided) and the gathers of KU steps are issued
// together before the first reduction -- the kernel is bound by the L2 latency of the gathers, not by bytes.  Per-step
// loss / hit sums are reduced in shared memory first: one global atomic per (block, step).
template <int J>
__global__ void __launch_bounds__(256) nce_fwd_kernel(const NceArgs a) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  constexpr int KU = 4;
  extern __shared__ float acc[];  // [2][K]
  const int lane = threadIdx.x & 31;
  const int T = a.B * a.W;
  const int chunks = (a.K + KU - 1) / KU;
  for (int i = threadIdx.x; i < 2 * a.K; i += blockDim.x) acc[i] = 0.f;
  __syncthreads();
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long idx = warp; idx < (long long)chunks * T; idx += nwarps) {
    const int row = (int)(idx % T);
    const int k0 = (int)(idx / T) * KU;
    const int b = row / a.W, t = row % a.W;
    long long col[J];
#pragma unroll
    for (int j = 0; j < J; ++j) {
      const int n = lane + 32 * j;
      col[j] = n < a.NEG ? a.ext[((long long)b * a.NEG + n) * a.W + t] : -1;
    }
    float v[KU][J], lpos[KU];
#pragma unroll
    for (int u = 0; u < KU; ++u) {
      const int k = k0 + u < a.K ? k0 + u : a.K - 1;
      const float* L = a.logits + ((long long)k * T + row) * a.ld;
      lpos[u] = L[(long long)b * a.S + t + k + 1];
#pragma unroll
      for (int j = 0; j < J; ++j) v[u][j] = col[j] >= 0 ? L[col[j]] : -INFINITY;
    }
    const float w = a.weights != nullptr ? a.weights[row] : 1.f;
#pragma unroll
    for (int u = 0; u < KU; ++u) {
      const int k = k0 + u;
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < J; ++j) mx = fmaxf(mx, v[u][j]);
      mx = warp_max(mx);
      const float m = fmaxf(mx, lpos[u]);
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < J; ++j) s += expf(v[u][j] - m);  // exp(-inf) = 0 for the lanes past NEG
      s = warp_sum(s) + expf(lpos[u] - m);
      if (lane == 0 && k < a.K) {
        const float lse = m + logf(s);
        a.lse[(long long)k * T + row] = lse;
        atomicAdd(acc + k, w * (lse - lpos[u]));
        if (lpos[u] >= mx) atomicAdd(acc + a.K + k, 1.f);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * a.K; i += blockDim.x) {
    const float x = acc[i];
    if (x != 0.f) red_add_f32((i < a.K ? a.loss_sum + i : a.hits + (i - a.K)), x * a.out_scale);
  }
}

// Same work split as the forward: one warp per (row, chunk of KU steps), the row's sampled columns held in
// registers across the steps of the chunk.
template <int J>
__global__ void __launch_bounds__(256) nce_bwd_kernel(const NceArgs a) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  constexpr int KU = 4;
  extern __shared__ float rowbuf[];  // [warps][ld]
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5;
  float* buf = rowbuf + (long long)wib * a.ld;
  const int T = a.B * a.W;
  const int chunks = (a.K + KU - 1) / KU;
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + wib;
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long idx = warp; idx < (long long)chunks * T; idx += nwarps) {
    const int row = (int)(idx % T);
    const int k0 = (int)(idx / T) * KU;
    const int b = row / a.W, t = row % a.W;
    long long col[J];
#pragma unroll
    for (int j = 0; j < J; ++j) {
      const int n = lane + 32 * j;
      col[j] = n < a.NEG ? a.ext[((long long)b * a.NEG + n) * a.W + t] : -1;
    }
    const float w = a.weights != nullptr ? a.weights[row] : 1.f;
    for (int k = k0; k < k0 + KU && k < a.K; ++k) {
      const float* L = a.logits + ((long long)k * T + row) * a.ld;
      float v[J];
#pragma unroll
      for (int j = 0; j < J; ++j) v[j] = col[j] >= 0 ? L[col[j]] : 0.f;
      const long long cpos = (long long)b * a.S + t + k + 1;
      const float lpos = L[cpos];
      const float lse = a.lse[(long long)k * T + row];
      const float coef = a.gscale[k] * w / (float)T;
      for (long long c = lane * 4; c < a.ld; c += 128) *reinterpret_cast<float4*>(buf + c) = make_float4(0.f, 0.f, 0.f, 0.f);
      __syncwarp();
#pragma unroll
      for (int j = 0; j < J; ++j)
        if (col[j] >= 0) atomicAdd(buf + col[j], coef * expf(v[j] - lse));  // a column can be drawn more than once
      if (lane == 0) atomicAdd(buf + cpos, coef * (expf(lpos - lse) - 1.f));
      __syncwarp();
      __nv_bfloat16* o = a.dlogits + ((long long)k * T + row) * a.ld;
      __nv_bfloat16* ol = a.dlogits_lo != nullptr ? a.dlogits_lo + ((long long)k * T + row) * a.ld : nullptr;
      for (long long c = lane * 4; c < a.ld; c += 128) {
        const float4 x = *reinterpret_cast<const float4*>(buf + c);
        const __nv_bfloat162 h0 = __floats2bfloat162_rn(x.x, x.y), h1 = __floats2bfloat162_rn(x.z, x.w);
        uint2 hi;
        hi.x = *reinterpret_cast<const uint32_t*>(&h0);
        hi.y = *reinterpret_cast<const uint32_t*>(&h1);
        *reinterpret_cast<uint2*>(o + c) = hi;
        // The softmax mass of the negatives clusters around 1/(NEG+1): rounding it to bf16 is a COHERENT
        // error over all rows, which survives the cancellation-heavy reductions downstream.  The residual
        // is kept as a second bf16 matrix and the two backward GEMMs run a second, accumulating pass on it.
        if (ol != nullptr) {
          const __nv_bfloat162 l0 = __floats2bfloat162_rn(x.x - __low2float(h0), x.y - __high2float(h0));
          const __nv_bfloat162 l1 = __floats2bfloat162_rn(x.z - __low2float(h1), x.w - __high2float(h1));
          uint2 lo;
          lo.x = *reinterpret_cast<const uint32_t*>(&l0);
          lo.y = *reinterpret_cast<const uint32_t*>(&l1);
          *reinterpret_cast<uint2*>(ol + c) = lo;
        }
      }
      __syncwarp();
    }
  }
}

}  // namespace cpc

using namespace cpc;

static int fill_nce(NceArgs& a, const CpcNce* d) {
  if (d == nullptr || d->K <= 0 || d->B <= 0 || d->W <= 0 || d->S <= 0 || d->NEG <= 0) return CPC_ERR_ARG;
  if (d->logits == nullptr || d->ext == nullptr || d->lse == nullptr) return CPC_ERR_ARG;
  if (d->W + d->K > d->S || d->ld < (long long)d->B * d->S || d->ld % 2 != 0) return CPC_ERR_ARG;
  a.logits = static_cast<const float*>(d->logits); a.ld = d->ld;
  a.ext = static_cast<const long long*>(d->ext);
  a.weights = static_cast<const float*>(d->weights);
  a.lse = static_cast<float*>(d->lse);
  a.loss_sum = static_cast<float*>(d->loss_sum); a.hits = static_cast<float*>(d->hits);
  a.out_scale = d->out_scale != 0.f ? d->out_scale : 1.f;
  a.gscale = static_cast<const float*>(d->gscale);
  a.dlogits = static_cast<__nv_bfloat16*>(d->dlogits);
  a.dlogits_lo = static_cast<__nv_bfloat16*>(d->dlogits_lo);
  a.K = d->K; a.B = d->B; a.W = d->W; a.S = d->S; a.NEG = d->NEG;
  return CPC_OK;
}

extern "C" int cpc_nce_fwd(const CpcNce* d, void* stream_) {
  NceArgs a;
  int rc = fill_nce(a, d);
  if (rc != CPC_OK) return rc;
  if (a.loss_sum == nullptr || a.hits == nullptr) return CPC_ERR_ARG;
  const long long items = (long long)((a.K + 3) / 4) * a.B * a.W;  // (row, chunk of 4 steps) per warp
  long long blocks = (items + 7) / 8;
  if (blocks > 148 * 4) blocks = 148 * 4;
  const size_t smem = (size_t)2 * a.K * sizeof(float);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int J = (a.NEG + 31) / 32;
  if (J <= 1) launch_k(nce_fwd_kernel<1>, dim3((unsigned)blocks), dim3(256), smem, stream, a);
  else if (J <= 2) launch_k(nce_fwd_kernel<2>, dim3((unsigned)blocks), dim3(256), smem, stream, a);
  else if (J <= 4) launch_k(nce_fwd_kernel<4>, dim3((unsigned)blocks), dim3(256), smem, stream, a);
  else if (J <= 8) launch_k(nce_fwd_kernel<8>, dim3((unsigned)blocks), dim3(256), smem, stream, a);
  else if (J <= 16) launch_k(nce_fwd_kernel<16>, dim3((unsigned)blocks), dim3(256), smem, stream, a);
  else if (J <= 32) launch_k(nce_fwd_kernel<32>, dim3((unsigned)blocks), dim3(256), smem, stream, a);
  else return CPC_ERR_UNSUPPORTED;  // more than 1024 negatives per token
  return launch_status();
}

extern "C" int cpc_nce_bwd(const CpcNce* d, void* stream_) {
  NceArgs a;
  int rc = fill_nce(a, d);
  if (rc != CPC_OK) return rc;
  if (a.gscale == nullptr || a.dlogits == nullptr) return CPC_ERR_ARG;
  const size_t smem = (size_t)8 * a.ld * sizeof(float);
  if (smem > 200 * 1024) return CPC_ERR_UNSUPPORTED;
  const long long items = (long long)((a.K + 3) / 4) * a.B * a.W;
  long long blocks = (items + 7) / 8;
  if (blocks > 148 * 4) blocks = 148 * 4;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int J = (a.NEG + 31) / 32;
  if (J > 32 || a.ld % 4 != 0) return CPC_ERR_UNSUPPORTED;  // > 1024 negatives; rows of B*S logits move as float4
#define CPC_NCE_BWD(JJ)                                                                                              \
  do {                                                                                                               \
    if (smem > 48 * 1024 &&                                                                                          \
        cudaFuncSetAttribute(nce_bwd_kernel<JJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) \
      return CPC_ERR_LAUNCH;                                                                                         \
    launch_k(nce_bwd_kernel<JJ>, dim3((unsigned)blocks), dim3(256), smem, stream, a);                                                  \
  } while (0)
  if (J <= 1) CPC_NCE_BWD(1);
  else if (J <= 2) CPC_NCE_BWD(2);
  else if (J <= 4) CPC_NCE_BWD(4);
  else if (J <= 8) CPC_NCE_BWD(8);
  else if (J <= 16) CPC_NCE_BWD(16);
  else CPC_NCE_BWD(32);
#undef CPC_NCE_BWD
  return launch_status();
}
