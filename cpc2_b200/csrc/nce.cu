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
  const float* gscale;   // [K]  upstream gradient of each loss_k (backward)
  __nv_bfloat16* dlogits;  // [K][T][ld]
  __nv_bfloat16* dlogits_lo;  // optional residual: dlogits + dlogits_lo carries ~16 mantissa bits
  int K, B, W, S, NEG;
};

__global__ void __launch_bounds__(256) nce_fwd_kernel(const NceArgs a) {
  const int lane = threadIdx.x & 31;
  const int T = a.B * a.W;
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long idx = warp; idx < (long long)a.K * T; idx += nwarps) {
    const int k = (int)(idx / T);
    const int row = (int)(idx % T);
    const int b = row / a.W, t = row % a.W;
    const float* L = a.logits + ((long long)k * T + row) * a.ld;
    const float lpos = L[(long long)b * a.S + t + k + 1];
    float mx = -INFINITY;
    for (int n = lane; n < a.NEG; n += 32) {
      const long long col = a.ext[((long long)b * a.NEG + n) * a.W + t];
      mx = fmaxf(mx, L[col]);
    }
    mx = warp_max(mx);
    const float m = fmaxf(mx, lpos);
    float s = 0.f;
    for (int n = lane; n < a.NEG; n += 32) {
      const long long col = a.ext[((long long)b * a.NEG + n) * a.W + t];
      s += expf(L[col] - m);
    }
    s = warp_sum(s) + expf(lpos - m);
    if (lane == 0) {
      const float lse = m + logf(s);
      a.lse[(long long)k * T + row] = lse;
      const float w = a.weights != nullptr ? a.weights[row] : 1.f;
      red_add_f32(a.loss_sum + k, w * (lse - lpos));
      if (lpos >= mx) red_add_f32(a.hits + k, 1.f);
    }
  }
}

__global__ void __launch_bounds__(256) nce_bwd_kernel(const NceArgs a) {
  extern __shared__ float rowbuf[];  // [warps][ld]
  const int lane = threadIdx.x & 31;
  const int wib = threadIdx.x >> 5;
  float* buf = rowbuf + (long long)wib * a.ld;
  const int T = a.B * a.W;
  const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + wib;
  const long long nwarps = (long long)gridDim.x * (blockDim.x >> 5);
  for (long long idx = warp; idx < (long long)a.K * T; idx += nwarps) {
    const int k = (int)(idx / T);
    const int row = (int)(idx % T);
    const int b = row / a.W, t = row % a.W;
    const float* L = a.logits + ((long long)k * T + row) * a.ld;
    for (long long c = lane; c < a.ld; c += 32) buf[c] = 0.f;
    __syncwarp();
    const float lse = a.lse[(long long)k * T + row];
    const float w = a.weights != nullptr ? a.weights[row] : 1.f;
    const float coef = a.gscale[k] * w / (float)T;
    for (int n = lane; n < a.NEG; n += 32) {
      const long long col = a.ext[((long long)b * a.NEG + n) * a.W + t];
      atomicAdd(buf + col, coef * expf(L[col] - lse));
    }
    if (lane == 0) {
      const long long col = (long long)b * a.S + t + k + 1;
      atomicAdd(buf + col, coef * (expf(L[col] - lse) - 1.f));
    }
    __syncwarp();
    __nv_bfloat16* o = a.dlogits + ((long long)k * T + row) * a.ld;
    __nv_bfloat16* ol = a.dlogits_lo != nullptr ? a.dlogits_lo + ((long long)k * T + row) * a.ld : nullptr;
    for (long long c = lane * 2; c < a.ld; c += 64) {
      const __nv_bfloat162 hi = __floats2bfloat162_rn(buf[c], buf[c + 1]);
      *reinterpret_cast<__nv_bfloat162*>(o + c) = hi;
      // The softmax mass of the negatives clusters around 1/(NEG+1): rounding it to bf16 is a COHERENT
      // error over all rows, which survives the cancellation-heavy reductions downstream.  The residual
      // is kept as a second bf16 matrix and the two backward GEMMs run a second, accumulating pass on it.
      if (ol != nullptr)
        *reinterpret_cast<__nv_bfloat162*>(ol + c) =
            __floats2bfloat162_rn(buf[c] - __low2float(hi), buf[c + 1] - __high2float(hi));
    }
    __syncwarp();
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
  const long long items = (long long)a.K * a.B * a.W;
  long long blocks = (items + 7) / 8;
  if (blocks > 148 * 8) blocks = 148 * 8;
  nce_fwd_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream_)>>>(a);
  return launch_status();
}

extern "C" int cpc_nce_bwd(const CpcNce* d, void* stream_) {
  NceArgs a;
  int rc = fill_nce(a, d);
  if (rc != CPC_OK) return rc;
  if (a.gscale == nullptr || a.dlogits == nullptr) return CPC_ERR_ARG;
  const size_t smem = (size_t)8 * a.ld * sizeof(float);
  if (smem > 200 * 1024) return CPC_ERR_UNSUPPORTED;
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    if (cudaFuncSetAttribute(nce_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
      return CPC_ERR_LAUNCH;
    configured = smem;
  }
  const long long items = (long long)a.K * a.B * a.W;
  long long blocks = (items + 7) / 8;
  if (blocks > 148 * 4) blocks = 148 * 4;
  nce_bwd_kernel<<<(unsigned)blocks, 256, smem, static_cast<cudaStream_t>(stream_)>>>(a);
  return launch_status();
}
