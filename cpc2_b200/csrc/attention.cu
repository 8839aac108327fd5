// Causal, block-local multi-head attention with learned relative-position keys
// (cpc/transformers.py:10-70 ScaledDotProductAttention, :73-104 MultiHeadAttention), forward and backward.
//
//   score[i,j] = (q_i.k_j + q_i.Krelpos[:, S-1-(i-j)]) / sqrt(dk)   for j <= i, -inf otherwise
//   out_i      = sum_j dropout(softmax_j(score[i,:])) v_j
//
// The reference builds the relative term with a concatenate/view "skew" of a full [S,S] product; here the
// shifted column is simply indexed.  One CTA per (group, sequence, head); the whole S x dk problem
// (S <= 256, dk <= 64) sits in shared memory, one warp per query row.  Total work is ~2 GFLOP per step
//  so CUDA cores suffice; the probabilities are kept (bf16, lower triangle) for the
// backward instead of being recomputed.  Dropout masks come from a counter RNG keyed on
// (seed, site, element index), so forward and backward regenerate the same mask with no mask tensor.
#include "cpc_internal.h"
#include "ptx.cuh"
#include "attn_args.h"

namespace cpc {


__device__ __forceinline__ float ldf(const float* p) { return *p; }
__device__ __forceinline__ float ldf(const __nv_bfloat16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void stf(float* p, float v) { *p = v; }
__device__ __forceinline__ void stf(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }

template <typename T>
__global__ void __launch_bounds__(256) attn_fwd_kernel(const AttnArgs a) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  extern __shared__ __align__(16) uint8_t smem_dyn[];
  const int S = a.S, dk = a.dk, SP = S | 1, DP = dk | 1;
  T* Kt = reinterpret_cast<T*>(smem_dyn);  // [dk][SP]
  T* Kr = Kt + dk * SP;                    // [dk][SP]
  T* V = Kr + dk * SP;                     // [S][DP]
  float* qrow = reinterpret_cast<float*>(V + S * DP + ((S * DP) & 1));  // [8][dk]
  float* prow = qrow + 8 * dk;                                          // [8][S]
  const int h = blockIdx.x, b = blockIdx.y, g = blockIdx.z;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const __nv_bfloat16* base = a.qkv + (long long)g * a.gs_qkv + (long long)b * S * 3 * a.d + h * dk;
  for (int i = threadIdx.x; i < S * dk; i += blockDim.x) {
    const int s = i / dk, c = i % dk;
    stf(&Kt[c * SP + s], __bfloat162float(base[(long long)s * 3 * a.d + a.d + c]));
    stf(&V[s * DP + c], __bfloat162float(base[(long long)s * 3 * a.d + 2 * a.d + c]));
    if (a.krel != nullptr) {
      const int cc = i / S, ss = i % S;
      stf(&Kr[cc * SP + ss], a.krel[((long long)g * dk + cc) * S + ss]);
    }
  }
  __syncthreads();
  unsigned long long seed = 0;
  uint32_t thresh = 0;
  float dscale = 1.f;
  if (a.rng != nullptr && a.drop_p > 0.f) {
    seed = a.rng[0] + a.rng[1] * 0x632BE59BD9B4E019ull;
    thresh = static_cast<uint32_t>(fminf(a.drop_p, 0.999999f) * 4294967296.0f);
    dscale = 1.f / (1.f - a.drop_p);
  }
  const long long pbase = (((long long)g * a.Bt + b) * a.Hh + h) * S * S;
  float* q = qrow + warp * dk;
  float* p = prow + warp * S;
  for (int i = warp; i < S; i += 8) {
    for (int c = lane; c < dk; c += 32) q[c] = __bfloat162float(base[(long long)i * 3 * a.d + c]);
    __syncwarp();
    float mx = -INFINITY;
    for (int j = lane; j <= i; j += 32) {
      float s = 0.f;
      if (a.krel != nullptr) {
        const int r = S - 1 - i + j;
        for (int c = 0; c < dk; ++c) s = fmaf(q[c], ldf(&Kt[c * SP + j]) + ldf(&Kr[c * SP + r]), s);
      } else {
        for (int c = 0; c < dk; ++c) s = fmaf(q[c], ldf(&Kt[c * SP + j]), s);
      }
      s *= a.scale;
      p[j] = s;
      mx = fmaxf(mx, s);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int j = lane; j <= i; j += 32) {
      const float e = __expf(p[j] - mx);
      p[j] = e;
      sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    for (int j = lane; j <= i; j += 32) {
      float pr = p[j] * inv;
      if (a.probs != nullptr) a.probs[pbase + (long long)i * S + j] = __float2bfloat16(pr);
      if (thresh != 0) pr = rng_u32(seed, a.site, pbase + (long long)i * S + j) >= thresh ? pr * dscale : 0.f;
      p[j] = pr;
    }
    __syncwarp();
    for (int c = lane; c < dk; c += 32) {
      float acc = 0.f;
      for (int j = 0; j <= i; ++j) acc = fmaf(p[j], ldf(&V[j * DP + c]), acc);
      a.out[(long long)g * a.gs_out + ((long long)b * S + i) * a.d + h * dk + c] = __float2bfloat16(acc);
      if (a.out_f32 != nullptr) a.out_f32[(long long)g * a.gs_out + ((long long)b * S + i) * a.d + h * dk + c] = acc;
    }
    __syncwarp();
  }
}

template <typename T>
__global__ void __launch_bounds__(256) attn_bwd_kernel(const AttnArgs a) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  extern __shared__ __align__(16) uint8_t smem_dyn[];
  const int S = a.S, dk = a.dk, SP = S | 1, DP = dk | 1;
  T* Kt = reinterpret_cast<T*>(smem_dyn);  // [dk][SP]
  T* Kr = Kt + dk * SP;                    // [dk][SP]
  T* Vt = Kr + dk * SP;                    // [dk][SP]
  T* Q = Vt + dk * SP;                     // [S][DP]
  T* dO = Q + S * DP;                      // [S][DP]
  float* dsrow = reinterpret_cast<float*>(dO + S * DP + ((3 * dk * SP + 2 * S * DP) & 1));  // [8][S]
  const int h = blockIdx.x, b = blockIdx.y, g = blockIdx.z;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const __nv_bfloat16* base = a.qkv + (long long)g * a.gs_qkv + (long long)b * S * 3 * a.d + h * dk;
  const long long obase = (long long)g * a.gs_out + (long long)b * S * a.d + h * dk;
  for (int i = threadIdx.x; i < S * dk; i += blockDim.x) {
    const int s = i / dk, c = i % dk;
    stf(&Q[s * DP + c], __bfloat162float(base[(long long)s * 3 * a.d + c]));
    stf(&Kt[c * SP + s], __bfloat162float(base[(long long)s * 3 * a.d + a.d + c]));
    stf(&Vt[c * SP + s], __bfloat162float(base[(long long)s * 3 * a.d + 2 * a.d + c]));
    stf(&dO[s * DP + c], __bfloat162float(a.dout[obase + (long long)s * a.d + c]));
    if (a.krel != nullptr) {
      const int cc = i / S, ss = i % S;
      stf(&Kr[cc * SP + ss], a.krel[((long long)g * dk + cc) * S + ss]);
    }
  }
  __syncthreads();
  unsigned long long seed = 0;
  uint32_t thresh = 0;
  float dscale = 1.f;
  if (a.rng != nullptr && a.drop_p > 0.f) {
    seed = a.rng[0] + a.rng[1] * 0x632BE59BD9B4E019ull;
    thresh = static_cast<uint32_t>(fminf(a.drop_p, 0.999999f) * 4294967296.0f);
    dscale = 1.f / (1.f - a.drop_p);
  }
  const long long pbase = (((long long)g * a.Bt + b) * a.Hh + h) * S * S;
  float* ds = dsrow + warp * S;
  __nv_bfloat16* dq_out = a.dqkv + (long long)g * a.gs_qkv + (long long)b * S * 3 * a.d + h * dk;

  // ---- phase A: one warp per query row: dS row, dQ row
  for (int i = warp; i < S; i += 8) {
    // D_i = sum_j p_ij dp~_ij / sum_j p_ij from the same (bf16) p and the same dp the row's dS uses: the row of dS
    // then sums to zero to fp32 rounding (see attention_tc.cu: dO_i . O_i is only as accurate as the stored O)
    float dsum = 0.f, psum = 0.f;
    for (int j = lane; j <= i; j += 32) {
      float dp = 0.f;
      for (int c = 0; c < dk; ++c) dp = fmaf(ldf(&dO[i * DP + c]), ldf(&Vt[c * SP + j]), dp);
      const long long e = pbase + (long long)i * S + j;
      if (thresh != 0) dp = rng_u32(seed, a.site, e) >= thresh ? dp * dscale : 0.f;
      const float pr = __bfloat162float(a.probs[e]);
      ds[j] = dp;
      dsum = fmaf(pr, dp, dsum);
      psum += pr;
    }
    dsum = warp_sum(dsum);
    psum = warp_sum(psum);
    dsum = psum > 0.f ? dsum / psum : 0.f;
    for (int j = lane; j <= i; j += 32) {
      const long long e = pbase + (long long)i * S + j;
      const float v = __bfloat162float(a.probs[e]) * (ds[j] - dsum) * a.scale;
      ds[j] = v;
      a.dscore[e] = __float2bfloat16(v);
    }
    __syncwarp();
    for (int c = lane; c < dk; c += 32) {
      float acc = 0.f;
      if (a.krel != nullptr) {
        for (int j = 0; j <= i; ++j) acc = fmaf(ds[j], ldf(&Kt[c * SP + j]) + ldf(&Kr[c * SP + S - 1 - i + j]), acc);
      } else {
        for (int j = 0; j <= i; ++j) acc = fmaf(ds[j], ldf(&Kt[c * SP + j]), acc);
      }
      dq_out[(long long)i * 3 * a.d + c] = __float2bfloat16(acc);
    }
    __syncwarp();
  }
  __syncthreads();  // dscore rows written by other warps of this CTA are now visible

  // ---- phase B: one warp per key row j: dK_j, dV_j
  for (int j = warp; j < S; j += 8) {
    for (int c = lane; c < dk; c += 32) {
      float dkacc = 0.f, dvacc = 0.f;
      for (int i = j; i < S; ++i) {
        const long long e = pbase + (long long)i * S + j;
        const float dsv = __bfloat162float(a.dscore[e]);
        float pr = __bfloat162float(a.probs[e]);
        if (thresh != 0) pr = rng_u32(seed, a.site, e) >= thresh ? pr * dscale : 0.f;
        dkacc = fmaf(dsv, ldf(&Q[i * DP + c]), dkacc);
        dvacc = fmaf(pr, ldf(&dO[i * DP + c]), dvacc);
      }
      dq_out[(long long)j * 3 * a.d + a.d + c] = __float2bfloat16(dkacc);
      dq_out[(long long)j * 3 * a.d + 2 * a.d + c] = __float2bfloat16(dvacc);
    }
  }
  // ---- phase C: relative-position keys: dKrel[c][r] += sum_{i >= delta} dS[i][i-delta] * q_i[c], delta = S-1-r
  if (a.krel != nullptr && a.dkrel != nullptr) {
    for (int r = warp; r < S; r += 8) {
      const int delta = S - 1 - r;
      for (int c = lane; c < dk; c += 32) {
        float acc = 0.f;
        for (int i = delta; i < S; ++i)
          acc = fmaf(__bfloat162float(a.dscore[pbase + (long long)i * S + (i - delta)]), ldf(&Q[i * DP + c]), acc);
        red_add_f32(a.dkrel + ((long long)g * dk + c) * S + r, acc);
      }
    }
  }
}

template <typename T>
static size_t attn_smem(int S, int dk, bool bwd) {
  const int SP = S | 1, DP = dk | 1;
  if (!bwd) return (size_t)(2 * dk * SP + S * DP + 1) * sizeof(T) + (size_t)(8 * dk + 8 * S) * 4 + 16;
  return (size_t)(3 * dk * SP + 2 * S * DP + 1) * sizeof(T) + (size_t)(8 * S) * 4 + 16;
}

}  // namespace cpc

using namespace cpc;

static int fill_attn(AttnArgs& a, const CpcAttn* d) {
  if (d == nullptr || d->S <= 0 || d->S > 256 || d->dk <= 0 || d->dk > 64 || d->Hh <= 0 || d->Bt <= 0 || d->G <= 0)
    return CPC_ERR_ARG;
  if (d->qkv == nullptr || d->out == nullptr || d->probs == nullptr) return CPC_ERR_ARG;
  a.qkv = static_cast<const __nv_bfloat16*>(d->qkv); a.gs_qkv = d->gs_qkv;
  a.krel = static_cast<const float*>(d->krel);
  a.out = static_cast<__nv_bfloat16*>(d->out); a.gs_out = d->gs_out;
  a.out_f32 = static_cast<float*>(d->out_f32);
  a.rowsum = static_cast<float*>(d->rowsum);
  a.probs = static_cast<__nv_bfloat16*>(d->probs);
  a.S = d->S; a.dk = d->dk; a.d = d->dk * d->Hh; a.Bt = d->Bt; a.Hh = d->Hh;
  a.scale = d->scale;
  a.rng = static_cast<const unsigned long long*>(d->rng); a.site = d->rng_site; a.drop_p = d->drop_p;
  a.dout = static_cast<const __nv_bfloat16*>(d->dout); a.dqkv = static_cast<__nv_bfloat16*>(d->dqkv);
  a.dscore = static_cast<__nv_bfloat16*>(d->dscore); a.dkrel = static_cast<float*>(d->dkrel);
  return CPC_OK;
}

template <typename KernT>
static int set_smem(KernT kern, size_t bytes) {
  if (bytes > 48 * 1024)
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) != cudaSuccess)
      return CPC_ERR_LAUNCH;
  return CPC_OK;
}

extern "C" int cpc_attn_fwd(const CpcAttn* d, void* stream_) {
  AttnArgs a;
  int rc = fill_attn(a, d);
  if (rc != CPC_OK) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (attn_tc_supported(a)) return attn_fwd_tc(a, d->G, stream);
  dim3 grid(a.Hh, a.Bt, d->G);
  size_t sm = attn_smem<float>(a.S, a.dk, false);
  if (sm <= 100 * 1024) {
    if ((rc = set_smem(attn_fwd_kernel<float>, sm)) != CPC_OK) return rc;
    launch_k(attn_fwd_kernel<float>, dim3(grid), dim3(256), sm, stream, a);
  } else {
    sm = attn_smem<__nv_bfloat16>(a.S, a.dk, false);
    if ((rc = set_smem(attn_fwd_kernel<__nv_bfloat16>, sm)) != CPC_OK) return rc;
    launch_k(attn_fwd_kernel<__nv_bfloat16>, dim3(grid), dim3(256), sm, stream, a);
  }
  return launch_status();
}

extern "C" int cpc_attn_bwd(const CpcAttn* d, void* stream_) {
  AttnArgs a;
  int rc = fill_attn(a, d);
  if (rc != CPC_OK) return rc;
  if (a.dout == nullptr || a.dqkv == nullptr) return CPC_ERR_ARG;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (attn_tc_supported(a)) return attn_bwd_tc(a, d->G, stream);
  if (a.dscore == nullptr) return CPC_ERR_ARG;
  dim3 grid(a.Hh, a.Bt, d->G);
  size_t sm = attn_smem<float>(a.S, a.dk, true);
  if (sm <= 100 * 1024) {
    if ((rc = set_smem(attn_bwd_kernel<float>, sm)) != CPC_OK) return rc;
    launch_k(attn_bwd_kernel<float>, dim3(grid), dim3(256), sm, stream, a);
  } else {
    sm = attn_smem<__nv_bfloat16>(a.S, a.dk, true);
    if (sm > 220 * 1024) return CPC_ERR_UNSUPPORTED;
    if ((rc = set_smem(attn_bwd_kernel<__nv_bfloat16>, sm)) != CPC_OK) return rc;
    launch_k(attn_bwd_kernel<__nv_bfloat16>, dim3(grid), dim3(256), sm, stream, a);
  }
  return launch_status();
}

// ------------------------------------------------------------------------------------------------------------------
// cpc_attn_bwd_noalloc: the scratch-using (general, CUDA-core) backward above with its `dscore` scratch in a file-scope
// device array instead of a caller-allocated buffer -- for callers that cannot allocate (code under CUDA-graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Nothing here (or below it) calls cudaMalloc / cudaFree or an allocating library routine.
//   * CPC_ATTN_NOALLOC_MAX_ELEMS bf16 elements (8 MiB) are held by every GPU that loads the library, until exit;
//   * every call SHARES the array: calls that overlap on the GPU (two streams, two graph branches) race on it;
//   * it keeps the previous call's contents.  Nothing has to start at zero: phase A of the kernel writes every
//     lower-triangle element (i >= j) of every (group, window, head) block before phases B / C read exactly those.
// In host code `g_attn_dscore_static` is a host-side stand-in for the symbol, not a device address: the address the
// kernel gets comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) __nv_bfloat16 g_attn_dscore_static[CPC_ATTN_NOALLOC_MAX_ELEMS];

static int attn_static_scratch(__nv_bfloat16** out) {
  constexpr int MAX_DEV = 64;
  static __nv_bfloat16* addr[MAX_DEV] = {};  // per device (benign race: every thread resolves the same address)
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEV) return CPC_ERR_LAUNCH;
  if (addr[dev] == nullptr) {
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, g_attn_dscore_static) != cudaSuccess || p == nullptr) {
      (void)cudaGetLastError();
      return CPC_ERR_LAUNCH;
    }
    addr[dev] = static_cast<__nv_bfloat16*>(p);
  }
  *out = addr[dev];
  return CPC_OK;
}

extern "C" int cpc_attn_bwd_noalloc(const CpcAttn* d, void* stream_) {
  AttnArgs a;
  int rc = fill_attn(a, d);
  if (rc != CPC_OK) return rc;
  if (a.dout == nullptr || a.dqkv == nullptr) return CPC_ERR_ARG;
  const long long elems = (long long)d->G * d->Bt * d->Hh * d->S * d->S;
  if (elems > (long long)CPC_ATTN_NOALLOC_MAX_ELEMS) return CPC_ERR_TOO_LARGE;
  if ((rc = attn_static_scratch(&a.dscore)) != CPC_OK) return rc;  // whatever the caller put in d->dscore is ignored
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  dim3 grid(a.Hh, a.Bt, d->G);
  size_t sm = attn_smem<float>(a.S, a.dk, true);
  if (sm <= 100 * 1024) {
    if ((rc = set_smem(attn_bwd_kernel<float>, sm)) != CPC_OK) return rc;
    launch_k(attn_bwd_kernel<float>, dim3(grid), dim3(256), sm, stream, a);
  } else {
    sm = attn_smem<__nv_bfloat16>(a.S, a.dk, true);
    if (sm > 220 * 1024) return CPC_ERR_UNSUPPORTED;
    if ((rc = set_smem(attn_bwd_kernel<__nv_bfloat16>, sm)) != CPC_OK) return rc;
    launch_k(attn_bwd_kernel<__nv_bfloat16>, dim3(grid), dim3(256), sm, stream, a);
  }
  return launch_status();
}
