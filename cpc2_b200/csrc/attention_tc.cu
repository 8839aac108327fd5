// Tensor-core path of the causal relative-position attention (cpc/transformers.py:52-70), S <= 256, dk in {32, 64}:
// the shapes of the default CPC step (12 prediction heads x 8 windows x 8 attention heads, S = 116, dk = 32), of
// the transformer AR (S = 128) and of the long-window sizing (S = 244 / 256, dk = 64).  One CTA per (head group,
// window, attention head), four warps; the sequence is padded to SP in {128, 256} and cut in 16-row blocks that
// are dealt to the warps in causal-balanced pairs {4q + w, SP/16 - 1 - 4q - w}; all products run on
// mma.sync.m16n8k16 bf16 with operands fetched by ldmatrix from padded (conflict-free) shared-memory tiles that
// cp.async fills while the fp32 Krel slice is converted and the row statistics D are computed.
//
//   forward   R = Q.Krel (only the r >= S-16-i0 columns a block can address), staged per warp in shared memory;
//             scores = (Q.K^T + skew(R)) * scale with skew(R)[i,j] = R[i, S-1-i+j]  -- the reference's
//             cat/view "skew" is an index shift here; softmax in registers (quad shuffles); P (bf16) is kept
//             for the backward; O = dropout(P).V with the score accumulators re-used as A fragments.
//   backward  row pass (query blocks): dP = dO.V^T, dS = P o (dP - D) * scale, dQ = dS.K + unskew(dS).Krel^T,
//             dKrel += Q^T.unskew(dS);
//             column pass (key blocks): the same [i][j] tiles are recomputed, P~ and dS are transposed in
//             registers (movmatrix) into the A fragments of dV = P~^T.dO and dK = dS^T.Q.
//             Every P element a pass needs is fetched into registers BEFORE its loop (one exposed global
//             latency per pass instead of one per 16-key step: the kernel was latency-bound, ncu r01), and
//             nothing but P ever round-trips through memory.
// Dropout masks are regenerated from a counter hash of the element index in both directions.
// Larger S / odd head sizes fall back to the CUDA-core kernels of attention.cu.
#include <cstdlib>

#include "cpc_internal.h"
#include "ptx.cuh"
#include "attn_args.h"

namespace cpc {

namespace {

// warps per CTA: 8 at SP = 128 (one row block + one column block each, 2 CTAs/SM -> 16 warps/SM: the kernels are
// bound by dependent-issue latency, not by the tensor pipe), 4 at SP = 256 (shared memory for the per-warp stage)
__host__ __device__ constexpr int att_warps(int SP) { return SP == 128 ? 8 : 4; }

struct DropCfg {
  uint32_t seed_lo, seed_hi, thresh;
  float scale;
};
__device__ __forceinline__ DropCfg make_drop(const AttnArgs& a) {
  DropCfg d{0u, 0u, 0u, 1.f};
  if (a.rng != nullptr && a.drop_p > 0.f) {
    const unsigned long long seed = a.rng[0] + a.rng[1] * 0x632BE59BD9B4E019ull;
    d.seed_lo = static_cast<uint32_t>(seed);
    d.seed_hi = static_cast<uint32_t>(seed >> 32) + a.site * 0x85EBCA77u;
    d.thresh = static_cast<uint32_t>(fminf(a.drop_p, 0.9999f) * 65536.0f);  // 16-bit draws, two per hash
    d.scale = 1.f / (1.f - a.drop_p);
  }
  return d;
}
// keep flags of the element pair (row, j), (row, j + 1), j even: one hash, two 16-bit draws.  `row` is the global
// row index ((group, window, head), i); the pair index is unique per (row, j / 2) for any S <= 512.
__device__ __forceinline__ void keep_pair(const DropCfg& d, long long row, int j, bool& k0, bool& k1) {
  const uint32_t h = rng_pair_u16(d.seed_lo, d.seed_hi, static_cast<unsigned long long>(row) * 256ull + (j >> 1));
  k0 = (h & 0xFFFFu) >= d.thresh;
  k1 = (h >> 16) >= d.thresh;
}

// rows [0, S) of a [S][dk] slice (row stride `ld` elements) -> shared [SP][PD] bf16 by cp.async, rows >= S zero.
// The caller commits / waits.
template <int DK, int SP, int ATT_THREADS>
__device__ __forceinline__ void load_rows_async(__nv_bfloat16* dst, const __nv_bfloat16* src, long long ld, int S) {
  constexpr int PD = DK + 8, CH = DK / 8;
#pragma unroll 4
  for (int i = threadIdx.x; i < SP * CH; i += ATT_THREADS) {
    const int s = i / CH, ch = i % CH;
    if (s < S) cp_async_16(dst + s * PD + ch * 8, src + (long long)s * ld + ch * 8);
    else *reinterpret_cast<uint4*>(dst + s * PD + ch * 8) = make_uint4(0u, 0u, 0u, 0u);
  }
}

// the 16-row block warp `w` owns in pass `p` (NB blocks, two per (warp, q): one short, one long causal row)
template <int NB, int NW>
__device__ __forceinline__ int block_of(int warp, int p) {
  const int q = p >> 1;
  return (p & 1) ? NB - 1 - NW * q - warp : NW * q + warp;
}

// P[e], P[e+1] (bf16, probabilities before dropout) as one packed register; `second` = element e+1 is inside
// the causal triangle.  The paired load needs an even S (4-byte alignment of every row start).
__device__ __forceinline__ uint32_t load_p_pair(const __nv_bfloat16* probs, long long e, bool pair_ok, bool second) {
  if (pair_ok && second) return *reinterpret_cast<const uint32_t*>(probs + e);
  uint32_t lo = *reinterpret_cast<const unsigned short*>(probs + e);
  if (second) lo |= static_cast<uint32_t>(*reinterpret_cast<const unsigned short*>(probs + e + 1)) << 16;
  return lo;
}
__device__ __forceinline__ float bf16lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16hi(uint32_t v) { return __uint_as_float(v & 0xFFFF0000u); }

// transpose of an 8x8 b16 matrix held one row per quad (the C / A fragment layout of mma.m16n8k16)
__device__ __forceinline__ uint32_t movmatrix_t(uint32_t v) {
  uint32_t r;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(r) : "r"(v));
  return r;
}

}  // namespace

// ================================================================================================ forward
template <int DK, int SP, int NW>
__global__ void __launch_bounds__(NW * 32) attn_fwd_tc_kernel(const AttnArgs a) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  constexpr int ATT_THREADS = NW * 32, ATT_WARPS = NW;
  constexpr int PD = DK + 8, KS = DK / 16, NT = DK / 8, RP = SP + 8, NB = SP / 16, NJ = SP / 8;
  extern __shared__ __align__(16) uint8_t smem_dyn[];
  __nv_bfloat16* Qs = reinterpret_cast<__nv_bfloat16*>(smem_dyn);
  __nv_bfloat16* Ks = Qs + SP * PD;
  __nv_bfloat16* Vs = Ks + SP * PD;
  __nv_bfloat16* KrT = Vs + SP * PD;                          // Krel^T [r][c]
  float* Rs_all = reinterpret_cast<float*>(KrT + SP * PD);    // per warp [16][RP]
  const int S = a.S, d = a.d;
  const int h = blockIdx.x, b = blockIdx.y, grp = blockIdx.z;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const bool rel = a.krel != nullptr;
  const __nv_bfloat16* base = a.qkv + (long long)grp * a.gs_qkv + (long long)b * S * 3 * d + h * DK;
  load_rows_async<DK, SP, ATT_THREADS>(Qs, base, 3 * d, S);
  load_rows_async<DK, SP, ATT_THREADS>(Ks, base + d, 3 * d, S);
  load_rows_async<DK, SP, ATT_THREADS>(Vs, base + 2 * d, 3 * d, S);
  cp_async_commit();
  if (rel) {
    const float* kr = a.krel + (long long)grp * DK * S;
#pragma unroll 8
    for (int i = threadIdx.x; i < DK * SP; i += ATT_THREADS) {
      const int c = i / SP, r = i % SP;
      KrT[r * PD + c] = __float2bfloat16(r < S ? kr[c * S + r] : 0.f);
    }
  }
  cp_async_wait<0>();
  __syncthreads();
  const DropCfg drop = make_drop(a);
  const long long rbase = (((long long)grp * a.Bt + b) * a.Hh + h) * S;  // global row index of query 0
  const long long pbase = rbase * S;
  float* Rs = Rs_all + warp * 16 * RP;
  const int ntile_s = (S + 7) >> 3;
  const bool pair_ok = (S & 1) == 0;

  for (int pass = 0; pass < NB / ATT_WARPS; ++pass) {
    const int rb = block_of<NB, ATT_WARPS>(warp, pass);
    const int i0 = rb * 16;
    if (i0 >= S) continue;
    uint32_t aq[KS][4];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) ldsm_x4(aq[ks], smem_u32(Qs + (i0 + (lane & 15)) * PD + ks * 16 + (lane >> 4) * 8));
    if (rel) {
      __syncwarp();
      for (int r0 = max(0, S - 16 - i0) & ~7; r0 < S; r0 += 8) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          uint32_t bb[2];
          ldsm_x2(bb, smem_u32(KrT + (r0 + (lane & 7)) * PD + ks * 16 + ((lane >> 3) & 1) * 8));
          mma_bf16_16816(acc, aq[ks], bb[0], bb[1]);
        }
        *reinterpret_cast<float2*>(Rs + g * RP + r0 + 2 * t) = make_float2(acc[0], acc[1]);
        *reinterpret_cast<float2*>(Rs + (g + 8) * RP + r0 + 2 * t) = make_float2(acc[2], acc[3]);
      }
      __syncwarp();
    }
    const int njt = min(2 * rb + 2, ntile_s);
    float sc[NJ][4];
#pragma unroll
    for (int jt = 0; jt < NJ; ++jt) {
      if (jt < njt) {
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
          uint32_t bb[2];
          ldsm_x2(bb, smem_u32(Ks + (jt * 8 + (lane & 7)) * PD + ks * 16 + ((lane >> 3) & 1) * 8));
          mma_bf16_16816(acc, aq[ks], bb[0], bb[1]);
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int ii = g + 8 * (e >> 1), i = i0 + ii, j = jt * 8 + 2 * t + (e & 1);
          const bool valid = j <= i && i < S;
          float s = acc[e];
          if (rel && valid) s += Rs[ii * RP + S - 1 - i + j];
          sc[jt][e] = valid ? s * a.scale : -INFINITY;
        }
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) sc[jt][e] = -INFINITY;
      }
    }
    // softmax over each row (a row lives in the 4 lanes of a quad)
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int jt = 0; jt < NJ; ++jt)
#pragma unroll
      for (int e = 0; e < 4; ++e) mx[e >> 1] = fmaxf(mx[e >> 1], sc[jt][e]);
    float sum[2] = {0.f, 0.f};
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      if (mx[r] == -INFINITY) mx[r] = 0.f;  // padding row
    }
#pragma unroll
    for (int jt = 0; jt < NJ; ++jt)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float p = __expf(sc[jt][e] - mx[e >> 1]);
        sc[jt][e] = p;
        sum[e >> 1] += p;
      }
    float inv[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      sum[r] += __shfl_xor_sync(0xffffffffu, sum[r], 1);
      sum[r] += __shfl_xor_sync(0xffffffffu, sum[r], 2);
      inv[r] = sum[r] > 0.f ? 1.f / sum[r] : 0.f;
    }
    // probabilities out (bf16, lower triangle).  P.V is formed from the SAME bf16-rounded values (the dropout scale
    // is applied to the accumulator afterwards, so the operand stays exactly the stored P^): the backward's row
    // statistic D_i = dO_i . O_i / sum_j P^_ij is then algebraically sum_j P^_ij dP~_ij / sum_j P^_ij for the very P^
    // and V it multiplies -- rows of dS sum to zero to fp32 rounding (CpcAttn.out_f32 / rowsum).
    float rs[2] = {0.f, 0.f};
#pragma unroll
    for (int jt = 0; jt < NJ; ++jt) {
      if (jt < njt) {
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const int i = i0 + g + 8 * r, j = jt * 8 + 2 * t;
          float p0 = sc[jt][2 * r] * inv[r], p1 = sc[jt][2 * r + 1] * inv[r];
          if (i < S && j <= i) {
            const long long e = pbase + (long long)i * S + j;
            const __nv_bfloat162 pb = __floats2bfloat162_rn(p0, p1);
            p0 = __low2float(pb);
            p1 = j + 1 <= i ? __high2float(pb) : 0.f;
            rs[r] += p0 + p1;
            if (a.probs != nullptr) {
              if (pair_ok && j + 1 <= i) *reinterpret_cast<__nv_bfloat162*>(a.probs + e) = pb;
              else {
                a.probs[e] = __low2bfloat16(pb);
                if (j + 1 <= i) a.probs[e + 1] = __high2bfloat16(pb);
              }
            }
            if (drop.thresh != 0) {
              bool k0, k1;
              keep_pair(drop, rbase + i, j, k0, k1);
              p0 = k0 ? p0 : 0.f;
              p1 = k1 ? p1 : 0.f;
            }
          } else {
            p0 = p1 = 0.f;
          }
          sc[jt][2 * r] = p0;
          sc[jt][2 * r + 1] = p1;
        }
      }
    }
    if (a.rowsum != nullptr) {
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        rs[r] += __shfl_xor_sync(0xffffffffu, rs[r], 1);
        rs[r] += __shfl_xor_sync(0xffffffffu, rs[r], 2);
        if (t == 0 && i0 + g + 8 * r < S) a.rowsum[rbase + i0 + g + 8 * r] = rs[r];
      }
    }
    // O = dropout_scale * (keep o P^) . V : two adjacent score tiles form the A fragment of one 16-key step
    float o[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) o[nt][e] = 0.f;
#pragma unroll
    for (int kk = 0; kk < NB; ++kk) {
      if (2 * kk < njt) {
        uint32_t ap[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)  // (exact: the values are bf16 already)
          ap[q] = pack_bf16x2(sc[2 * kk + (q >> 1)][2 * (q & 1)], sc[2 * kk + (q >> 1)][2 * (q & 1) + 1]);
#pragma unroll
        for (int n2 = 0; n2 < NT / 2; ++n2) {
          uint32_t bb[4];
          ldsm_x4_t(bb, smem_u32(Vs + (kk * 16 + (lane & 15)) * PD + n2 * 16 + (lane >> 4) * 8));
          mma_bf16_16816(o[2 * n2], ap, bb[0], bb[1]);
          mma_bf16_16816(o[2 * n2 + 1], ap, bb[2], bb[3]);
        }
      }
    }
    if (drop.thresh != 0) {
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) o[nt][e] *= drop.scale;
    }
    __nv_bfloat16* ob = a.out + (long long)grp * a.gs_out + (long long)b * S * d + h * DK;
    float* of = a.out_f32 != nullptr ? a.out_f32 + (long long)grp * a.gs_out + (long long)b * S * d + h * DK : nullptr;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int c = nt * 8 + 2 * t;
      if (i0 + g < S) {
        *reinterpret_cast<__nv_bfloat162*>(ob + (long long)(i0 + g) * d + c) = __floats2bfloat162_rn(o[nt][0], o[nt][1]);
        if (of != nullptr) *reinterpret_cast<float2*>(of + (long long)(i0 + g) * d + c) = make_float2(o[nt][0], o[nt][1]);
      }
      if (i0 + g + 8 < S) {
        *reinterpret_cast<__nv_bfloat162*>(ob + (long long)(i0 + g + 8) * d + c) = __floats2bfloat162_rn(o[nt][2], o[nt][3]);
        if (of != nullptr) *reinterpret_cast<float2*>(of + (long long)(i0 + g + 8) * d + c) = make_float2(o[nt][2], o[nt][3]);
      }
    }
  }
}

// ================================================================================================ backward
template <int DK, int SP>
__global__ void __launch_bounds__(att_warps(SP) * 32) attn_bwd_tc_kernel(const AttnArgs a) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  constexpr int ATT_WARPS = att_warps(SP), ATT_THREADS = ATT_WARPS * 32;
  constexpr int PD = DK + 8, KS = DK / 16, NT = DK / 8, PS = SP + 8, NB = SP / 16, NJ = SP / 8;
  constexpr bool FULL_U = NB == att_warps(SP);  // one row block per warp: the stages hold all of unskew(dS)
  extern __shared__ __align__(16) uint8_t smem_dyn[];
  __nv_bfloat16* Qs = reinterpret_cast<__nv_bfloat16*>(smem_dyn);
  __nv_bfloat16* Ks = Qs + SP * PD;
  __nv_bfloat16* Vs = Ks + SP * PD;
  __nv_bfloat16* dOs = Vs + SP * PD;
  __nv_bfloat16* Krs = dOs + SP * PD;                                      // Krel [c][r] bf16, pitch PS
  __nv_bfloat16* stage_all = Krs + DK * PS;                                // per warp unskew(dS) [16][PS]
  float* Dsm = reinterpret_cast<float*>(stage_all + ATT_WARPS * 16 * PS);  // [SP]
  const int S = a.S, d = a.d;
  const int h = blockIdx.x, b = blockIdx.y, grp = blockIdx.z;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const bool rel = a.krel != nullptr;
  const __nv_bfloat16* base = a.qkv + (long long)grp * a.gs_qkv + (long long)b * S * 3 * d + h * DK;
  const long long obase = (long long)grp * a.gs_out + (long long)b * S * d + h * DK;
  load_rows_async<DK, SP, ATT_THREADS>(Qs, base, 3 * d, S);
  load_rows_async<DK, SP, ATT_THREADS>(Ks, base + d, 3 * d, S);
  load_rows_async<DK, SP, ATT_THREADS>(Vs, base + 2 * d, 3 * d, S);
  load_rows_async<DK, SP, ATT_THREADS>(dOs, a.dout + obase, d, S);
  cp_async_commit();
  // D_i = sum_j P_ij dP~_ij / sum_j P_ij is formed in the row pass from the SAME bf16 P and the SAME dP (identical
  // MMA sequence) that dS = P o (dP~ - D) uses, so every row of dS sums to zero to fp32 rounding.  The flash-attention
  // shortcut D_i = dO_i . O_i is only as good as O: in trained models K and V carry a large component common to all
  // keys, dP_ij is nearly constant along a row, and the 2^-17 mismatch between O and the recomputed dP leaked,
  // multiplied by that common component, into dQ (cosine 0.4-0.95 on Wq gradients away from init).
  const bool fast_d = a.out_f32 != nullptr && a.rowsum != nullptr;  // (a) of CpcAttn: D_i = dO_i . O_i / rowsum_i
  const long long rbase0 = (((long long)grp * a.Bt + b) * a.Hh + h) * S;
  for (int i = threadIdx.x; i < SP; i += ATT_THREADS) {
    float acc = 0.f;
    if (fast_d && i < S) {  // straight from global memory while the tiles are in flight
      const float* orow = a.out_f32 + obase + (long long)i * d;
      const __nv_bfloat16* drow = a.dout + obase + (long long)i * d;
      float4 ov[DK / 4];
      uint4 dv[DK / 8];
#pragma unroll
      for (int ch = 0; ch < DK / 4; ++ch) ov[ch] = *reinterpret_cast<const float4*>(orow + ch * 4);
#pragma unroll
      for (int ch = 0; ch < DK / 8; ++ch) dv[ch] = *reinterpret_cast<const uint4*>(drow + ch * 8);
      const float rsum = a.rowsum[rbase0 + i];
#pragma unroll
      for (int ch = 0; ch < DK / 8; ++ch) {
        const uint32_t* d2 = reinterpret_cast<const uint32_t*>(&dv[ch]);
        acc += ov[2 * ch].x * bf16lo(d2[0]) + ov[2 * ch].y * bf16hi(d2[0]) + ov[2 * ch].z * bf16lo(d2[1]) + ov[2 * ch].w * bf16hi(d2[1]);
        acc += ov[2 * ch + 1].x * bf16lo(d2[2]) + ov[2 * ch + 1].y * bf16hi(d2[2]) + ov[2 * ch + 1].z * bf16lo(d2[3]) +
               ov[2 * ch + 1].w * bf16hi(d2[3]);
      }
      acc = rsum > 0.f ? acc / rsum : 0.f;
    }
    Dsm[i] = acc;
  }
  if (rel) {
    const float* kr = a.krel + (long long)grp * DK * S;
#pragma unroll 8
    for (int i = threadIdx.x; i < DK * SP; i += ATT_THREADS) {
      const int c = i / SP, r = i % SP;
      Krs[c * PS + r] = __float2bfloat16(r < S ? kr[c * S + r] : 0.f);
    }
  }
  cp_async_wait<0>();
  __syncthreads();
  const DropCfg drop = make_drop(a);
  const long long rbase = (((long long)grp * a.Bt + b) * a.Hh + h) * S;  // global row index of query 0
  const long long pbase = rbase * S;
  const int ntile_s = (S + 7) >> 3;
  const int nblk_s = (S + 15) >> 4;
  const bool pair_ok = (S & 1) == 0;
  __nv_bfloat16* stage = stage_all + warp * 16 * PS;
  __nv_bfloat16* dq_out = a.dqkv + (long long)grp * a.gs_qkv + (long long)b * S * 3 * d + h * DK;
  float* dkr_out = a.dkrel != nullptr ? a.dkrel + (long long)grp * DK * S : nullptr;

  // ------------------------------------------------------------------ row pass: dQ, dKrel
  for (int pass = 0; pass < NB / ATT_WARPS; ++pass) {
    const int rb = block_of<NB, ATT_WARPS>(warp, pass);
    const int i0 = rb * 16;
    if (i0 >= S) continue;
    const int njt = min(2 * rb + 2, ntile_s);
    // every P pair of this row block, in flight together
    uint32_t pv[NJ][2];
#pragma unroll
    for (int jt = 0; jt < NJ; ++jt)
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int i = i0 + g + 8 * r, j = jt * 8 + 2 * t;
        pv[jt][r] = 0u;
        if (jt < njt && i < S && j <= i) pv[jt][r] = load_p_pair(a.probs, pbase + (long long)i * S + j, pair_ok, j + 1 <= i);
      }
    uint32_t ado[KS][4];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) ldsm_x4(ado[ks], smem_u32(dOs + (i0 + (lane & 15)) * PD + ks * 16 + (lane >> 4) * 8));
    if (rel) {
      __syncwarp();
      for (int i = lane; i < 16 * PS / 8; i += 32) reinterpret_cast<uint4*>(stage)[i] = make_uint4(0u, 0u, 0u, 0u);
      __syncwarp();
    }
    float dq[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) dq[nt][e] = 0.f;
    float Di[2] = {Dsm[i0 + g], Dsm[i0 + g + 8]};
    if (!fast_d) {  // (b) of CpcAttn: one more dP sweep, D from the same P^ and dP~ the row's dS uses
      float dsum[2] = {0.f, 0.f}, psum[2] = {0.f, 0.f};
#pragma unroll
      for (int jt = 0; jt < NJ; ++jt) {
        if (jt < njt) {
          float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int ks = 0; ks < KS; ++ks) {
            uint32_t bb[2];
            ldsm_x2(bb, smem_u32(Vs + (jt * 8 + (lane & 7)) * PD + ks * 16 + ((lane >> 3) & 1) * 8));
            mma_bf16_16816(acc, ado[ks], bb[0], bb[1]);
          }
#pragma unroll
          for (int r = 0; r < 2; ++r) {
            const int i = i0 + g + 8 * r, j = jt * 8 + 2 * t;
            if (i < S && j <= i) {
              const float p0 = bf16lo(pv[jt][r]), p1 = bf16hi(pv[jt][r]);  // p1 = 0 outside the triangle
              float dp0 = acc[2 * r], dp1 = j + 1 <= i ? acc[2 * r + 1] : 0.f;
              if (drop.thresh != 0) {
                bool k0, k1;
                keep_pair(drop, rbase + i, j, k0, k1);
                dp0 = k0 ? dp0 * drop.scale : 0.f;
                dp1 = k1 ? dp1 * drop.scale : 0.f;
              }
              dsum[r] += p0 * dp0 + p1 * dp1;
              psum[r] += p0 + p1;
            }
          }
        }
      }
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        dsum[r] += __shfl_xor_sync(0xffffffffu, dsum[r], 1);
        dsum[r] += __shfl_xor_sync(0xffffffffu, dsum[r], 2);
        psum[r] += __shfl_xor_sync(0xffffffffu, psum[r], 1);
        psum[r] += __shfl_xor_sync(0xffffffffu, psum[r], 2);
        Di[r] = psum[r] > 0.f ? dsum[r] / psum[r] : 0.f;
        if (t == 0) Dsm[i0 + g + 8 * r] = Di[r];  // for the column pass
      }
    }
#pragma unroll
    for (int kk = 0; kk < NB; ++kk) {
      if (2 * kk < njt) {
        float ds[2][4];
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          const int jt = 2 * kk + hf;
#pragma unroll
          for (int e = 0; e < 4; ++e) ds[hf][e] = 0.f;
          if (jt < njt) {
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int ks = 0; ks < KS; ++ks) {
              uint32_t bb[2];
              ldsm_x2(bb, smem_u32(Vs + (jt * 8 + (lane & 7)) * PD + ks * 16 + ((lane >> 3) & 1) * 8));
              mma_bf16_16816(acc, ado[ks], bb[0], bb[1]);
            }
#pragma unroll
            for (int r = 0; r < 2; ++r) {
              const int ii = g + 8 * r, i = i0 + ii, j = jt * 8 + 2 * t;
              if (i < S && j <= i) {
                const float p0 = bf16lo(pv[jt][r]), p1 = bf16hi(pv[jt][r]);
                float dp0 = acc[2 * r], dp1 = acc[2 * r + 1];
                if (drop.thresh != 0) {
                  bool k0, k1;
                  keep_pair(drop, rbase + i, j, k0, k1);
                  dp0 = k0 ? dp0 * drop.scale : 0.f;
                  dp1 = k1 ? dp1 * drop.scale : 0.f;
                }
                const float s0 = p0 * (dp0 - Di[r]) * a.scale;
                const float s1 = j + 1 <= i ? p1 * (dp1 - Di[r]) * a.scale : 0.f;
                ds[hf][2 * r] = s0;
                ds[hf][2 * r + 1] = s1;
                if (rel) {
                  const int rr = S - 1 - i + j;
                  stage[ii * PS + rr] = __float2bfloat16(s0);
                  if (j + 1 <= i) stage[ii * PS + rr + 1] = __float2bfloat16(s1);
                }
              }
            }
          }
        }
        uint32_t ap[4] = {pack_bf16x2(ds[0][0], ds[0][1]), pack_bf16x2(ds[0][2], ds[0][3]),
                          pack_bf16x2(ds[1][0], ds[1][1]), pack_bf16x2(ds[1][2], ds[1][3])};
#pragma unroll
        for (int n2 = 0; n2 < NT / 2; ++n2) {
          uint32_t bb[4];
          ldsm_x4_t(bb, smem_u32(Ks + (kk * 16 + (lane & 15)) * PD + n2 * 16 + (lane >> 4) * 8));
          mma_bf16_16816(dq[2 * n2], ap, bb[0], bb[1]);
          mma_bf16_16816(dq[2 * n2 + 1], ap, bb[2], bb[3]);
        }
      }
    }
    if (rel) {
      __syncwarp();
      const int r_lo = max(0, S - 16 - i0);
      // dQ += unskew(dS) . Krel^T   (reduction over r)
      for (int r0 = r_lo & ~15; r0 < S; r0 += 16) {
        uint32_t as[4];
        ldsm_x4(as, smem_u32(stage + (lane & 15) * PS + r0 + (lane >> 4) * 8));
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          uint32_t bb[2];
          ldsm_x2(bb, smem_u32(Krs + (nt * 8 + (lane & 7)) * PS + r0 + ((lane >> 3) & 1) * 8));
          mma_bf16_16816(dq[nt], as, bb[0], bb[1]);
        }
      }
      // dKrel[c][r] += sum_ii Q[i0+ii][c] * unskew(dS)[ii][r]: with one row block per warp (FULL_U) the stages of
      // the CTA form the whole unskew(dS) matrix and the product is taken once after the passes; otherwise per
      // row block, reduced straight into global memory
      if (!FULL_U && dkr_out != nullptr) {
        uint32_t aqt[KS][4];
#pragma unroll
        for (int mt = 0; mt < KS; ++mt)
          ldsm_x4_t(aqt[mt], smem_u32(Qs + (i0 + (lane & 7) + 8 * (lane >> 4)) * PD + mt * 16 + 8 * ((lane >> 3) & 1)));
        for (int r0 = r_lo & ~7; r0 < S; r0 += 8) {
          uint32_t bb[2];
          ldsm_x2_t(bb, smem_u32(stage + (lane & 15) * PS + r0));
#pragma unroll
          for (int mt = 0; mt < KS; ++mt) {
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            mma_bf16_16816(acc, aqt[mt], bb[0], bb[1]);
            const int c = mt * 16 + g, r = r0 + 2 * t;
            if (r < S) {
              red_add_f32(dkr_out + c * S + r, acc[0]);
              red_add_f32(dkr_out + (c + 8) * S + r, acc[2]);
            }
            if (r + 1 < S) {
              red_add_f32(dkr_out + c * S + r + 1, acc[1]);
              red_add_f32(dkr_out + (c + 8) * S + r + 1, acc[3]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int c = nt * 8 + 2 * t;
      if (i0 + g < S)
        *reinterpret_cast<__nv_bfloat162*>(dq_out + (long long)(i0 + g) * 3 * d + c) = __floats2bfloat162_rn(dq[nt][0], dq[nt][1]);
      if (i0 + g + 8 < S)
        *reinterpret_cast<__nv_bfloat162*>(dq_out + (long long)(i0 + g + 8) * 3 * d + c) = __floats2bfloat162_rn(dq[nt][2], dq[nt][3]);
    }
  }

  __syncthreads();  // every row block's D is in Dsm
  // ------------------------------------------------------------------ column pass: dK, dV
  for (int pass = 0; pass < NB / ATT_WARPS; ++pass) {
    // multi-pass: long columns first; single pass: column block w costs NB - w, row block w cost w + 1 (balanced)
    const int jb = block_of<NB, ATT_WARPS>(warp, NB / ATT_WARPS == 1 ? pass : pass ^ 1);
    const int j0 = jb * 16;
    if (j0 >= S) continue;
    // P pairs of the strip [i >= j0][j0 .. j0+15]: (query block, 8-key half, row half)
    uint32_t pc[NB][2][2];
#pragma unroll
    for (int ib = 0; ib < NB; ++ib)
#pragma unroll
      for (int hf = 0; hf < 2; ++hf)
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const int i = ib * 16 + g + 8 * r, j = j0 + 8 * hf + 2 * t;
          pc[ib][hf][r] = 0u;
          if (ib >= jb && i < S && j <= i)
            pc[ib][hf][r] = load_p_pair(a.probs, pbase + (long long)i * S + j, pair_ok, j + 1 <= i);
        }
    uint32_t bv[2][KS][2];  // V rows of this key block as B fragments of dP = dO . V^T
#pragma unroll
    for (int hf = 0; hf < 2; ++hf)
#pragma unroll
      for (int ks = 0; ks < KS; ++ks)
        ldsm_x2(bv[hf][ks], smem_u32(Vs + (j0 + hf * 8 + (lane & 7)) * PD + ks * 16 + ((lane >> 3) & 1) * 8));
    float dk[NT][4], dv[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) dk[nt][e] = dv[nt][e] = 0.f;
#pragma unroll
    for (int ib = 0; ib < NB; ++ib) {
      if (ib >= jb && ib < nblk_s) {
        uint32_t ado[KS][4];
#pragma unroll
        for (int ks = 0; ks < KS; ++ks)
          ldsm_x4(ado[ks], smem_u32(dOs + (ib * 16 + (lane & 15)) * PD + ks * 16 + (lane >> 4) * 8));
        const float Di[2] = {Dsm[ib * 16 + g], Dsm[ib * 16 + g + 8]};
        uint32_t pk[2][2], sk[2][2];
#pragma unroll
        for (int hf = 0; hf < 2; ++hf) {
          float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int ks = 0; ks < KS; ++ks) mma_bf16_16816(acc, ado[ks], bv[hf][ks][0], bv[hf][ks][1]);
#pragma unroll
          for (int r = 0; r < 2; ++r) {
            const int i = ib * 16 + g + 8 * r, j = j0 + 8 * hf + 2 * t;
            float pd0 = 0.f, pd1 = 0.f, s0 = 0.f, s1 = 0.f;
            if (i < S && j <= i) {
              const float p0 = bf16lo(pc[ib][hf][r]), p1 = bf16hi(pc[ib][hf][r]);
              float dp0 = acc[2 * r], dp1 = acc[2 * r + 1];
              pd0 = p0;
              pd1 = p1;
              if (drop.thresh != 0) {
                bool k0, k1;
                keep_pair(drop, rbase + i, j, k0, k1);
                pd0 = k0 ? p0 * drop.scale : 0.f;
                pd1 = k1 ? p1 * drop.scale : 0.f;
                dp0 = k0 ? dp0 * drop.scale : 0.f;
                dp1 = k1 ? dp1 * drop.scale : 0.f;
              }
              s0 = p0 * (dp0 - Di[r]) * a.scale;
              s1 = p1 * (dp1 - Di[r]) * a.scale;  // p1 = 0 outside the triangle
            }
            pk[hf][r] = pack_bf16x2(pd0, pd1);
            sk[hf][r] = pack_bf16x2(s0, s1);
          }
        }
        // [i][j] tiles -> A fragments of the transposed [j][i] tile
        const uint32_t ap[4] = {movmatrix_t(pk[0][0]), movmatrix_t(pk[1][0]), movmatrix_t(pk[0][1]), movmatrix_t(pk[1][1])};
        const uint32_t as[4] = {movmatrix_t(sk[0][0]), movmatrix_t(sk[1][0]), movmatrix_t(sk[0][1]), movmatrix_t(sk[1][1])};
#pragma unroll
        for (int n2 = 0; n2 < NT / 2; ++n2) {
          uint32_t bb[4];
          ldsm_x4_t(bb, smem_u32(dOs + (ib * 16 + (lane & 15)) * PD + n2 * 16 + (lane >> 4) * 8));
          mma_bf16_16816(dv[2 * n2], ap, bb[0], bb[1]);
          mma_bf16_16816(dv[2 * n2 + 1], ap, bb[2], bb[3]);
          ldsm_x4_t(bb, smem_u32(Qs + (ib * 16 + (lane & 15)) * PD + n2 * 16 + (lane >> 4) * 8));
          mma_bf16_16816(dk[2 * n2], as, bb[0], bb[1]);
          mma_bf16_16816(dk[2 * n2 + 1], as, bb[2], bb[3]);
        }
      }
    }
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int c = nt * 8 + 2 * t;
      if (j0 + g < S) {
        __nv_bfloat16* o = dq_out + (long long)(j0 + g) * 3 * d + c;
        *reinterpret_cast<__nv_bfloat162*>(o + d) = __floats2bfloat162_rn(dk[nt][0], dk[nt][1]);
        *reinterpret_cast<__nv_bfloat162*>(o + 2 * d) = __floats2bfloat162_rn(dv[nt][0], dv[nt][1]);
      }
      if (j0 + g + 8 < S) {
        __nv_bfloat16* o = dq_out + (long long)(j0 + g + 8) * 3 * d + c;
        *reinterpret_cast<__nv_bfloat162*>(o + d) = __floats2bfloat162_rn(dk[nt][2], dk[nt][3]);
        *reinterpret_cast<__nv_bfloat162*>(o + 2 * d) = __floats2bfloat162_rn(dv[nt][2], dv[nt][3]);
      }
    }
  }
  if (FULL_U && rel && dkr_out != nullptr) {
    // dKrel[c][r] = sum_i Q[i][c] * U[i][r], U = unskew(dS) = the eight per-warp stages; warp w owns columns
    // [16w, 16w+16) and only the query blocks that can address them (U[i][r] = 0 for i < S-1-r)
    __syncthreads();
    const int r0 = warp * 16;
    if (r0 < S) {
      float acc[KS][2][4];
#pragma unroll
      for (int mt = 0; mt < KS; ++mt)
#pragma unroll
        for (int nn = 0; nn < 2; ++nn)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[mt][nn][e] = 0.f;
      for (int ib = max(0, S - 16 - r0) >> 4; ib < nblk_s; ++ib) {
        uint32_t aqt[KS][4];
#pragma unroll
        for (int mt = 0; mt < KS; ++mt)
          ldsm_x4_t(aqt[mt], smem_u32(Qs + (ib * 16 + (lane & 7) + 8 * (lane >> 4)) * PD + mt * 16 + 8 * ((lane >> 3) & 1)));
#pragma unroll
        for (int nn = 0; nn < 2; ++nn) {
          uint32_t bb[2];
          ldsm_x2_t(bb, smem_u32(stage_all + (ib * 16 + (lane & 15)) * PS + r0 + nn * 8));
#pragma unroll
          for (int mt = 0; mt < KS; ++mt) mma_bf16_16816(acc[mt][nn], aqt[mt], bb[0], bb[1]);
        }
      }
#pragma unroll
      for (int mt = 0; mt < KS; ++mt)
#pragma unroll
        for (int nn = 0; nn < 2; ++nn) {
          const int c = mt * 16 + g, r = r0 + nn * 8 + 2 * t;
          if (r < S) {
            red_add_f32(dkr_out + c * S + r, acc[mt][nn][0]);
            red_add_f32(dkr_out + (c + 8) * S + r, acc[mt][nn][2]);
          }
          if (r + 1 < S) {
            red_add_f32(dkr_out + c * S + r + 1, acc[mt][nn][1]);
            red_add_f32(dkr_out + (c + 8) * S + r + 1, acc[mt][nn][3]);
          }
        }
    }
  }
}

// ================================================================================================ host
constexpr int FWD_WARPS = 4;  // the forward holds a whole score row block in registers (~210): 4 warps, 2 CTAs/SM
template <int DK, int SP, int NW>
static size_t fwd_smem() { return (size_t)4 * SP * (DK + 8) * 2 + (size_t)NW * 16 * (SP + 8) * 4; }
template <int DK, int SP>
static size_t bwd_smem() {
  return (size_t)4 * SP * (DK + 8) * 2 + (size_t)DK * (SP + 8) * 2 + (size_t)att_warps(SP) * 16 * (SP + 8) * 2 + (size_t)SP * 4;
}

int attn_tc_supported(const AttnArgs& a) {
  static const bool enabled = [] {  // CPC_ATTN_TC=0 forces the general CUDA-core path (debugging / bisection)
    const char* e = getenv("CPC_ATTN_TC");
    return e == nullptr || e[0] != '0';
  }();
  return enabled && a.S <= 256 && a.S >= 1 && (a.dk == 32 || a.dk == 64) && a.d % 8 == 0;
}

template <int DK, int SP>
static int launch_fwd(const AttnArgs& a, int G, cudaStream_t stream) {
  constexpr int NW = FWD_WARPS;
  auto kern = attn_fwd_tc_kernel<DK, SP, NW>;
  static bool done = false;
  if (!done) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fwd_smem<DK, SP, NW>()) != cudaSuccess)
      return CPC_ERR_LAUNCH;
    done = true;
  }
  launch_k(kern, dim3(dim3(a.Hh, a.Bt, G)), dim3(NW * 32), fwd_smem<DK, SP, NW>(), stream, a);
  return launch_status();
}
template <int DK, int SP>
static int launch_bwd(const AttnArgs& a, int G, cudaStream_t stream) {
  auto kern = attn_bwd_tc_kernel<DK, SP>;
  static bool done = false;
  if (!done) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bwd_smem<DK, SP>()) != cudaSuccess)
      return CPC_ERR_LAUNCH;
    done = true;
  }
  launch_k(kern, dim3(dim3(a.Hh, a.Bt, G)), dim3(att_warps(SP) * 32), bwd_smem<DK, SP>(), stream, a);
  return launch_status();
}

int attn_fwd_tc(const AttnArgs& a, int G, cudaStream_t stream) {
  if (a.S <= 128) return a.dk == 32 ? launch_fwd<32, 128>(a, G, stream) : launch_fwd<64, 128>(a, G, stream);
  return a.dk == 32 ? launch_fwd<32, 256>(a, G, stream) : launch_fwd<64, 256>(a, G, stream);
}
int attn_bwd_tc(const AttnArgs& a, int G, cudaStream_t stream) {
  if (a.S <= 128) return a.dk == 32 ? launch_bwd<32, 128>(a, G, stream) : launch_bwd<64, 128>(a, G, stream);
  return a.dk == 32 ? launch_bwd<32, 256>(a, G, stream) : launch_bwd<64, 256>(a, G, stream);
}

}  // namespace cpc
