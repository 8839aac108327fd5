// Tensor-core path of the causal relative-position attention (cpc/transformers.py:52-70), S <= 128, dk in {32, 64}:
// the shapes of the default CPC step (12 prediction heads x 8 windows x 8 attention heads, S = 116, dk = 32) and
// of the transformer AR (S = 128).  One CTA per (head group, window, attention head), four warps; each warp owns
// 16-row query blocks {w, 7-w} (causal work is balanced), all products run on mma.sync.m16n8k16 bf16 with
// operands fetched by ldmatrix from padded (conflict-free) shared-memory tiles:
//
//   forward   R = Q.Krel (only the r >= S-16-i0 columns a block can address), staged per warp in shared memory;
//             scores = (Q.K^T + skew(R)) * scale with skew(R)[i,j] = R[i, S-1-i+j]  -- the reference's
//             cat/view "skew" is an index shift here; softmax in registers (quad shuffles); P (bf16) is kept
//             for the backward; O = dropout(P).V with the score accumulators re-used as A fragments.
//   backward  row pass (query blocks): dP = dO.V^T, dS = P o (dP - D) * scale, dQ = dS.K + unskew(dS).Krel^T,
//             dKrel += Q^T.unskew(dS) (shared-memory fp32 accumulation, one red.global per element per CTA);
//             column pass (key blocks, transposed tiles): dV = P~^T.dO, dK = dS^T.Q.
//             Both passes recompute dP / dS from registers, so nothing but P ever round-trips through memory.
// Dropout masks are regenerated from a counter hash of the element index in both directions.
// Larger S / odd head sizes fall back to the CUDA-core kernels of attention.cu.
#include <cstdlib>

#include "cpc_internal.h"
#include "ptx.cuh"
#include "attn_args.h"

namespace cpc {

namespace {

constexpr int SP = 128;      // padded sequence length of the shared-memory tiles
constexpr int PS = SP + 8;   // pitch (elements) of [.][SP] tiles: 272 B rows, conflict-free for ldmatrix
constexpr int ATT_THREADS = 128;

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
    d.thresh = static_cast<uint32_t>(fminf(a.drop_p, 0.999999f) * 4294967296.0f);
    d.scale = 1.f / (1.f - a.drop_p);
  }
  return d;
}
__device__ __forceinline__ bool keep_elem(const DropCfg& d, long long e) {
  return rng_pair_u16(d.seed_lo, d.seed_hi, static_cast<unsigned long long>(e)) >= d.thresh;
}

// rows [0, S) of a [S][dk] slice (row stride `ld` elements) -> shared [SP][PD] bf16, rows >= S zero
template <int DK>
__device__ __forceinline__ void load_rows(__nv_bfloat16* dst, const __nv_bfloat16* src, long long ld, int S) {
  constexpr int PD = DK + 8, CH = DK / 8;
  for (int i = threadIdx.x; i < SP * CH; i += ATT_THREADS) {
    const int s = i / CH, ch = i % CH;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (s < S) v = *reinterpret_cast<const uint4*>(src + (long long)s * ld + ch * 8);
    *reinterpret_cast<uint4*>(dst + s * PD + ch * 8) = v;
  }
}

}  // namespace

// ================================================================================================ forward
template <int DK>
__global__ void __launch_bounds__(ATT_THREADS) attn_fwd_tc_kernel(const AttnArgs a) {
  constexpr int PD = DK + 8, KS = DK / 16, NT = DK / 8, RP = SP + 8;
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
  load_rows<DK>(Qs, base, 3 * d, S);
  load_rows<DK>(Ks, base + d, 3 * d, S);
  load_rows<DK>(Vs, base + 2 * d, 3 * d, S);
  if (rel) {
    const float* kr = a.krel + (long long)grp * DK * S;
    for (int i = threadIdx.x; i < DK * SP; i += ATT_THREADS) {
      const int c = i / SP, r = i % SP;
      KrT[r * PD + c] = __float2bfloat16(r < S ? kr[c * S + r] : 0.f);
    }
  }
  __syncthreads();
  const DropCfg drop = make_drop(a);
  const long long pbase = (((long long)grp * a.Bt + b) * a.Hh + h) * S * S;
  float* Rs = Rs_all + warp * 16 * RP;
  const int ntile_s = (S + 7) >> 3;
  const bool pair_ok = (S & 1) == 0;

  for (int pass = 0; pass < 2; ++pass) {
    const int rb = pass ? 7 - warp : warp;
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
    float sc[16][4];
#pragma unroll
    for (int jt = 0; jt < 16; ++jt) {
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
    for (int jt = 0; jt < 16; ++jt)
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
    for (int jt = 0; jt < 16; ++jt)
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
    // probabilities out (bf16, lower triangle); the copy that feeds P.V keeps fp32 accuracy (hi + lo bf16 pair
    // below): the backward's D_i = dO_i.O_i must equal sum_j P_ij dP_ij for the UNROUNDED P, otherwise the
    // rounding error of P, weighted by the common component of dP, leaks into every dS of the row.
#pragma unroll
    for (int jt = 0; jt < 16; ++jt) {
      if (jt < njt) {
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const int i = i0 + g + 8 * r, j = jt * 8 + 2 * t;
          float p0 = sc[jt][2 * r] * inv[r], p1 = sc[jt][2 * r + 1] * inv[r];
          if (i < S && j <= i) {
            const long long e = pbase + (long long)i * S + j;
            if (a.probs != nullptr) {
              const __nv_bfloat162 pb = __floats2bfloat162_rn(p0, p1);
              if (pair_ok && j + 1 <= i) *reinterpret_cast<__nv_bfloat162*>(a.probs + e) = pb;
              else {
                a.probs[e] = __low2bfloat16(pb);
                if (j + 1 <= i) a.probs[e + 1] = __high2bfloat16(pb);
              }
            }
            if (drop.thresh != 0) {
              p0 = keep_elem(drop, e) ? p0 * drop.scale : 0.f;
              p1 = keep_elem(drop, e + 1) ? p1 * drop.scale : 0.f;
            }
          }
          sc[jt][2 * r] = p0;
          sc[jt][2 * r + 1] = p1;
        }
      }
    }
    // O = P~ . V : two adjacent score tiles form the A fragment of one 16-key step
    float o[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) o[nt][e] = 0.f;
#pragma unroll
    for (int kk = 0; kk < 8; ++kk) {
      if (2 * kk < njt) {
        uint32_t ap[4], al[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float x0 = sc[2 * kk + (q >> 1)][2 * (q & 1)], x1 = sc[2 * kk + (q >> 1)][2 * (q & 1) + 1];
          const __nv_bfloat162 hi = __floats2bfloat162_rn(x0, x1);
          ap[q] = *reinterpret_cast<const uint32_t*>(&hi);
          al[q] = pack_bf16x2(x0 - __low2float(hi), x1 - __high2float(hi));
        }
#pragma unroll
        for (int n2 = 0; n2 < NT / 2; ++n2) {
          uint32_t bb[4];
          ldsm_x4_t(bb, smem_u32(Vs + (kk * 16 + (lane & 15)) * PD + n2 * 16 + (lane >> 4) * 8));
          mma_bf16_16816(o[2 * n2], ap, bb[0], bb[1]);
          mma_bf16_16816(o[2 * n2 + 1], ap, bb[2], bb[3]);
          mma_bf16_16816(o[2 * n2], al, bb[0], bb[1]);
          mma_bf16_16816(o[2 * n2 + 1], al, bb[2], bb[3]);
        }
      }
    }
    __nv_bfloat16* ob = a.out + (long long)grp * a.gs_out + (long long)b * S * d + h * DK;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      const int c = nt * 8 + 2 * t;
      if (i0 + g < S)
        *reinterpret_cast<__nv_bfloat162*>(ob + (long long)(i0 + g) * d + c) = __floats2bfloat162_rn(o[nt][0], o[nt][1]);
      if (i0 + g + 8 < S)
        *reinterpret_cast<__nv_bfloat162*>(ob + (long long)(i0 + g + 8) * d + c) = __floats2bfloat162_rn(o[nt][2], o[nt][3]);
    }
  }
}

// ================================================================================================ backward
template <int DK>
__global__ void __launch_bounds__(ATT_THREADS) attn_bwd_tc_kernel(const AttnArgs a) {
  constexpr int PD = DK + 8, KS = DK / 16, NT = DK / 8;
  extern __shared__ __align__(16) uint8_t smem_dyn[];
  __nv_bfloat16* Qs = reinterpret_cast<__nv_bfloat16*>(smem_dyn);
  __nv_bfloat16* Ks = Qs + SP * PD;
  __nv_bfloat16* Vs = Ks + SP * PD;
  __nv_bfloat16* dOs = Vs + SP * PD;
  __nv_bfloat16* Krs = dOs + SP * PD;                              // Krel [c][r] bf16, pitch PS
  __nv_bfloat16* stage_all = Krs + DK * PS;                        // per warp unskew(dS) [16][PS]
  float* Dsm = reinterpret_cast<float*>(stage_all + 4 * 16 * PS);  // [SP]
  float* dKr = Dsm + SP;                                           // [DK][SP] fp32 accumulator
  const int S = a.S, d = a.d;
  const int h = blockIdx.x, b = blockIdx.y, grp = blockIdx.z;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const bool rel = a.krel != nullptr;
  const __nv_bfloat16* base = a.qkv + (long long)grp * a.gs_qkv + (long long)b * S * 3 * d + h * DK;
  const long long obase = (long long)grp * a.gs_out + (long long)b * S * d + h * DK;
  load_rows<DK>(Qs, base, 3 * d, S);
  load_rows<DK>(Ks, base + d, 3 * d, S);
  load_rows<DK>(Vs, base + 2 * d, 3 * d, S);
  load_rows<DK>(dOs, a.dout + obase, d, S);
  if (rel) {
    const float* kr = a.krel + (long long)grp * DK * S;
    for (int i = threadIdx.x; i < DK * SP; i += ATT_THREADS) {
      const int c = i / SP, r = i % SP;
      Krs[c * PS + r] = __float2bfloat16(r < S ? kr[c * S + r] : 0.f);
      dKr[i] = 0.f;
    }
  }
  __syncthreads();
  // D_i = dO_i . O_i  (= sum_j P~_ij dP~_ij)
  for (int i = threadIdx.x; i < SP; i += ATT_THREADS) {
    float acc = 0.f;
    if (i < S) {
      const __nv_bfloat16* orow = a.out + obase + (long long)i * d;
#pragma unroll
      for (int ch = 0; ch < DK / 8; ++ch) {
        const uint4 ov = *reinterpret_cast<const uint4*>(orow + ch * 8);
        const uint4 dv = *reinterpret_cast<const uint4*>(dOs + i * PD + ch * 8);
        const __nv_bfloat162* o2 = reinterpret_cast<const __nv_bfloat162*>(&ov);
        const __nv_bfloat162* d2 = reinterpret_cast<const __nv_bfloat162*>(&dv);
#pragma unroll
        for (int k = 0; k < 4; ++k) acc += __low2float(o2[k]) * __low2float(d2[k]) + __high2float(o2[k]) * __high2float(d2[k]);
      }
    }
    Dsm[i] = acc;
  }
  __syncthreads();
  const DropCfg drop = make_drop(a);
  const long long pbase = (((long long)grp * a.Bt + b) * a.Hh + h) * S * S;
  const int ntile_s = (S + 7) >> 3;
  const bool pair_ok = (S & 1) == 0;
  __nv_bfloat16* stage = stage_all + warp * 16 * PS;
  __nv_bfloat16* dq_out = a.dqkv + (long long)grp * a.gs_qkv + (long long)b * S * 3 * d + h * DK;

  // ------------------------------------------------------------------ row pass: dQ, dKrel
  for (int pass = 0; pass < 2; ++pass) {
    const int rb = pass ? 7 - warp : warp;
    const int i0 = rb * 16;
    if (i0 >= S) continue;
    uint32_t ado[KS][4];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) ldsm_x4(ado[ks], smem_u32(dOs + (i0 + (lane & 15)) * PD + ks * 16 + (lane >> 4) * 8));
    if (rel) {
      __syncwarp();
      for (int i = lane; i < 16 * PS / 8; i += 32) reinterpret_cast<uint4*>(stage)[i] = make_uint4(0u, 0u, 0u, 0u);
      __syncwarp();
    }
    const int njt = min(2 * rb + 2, ntile_s);
    float dq[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) dq[nt][e] = 0.f;
    const float Di[2] = {Dsm[i0 + g], Dsm[i0 + g + 8]};
    for (int kk = 0; 2 * kk < njt; ++kk) {
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
              const long long e = pbase + (long long)i * S + j;
              float p0, p1 = 0.f;
              if (pair_ok && j + 1 <= i) {
                const __nv_bfloat162 pb = *reinterpret_cast<const __nv_bfloat162*>(a.probs + e);
                p0 = __low2float(pb);
                p1 = __high2float(pb);
              } else {
                p0 = __bfloat162float(a.probs[e]);
                if (j + 1 <= i) p1 = __bfloat162float(a.probs[e + 1]);
              }
              float dp0 = acc[2 * r], dp1 = acc[2 * r + 1];
              if (drop.thresh != 0) {
                dp0 = keep_elem(drop, e) ? dp0 * drop.scale : 0.f;
                dp1 = keep_elem(drop, e + 1) ? dp1 * drop.scale : 0.f;
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
      // dKrel[c][r] += sum_ii Q[i0+ii][c] * unskew(dS)[ii][r]
      if (a.dkrel != nullptr) {
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
              atomicAdd(&dKr[c * SP + r], acc[0]);
              atomicAdd(&dKr[(c + 8) * SP + r], acc[2]);
            }
            if (r + 1 < S) {
              atomicAdd(&dKr[c * SP + r + 1], acc[1]);
              atomicAdd(&dKr[(c + 8) * SP + r + 1], acc[3]);
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

  // ------------------------------------------------------------------ column pass: dK, dV (transposed tiles [j][i])
  for (int pass = 0; pass < 2; ++pass) {
    const int jb = pass ? warp : 7 - warp;
    const int j0 = jb * 16;
    if (j0 >= S) continue;
    uint32_t av[KS][4];
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) ldsm_x4(av[ks], smem_u32(Vs + (j0 + (lane & 15)) * PD + ks * 16 + (lane >> 4) * 8));
    float dk[NT][4], dv[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) dk[nt][e] = dv[nt][e] = 0.f;
    for (int kk = jb; 2 * kk < ntile_s; ++kk) {
      float pt[2][4], dst[2][4];
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        const int it = 2 * kk + hf;
#pragma unroll
        for (int e = 0; e < 4; ++e) pt[hf][e] = dst[hf][e] = 0.f;
        if (it < ntile_s) {
          float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int ks = 0; ks < KS; ++ks) {
            uint32_t bb[2];
            ldsm_x2(bb, smem_u32(dOs + (it * 8 + (lane & 7)) * PD + ks * 16 + ((lane >> 3) & 1) * 8));
            mma_bf16_16816(acc, av[ks], bb[0], bb[1]);
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int j = j0 + g + 8 * (e >> 1), i = it * 8 + 2 * t + (e & 1);
            if (i < S && j <= i) {
              const long long el = pbase + (long long)i * S + j;
              const float p = __bfloat162float(a.probs[el]);
              float pd = p, dp = acc[e];
              if (drop.thresh != 0) {
                const bool keep = keep_elem(drop, el);
                pd = keep ? p * drop.scale : 0.f;
                dp = keep ? dp * drop.scale : 0.f;
              }
              pt[hf][e] = pd;
              dst[hf][e] = p * (dp - Dsm[i]) * a.scale;
            }
          }
        }
      }
      uint32_t ap[4] = {pack_bf16x2(pt[0][0], pt[0][1]), pack_bf16x2(pt[0][2], pt[0][3]),
                        pack_bf16x2(pt[1][0], pt[1][1]), pack_bf16x2(pt[1][2], pt[1][3])};
      uint32_t as[4] = {pack_bf16x2(dst[0][0], dst[0][1]), pack_bf16x2(dst[0][2], dst[0][3]),
                        pack_bf16x2(dst[1][0], dst[1][1]), pack_bf16x2(dst[1][2], dst[1][3])};
#pragma unroll
      for (int n2 = 0; n2 < NT / 2; ++n2) {
        uint32_t bb[4];
        ldsm_x4_t(bb, smem_u32(dOs + (kk * 16 + (lane & 15)) * PD + n2 * 16 + (lane >> 4) * 8));
        mma_bf16_16816(dv[2 * n2], ap, bb[0], bb[1]);
        mma_bf16_16816(dv[2 * n2 + 1], ap, bb[2], bb[3]);
        ldsm_x4_t(bb, smem_u32(Qs + (kk * 16 + (lane & 15)) * PD + n2 * 16 + (lane >> 4) * 8));
        mma_bf16_16816(dk[2 * n2], as, bb[0], bb[1]);
        mma_bf16_16816(dk[2 * n2 + 1], as, bb[2], bb[3]);
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
  if (rel && a.dkrel != nullptr) {
    __syncthreads();
    float* out = a.dkrel + (long long)grp * DK * S;
    for (int i = threadIdx.x; i < DK * S; i += ATT_THREADS) {
      const int c = i / S, r = i % S;
      red_add_f32(out + i, dKr[c * SP + r]);
    }
  }
}

// ================================================================================================ host
template <int DK>
static size_t fwd_smem() { return (size_t)4 * SP * (DK + 8) * 2 + (size_t)4 * 16 * (SP + 8) * 4; }
template <int DK>
static size_t bwd_smem() {
  return (size_t)4 * SP * (DK + 8) * 2 + (size_t)DK * PS * 2 + (size_t)4 * 16 * PS * 2 + (size_t)SP * 4 + (size_t)DK * SP * 4;
}

int attn_tc_supported(const AttnArgs& a) {
  static const bool enabled = [] {  // CPC_ATTN_TC=0 forces the general CUDA-core path (debugging / bisection)
    const char* e = getenv("CPC_ATTN_TC");
    return e == nullptr || e[0] != '0';
  }();
  return enabled && a.S <= SP && a.S >= 1 && (a.dk == 32 || a.dk == 64) && a.d % 8 == 0;
}

template <int DK>
static int launch_fwd(const AttnArgs& a, int G, cudaStream_t stream) {
  auto kern = attn_fwd_tc_kernel<DK>;
  static bool done = false;
  if (!done) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fwd_smem<DK>()) != cudaSuccess)
      return CPC_ERR_LAUNCH;
    done = true;
  }
  kern<<<dim3(a.Hh, a.Bt, G), ATT_THREADS, fwd_smem<DK>(), stream>>>(a);
  return launch_status();
}
template <int DK>
static int launch_bwd(const AttnArgs& a, int G, cudaStream_t stream) {
  auto kern = attn_bwd_tc_kernel<DK>;
  static bool done = false;
  if (!done) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bwd_smem<DK>()) != cudaSuccess)
      return CPC_ERR_LAUNCH;
    done = true;
  }
  kern<<<dim3(a.Hh, a.Bt, G), ATT_THREADS, bwd_smem<DK>(), stream>>>(a);
  return launch_status();
}

int attn_fwd_tc(const AttnArgs& a, int G, cudaStream_t stream) {
  return a.dk == 32 ? launch_fwd<32>(a, G, stream) : launch_fwd<64>(a, G, stream);
}
int attn_bwd_tc(const AttnArgs& a, int G, cudaStream_t stream) {
  return a.dk == 32 ? launch_bwd<32>(a, G, stream) : launch_bwd<64>(a, G, stream);
}

}  // namespace cpc
