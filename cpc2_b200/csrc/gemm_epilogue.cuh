// Epilogue of the tcgen05 GEMM (gemm.cu): TMEM -> registers -> global, specialised per MODE so that the inner
// loops carry no runtime feature tests.  For the K = 256 shapes that dominate the CPC step the epilogue, not the
// MMA main loop, bounds the kernel (a 128 x 256 tile needs ~0.5 us of tensor time), so the code here is written
// for instruction count and memory-level parallelism:
//   * every global load a chunk needs (residual / gate / previous fp32 value for 4 rows x 4 column groups) is
//     issued before the first dependent instruction, so one memory round trip covers the whole chunk;
//   * addresses are lane-constant bases plus compile-time-strided offsets;
//   * dropout folds its 1/(1-p) scale into alpha and the bias, and draws one 32-bit hash per two columns.
#pragma once
#include "cpc_internal.h"
#include "ptx.cuh"

namespace cpc {

// (timing probe only) true for the CTA's first tile: accumulator buffer 0, first use
__device__ __forceinline__ bool acc_full_bar_is_first(const uint64_t* bar) { return (reinterpret_cast<uintptr_t>(bar) & 8) == 0; }

enum EpiMode { EPI_PLAIN = 0, EPI_RES = 1, EPI_GATE = 2, EPI_DROP = 3, EPI_ACCUM = 4, EPI_SPLITK = 5, EPI_NORM = 6, EPI_NORM_CL = 7 };

struct EpiRng {
  uint32_t seed_lo, seed_hi, thresh;
  float scale;
};

__device__ __forceinline__ float bf16lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// 32-bit counter hash (lowbias32 finaliser): two 16-bit draws for a pair of consecutive columns
__device__ __forceinline__ uint32_t epi_hash(uint32_t idx2, uint32_t seed_lo, uint32_t seed_hi) {
  uint32_t h = idx2 * 0x9E3779B1u + seed_lo;
  h ^= seed_hi;
  h ^= h >> 16; h *= 0x7FEB352Du;
  h ^= h >> 15; h *= 0x846CA68Bu;
  h ^= h >> 16;
  return h;
}

// 4 x 4 transpose of 32-bit words inside each quad of lanes: on entry lane t (= lane % 4) holds w[j] = element (t, j),
// on exit w[k] = element (k, t).  tcgen05.ld.16x256b hands a lane two consecutive columns of every 8-column group, so
// a packed bf16x2 store covers 4 bytes per lane and 16 contiguous bytes per row -- HALF a 32-byte sector per row and
// instruction, which made the bf16 stores (not the TMEM reads) the cost of the epilogue (3.7 us per 128 x 256 tile
// against 1.1 us of TMEM read time, tools/gemm_phase_probe.py).  After the transpose lane t owns the whole 8-column
// group t of the chunk: one 16-byte store per lane, 64 contiguous bytes per row, a quarter of the store instructions.
__device__ __forceinline__ void quad_transpose4(uint32_t (&w)[4], int t) {
  const bool odd = (t & 1) != 0, hi = (t & 2) != 0;
  uint32_t s0 = odd ? w[0] : w[1], s1 = odd ? w[2] : w[3];
  uint32_t r0 = __shfl_xor_sync(0xffffffffu, s0, 1), r1 = __shfl_xor_sync(0xffffffffu, s1, 1);
  if (odd) { w[0] = r0; w[2] = r1; } else { w[1] = r0; w[3] = r1; }
  s0 = hi ? w[0] : w[2];
  s1 = hi ? w[1] : w[3];
  r0 = __shfl_xor_sync(0xffffffffu, s0, 2);
  r1 = __shfl_xor_sync(0xffffffffu, s1, 2);
  if (hi) { w[0] = r0; w[1] = r1; } else { w[2] = r0; w[3] = r1; }
}
__device__ __forceinline__ uint32_t epi_pack_bf16x2(float a, float b) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}

template <int MODE>
struct EpiLoads {  // everything one 32-column chunk reads from global memory
  float2 b2[4];
  uint32_t ax[(MODE == EPI_RES || MODE == EPI_GATE) ? 4 : 1][4];
  float2 old[MODE == EPI_ACCUM ? 4 : 1][4];
};

struct EpiLane {  // lane-constant addressing of one tile
  const float* bias;
  __nv_bfloat16* o16;
  float* o32;
  const __nv_bfloat16* aux;  // residual (EPI_RES) or gate (EPI_GATE)
  long long ld_aux;
  int col_first;  // first column of this warp's half (multiple of 8)
  bool rok[4];
};

template <int MODE>
__device__ __forceinline__ void epi_load_chunk(EpiLoads<MODE>& L, const GemmKernelArgs& p, const EpiLane& e, int c) {
  const int jmax = (p.N - (e.col_first + c * 32)) >> 3;
#pragma unroll
  for (int j = 0; j < 4; ++j)
    L.b2[j] = (e.bias != nullptr && j < jmax) ? *reinterpret_cast<const float2*>(e.bias + c * 32 + 8 * j) : make_float2(0.f, 0.f);
  if (MODE == EPI_RES || MODE == EPI_GATE) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        L.ax[i][j] = (e.rok[i] && j < jmax)
                         ? __ldg(reinterpret_cast<const unsigned int*>(e.aux + (long long)(8 * i) * e.ld_aux + c * 32 + 8 * j))
                         : 0u;
  }
  if (MODE == EPI_ACCUM) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        L.old[i][j] = (e.rok[i] && j < jmax)
                          ? *reinterpret_cast<const float2*>(e.o32 + (long long)(8 * i) * p.ldd_f32 + c * 32 + 8 * j)
                          : make_float2(0.f, 0.f);
  }
}

// One 128 x BN accumulator tile.  Warp (q = TMEM lane quarter, half = column half) drains rows q*32 .. +31 of
// columns half*BN/2 .. +BN/2 in chunks of 32 columns; tcgen05.ld.16x256b hands lane (rsub = lane/4, csub =
// 2*(lane%4)) columns csub, csub+1 of every 8-column group for rows rsub + 8*i, i = 0..3 (four lanes cover 32
// contiguous bytes of an fp32 row, 16 of a bf16 row).  The global loads of chunk 0 are issued BEFORE the wait for
// the accumulator (they overlap the MMA main loop), those of chunk c+1 before the arithmetic of chunk c.
template <int MODE, int BN>
__device__ __forceinline__ void epilogue_tile(const GemmKernelArgs& p, const TileCoord& tc, uint32_t taddr, int q,
                                              int half, int lane, const EpiRng& rng, uint64_t* acc_full_bar,
                                              uint32_t parity) {
  constexpr int NC = BN / 64;
  const int rsub = lane >> 2, csub = (lane & 3) * 2;
  const long long g = tc.grp;
  const int row0 = tc.m0 + q * 32 + rsub;
  const int col0 = tc.n0 + half * (BN / 2) + csub;
  EpiLane e;
#pragma unroll
  for (int i = 0; i < 4; ++i) e.rok[i] = row0 + 8 * i < p.M;
  e.col_first = col0 - csub;
  e.bias = p.bias != nullptr ? p.bias + g * p.gs_bias + col0 : nullptr;
  e.o16 = p.out_bf16 != nullptr ? p.out_bf16 + g * p.gs_bf16 + (long long)row0 * p.ldd_bf16 + col0 : nullptr;
  e.o32 = p.out_f32 != nullptr ? p.out_f32 + g * p.gs_f32 + (long long)row0 * p.ldd_f32 + col0 : nullptr;
  e.aux = nullptr;
  e.ld_aux = 0;
  if (MODE == EPI_RES) { e.aux = p.res + g * p.gs_res + (long long)row0 * p.ld_res + col0; e.ld_aux = p.ld_res; }
  if (MODE == EPI_GATE) { e.aux = p.gate + g * p.gs_gate + (long long)row0 * p.ld_gate + col0; e.ld_aux = p.ld_gate; }
  float alpha = p.alpha;
  if (MODE == EPI_DROP) alpha *= rng.scale;
  const float relu_floor = p.relu ? 0.f : -INFINITY;
  const uint32_t pair0 = (uint32_t)(((unsigned long long)g * p.M + row0) * (unsigned long long)p.N + col0) >> 1;
  // EPI_ACCUM keeps a two-deep buffer (its 32 registers per chunk do not fit four times); every other mode requests
  // the loads of ALL chunks before the accumulator is awaited: one L2 round trip per tile instead of one per chunk
  constexpr int LD = MODE == EPI_ACCUM ? 2 : NC;
  EpiLoads<MODE> L[LD];
  if (MODE == EPI_ACCUM) {
    epi_load_chunk<MODE>(L[0], p, e, 0);
  } else {
#pragma unroll
    for (int c = 0; c < NC; ++c) epi_load_chunk<MODE>(L[c % LD], p, e, c);
  }
  mbar_wait(acc_full_bar, parity);
#ifdef CPC_GEMM_TIMING
  if (parity == 0 && threadIdx.x == 64 && acc_full_bar_is_first(acc_full_bar)) GEMM_TSTAMP(7);
#endif
  tc_fence_after();
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int jmax = min(4, (p.N - (e.col_first + c * 32)) >> 3);  // valid 8-column groups (warp-uniform)
    if (jmax <= 0) break;
    uint32_t v[2][16];
    tmem_ld_16x256b_x4(taddr + c * 32, v[0]);
    tmem_ld_16x256b_x4(taddr + (16u << 16) + c * 32, v[1]);
    if (MODE == EPI_ACCUM && c + 1 < NC) epi_load_chunk<MODE>(L[(c + 1) % LD], p, e, c + 1);
    tmem_ld_wait();
    float2 b2[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) b2[j] = L[c % LD].b2[j];
    const bool* rok = e.rok;
    __nv_bfloat16* o16 = e.o16;
    float* o32 = e.o32;
    auto& ax = L[c % LD].ax;
    auto& old = L[c % LD].old;
    if (MODE == EPI_DROP) {
#pragma unroll
      for (int j = 0; j < 4; ++j) { b2[j].x *= rng.scale; b2[j].y *= rng.scale; }
    }
    // ---- arithmetic and stores, row by row (v[i >> 1] holds rows +0 / +8 in regs 0,1 / 2,3 of each group)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float x[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        x[2 * j] = fmaf(__uint_as_float(v[i >> 1][4 * j + 2 * (i & 1)]), alpha, b2[j].x);
        x[2 * j + 1] = fmaf(__uint_as_float(v[i >> 1][4 * j + 2 * (i & 1) + 1]), alpha, b2[j].y);
      }
      if (MODE == EPI_RES) {
#pragma unroll
        for (int j = 0; j < 4; ++j) { x[2 * j] += bf16lo(ax[i][j]); x[2 * j + 1] += bf16hi(ax[i][j]); }
      }
      if (MODE != EPI_SPLITK && MODE != EPI_ACCUM) {
#pragma unroll
        for (int k = 0; k < 8; ++k) x[k] = fmaxf(x[k], relu_floor);
      }
      if (MODE == EPI_GATE) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          x[2 * j] = bf16lo(ax[i][j]) > 0.f ? x[2 * j] * p.gate_scale : 0.f;
          x[2 * j + 1] = bf16hi(ax[i][j]) > 0.f ? x[2 * j + 1] * p.gate_scale : 0.f;
        }
      }
      if (MODE == EPI_DROP) {
        const uint32_t pr = pair0 + (uint32_t)(((long long)(8 * i) * p.N + c * 32) >> 1);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t bits = epi_hash(pr + 4 * j, rng.seed_lo, rng.seed_hi);
          x[2 * j] = (bits & 0xFFFFu) >= rng.thresh ? x[2 * j] : 0.f;
          x[2 * j + 1] = (bits >> 16) >= rng.thresh ? x[2 * j + 1] : 0.f;
        }
      }
      if (MODE == EPI_ACCUM) {
#pragma unroll
        for (int j = 0; j < 4; ++j) { x[2 * j] += old[i][j].x; x[2 * j + 1] += old[i][j].y; }
      }
      if (MODE != EPI_SPLITK && o16 != nullptr) {  // (warp-uniform)
        __nv_bfloat16* dst = o16 + (long long)(8 * i) * p.ldd_bf16 + c * 32;
        if (p.vec16) {
          uint32_t w[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) w[j] = epi_pack_bf16x2(x[2 * j], x[2 * j + 1]);
          quad_transpose4(w, lane & 3);  // before any row test: every lane takes part in the shuffles
          if (rok[i] && (lane & 3) < jmax)
            *reinterpret_cast<uint4*>(dst - csub + 8 * (lane & 3)) = make_uint4(w[0], w[1], w[2], w[3]);
        } else if (rok[i]) {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (j < jmax) *reinterpret_cast<__nv_bfloat162*>(dst + 8 * j) = __floats2bfloat162_rn(x[2 * j], x[2 * j + 1]);
        }
      }
      if (!rok[i]) continue;
      if (o32 != nullptr) {
        float* dst = o32 + (long long)(8 * i) * p.ldd_f32 + c * 32;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (j < jmax) {
            if (MODE == EPI_SPLITK) red_add_v2_f32(dst + 8 * j, x[2 * j], x[2 * j + 1]);
            else *reinterpret_cast<float2*>(dst + 8 * j) = make_float2(x[2 * j], x[2 * j + 1]);
          }
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------
// EPI_NORM: row normalisation fused into the epilogue (ChannelNorm of conv1-4, cpc/model.py:52-60; nn.LayerNorm of
// the transformer blocks, cpc/transformers.py:124,130).  N == BN == 256, so the accumulator tile holds WHOLE rows:
//   sweep 1   x = alpha*acc + bias (+ residual); pre-norm x -> out_bf16 (saved for the backward); per-row sum(x),
//             sum(x^2) in registers, folded over the four lanes that share a row (two shuffles) and over the two warps
//             that share it (column halves) through 2 KB of shared memory and one 64-thread named barrier;
//   sweep 2   the accumulator is read from TMEM a second time (reads are ~free next to the global stores):
//             y = (x - mean) * rstd * gamma + beta, ReLU, -> norm_bf16 / norm_f32; mean / rstd -> fp32 [M].
// fp32 sum / sum-of-squares over 256 values: the relative error of the variance is ~1e-6 * (1 + mean^2/var), far below
// the bf16 rounding of the outputs.  The zero halo rows the next strided conv reads around each sequence are written
// by the CTA that owns the first / last row tile of the group.
__device__ __forceinline__ void epi_pair_barrier(int q) {
  asm volatile("bar.sync %0, 64;" ::"r"(q + 1) : "memory");
}

template <int BN>
__device__ __forceinline__ void epilogue_tile_norm(const GemmKernelArgs& p, const TileCoord& tc, uint32_t taddr, int q,
                                                   int half, int lane, uint64_t* acc_full_bar, uint32_t parity,
                                                   float2* stat) {
  static_assert(BN == 256, "fused row norm: one 256-column tile per row");
  constexpr int NC = BN / 64;  // 32-column chunks per warp (this warp owns BN/2 columns)
  const int rsub = lane >> 2, csub = (lane & 3) * 2;
  const long long g = tc.grp;
  const int row0 = tc.m0 + q * 32 + rsub;
  const int col0 = half * (BN / 2) + csub;
  bool rok[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) rok[i] = row0 + 8 * i < p.M;
  const float* bias = p.bias != nullptr ? p.bias + g * p.gs_bias + col0 : nullptr;
  const float* gam = p.norm_gamma + g * p.gs_norm_param + col0;
  const float* bet = p.norm_beta + g * p.gs_norm_param + col0;
  const __nv_bfloat16* res = p.res != nullptr ? p.res + g * p.gs_res + (long long)row0 * p.ld_res + col0 : nullptr;
  __nv_bfloat16* o16 = p.out_bf16 != nullptr ? p.out_bf16 + g * p.gs_bf16 + (long long)row0 * p.ldd_bf16 + col0 : nullptr;
  __nv_bfloat16* n16 = p.norm_bf16 != nullptr ? p.norm_bf16 + g * p.gs_norm_bf16 + (long long)row0 * p.ld_norm_bf16 + col0 : nullptr;
  float* n32 = p.norm_f32 != nullptr ? p.norm_f32 + g * p.gs_norm_f32 + (long long)row0 * p.ld_norm_f32 + col0 : nullptr;
  const float alpha = p.alpha;

  float s[4] = {0.f, 0.f, 0.f, 0.f}, ss[4] = {0.f, 0.f, 0.f, 0.f};
  // the first chunk's bias / residual are requested before the accumulator is awaited (they overlap the main loop)
  // bias and residual of chunk c + 1 are requested while chunk c is processed (two register slots: all four chunks at
  // once put this kernel over the 168 registers a 10-warp CTA can have and it spilled)
  float2 b2[2][4];
  uint32_t rx[2][4][4];
  auto load_chunk = [&](int c, int slot) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      b2[slot][j] = bias != nullptr ? *reinterpret_cast<const float2*>(bias + c * 32 + 8 * j) : make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        rx[slot][i][j] = (res != nullptr && rok[i])
                             ? __ldg(reinterpret_cast<const unsigned int*>(res + (long long)(8 * i) * p.ld_res + c * 32 + 8 * j))
                             : 0u;
  };
  load_chunk(0, 0);
  mbar_wait(acc_full_bar, parity);
#ifdef CPC_GEMM_TIMING
  if (parity == 0 && threadIdx.x == 64 && acc_full_bar_is_first(acc_full_bar)) GEMM_TSTAMP(7);
#endif
  tc_fence_after();
  // ---------------------------------------------------------------- sweep 1: pre-norm values + row statistics
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    uint32_t v[2][16];
    tmem_ld_16x256b_x4(taddr + c * 32, v[0]);
    tmem_ld_16x256b_x4(taddr + (16u << 16) + c * 32, v[1]);
    if (c + 1 < NC) load_chunk(c + 1, (c + 1) & 1);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float x0 = fmaf(__uint_as_float(v[i >> 1][4 * j + 2 * (i & 1)]), alpha, b2[c & 1][j].x) + bf16lo(rx[c & 1][i][j]);
        float x1 = fmaf(__uint_as_float(v[i >> 1][4 * j + 2 * (i & 1) + 1]), alpha, b2[c & 1][j].y) + bf16hi(rx[c & 1][i][j]);
        s[i] += x0 + x1;
        ss[i] = fmaf(x0, x0, fmaf(x1, x1, ss[i]));
        w[j] = epi_pack_bf16x2(x0, x1);
      }
      if (o16 != nullptr) {  // (warp-uniform)
        __nv_bfloat16* dst = o16 + (long long)(8 * i) * p.ldd_bf16 + c * 32;
        if (p.vec16) {
          quad_transpose4(w, lane & 3);
          if (rok[i]) *reinterpret_cast<uint4*>(dst - csub + 8 * (lane & 3)) = make_uint4(w[0], w[1], w[2], w[3]);
        } else if (rok[i]) {
#pragma unroll
          for (int j = 0; j < 4; ++j) *reinterpret_cast<uint32_t*>(dst + 8 * j) = w[j];
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    s[i] += __shfl_xor_sync(0xffffffffu, s[i], 1);
    ss[i] += __shfl_xor_sync(0xffffffffu, ss[i], 1);
    s[i] += __shfl_xor_sync(0xffffffffu, s[i], 2);
    ss[i] += __shfl_xor_sync(0xffffffffu, ss[i], 2);
  }
  if ((lane & 3) == 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) stat[half * BM + q * 32 + rsub + 8 * i] = make_float2(s[i], ss[i]);
  }
  epi_pair_barrier(q);  // the two warps that share these 32 rows (column halves 0 / 1)
  float mean[4], rstd[4];
  const float inv_n = 1.f / static_cast<float>(BN);
  const float inv_d = 1.f / static_cast<float>(p.norm_unbiased ? BN - 1 : BN);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 o = stat[(half ^ 1) * BM + q * 32 + rsub + 8 * i];
    const float tot = s[i] + o.x, tots = ss[i] + o.y;
    mean[i] = tot * inv_n;
    const float var = fmaxf((tots - tot * mean[i]) * inv_d, 0.f);
    rstd[i] = rsqrtf(var + p.norm_eps);
  }
  if (half == 0 && (lane & 3) == 0 && p.norm_mean != nullptr) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (rok[i]) {
        p.norm_mean[g * p.gs_norm_stat + row0 + 8 * i] = mean[i];
        p.norm_rstd[g * p.gs_norm_stat + row0 + 8 * i] = rstd[i];
      }
    }
  }
  // ---------------------------------------------------------------- sweep 2: normalise, affine, ReLU, store
  const float relu_floor = p.relu ? 0.f : -INFINITY;
  load_chunk(0, 0);
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    uint32_t v[2][16];
    tmem_ld_16x256b_x4(taddr + c * 32, v[0]);
    tmem_ld_16x256b_x4(taddr + (16u << 16) + c * 32, v[1]);
    float2 g2[4], e2[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      g2[j] = *reinterpret_cast<const float2*>(gam + c * 32 + 8 * j);
      e2[j] = *reinterpret_cast<const float2*>(bet + c * 32 + 8 * j);
    }
    if (c + 1 < NC) load_chunk(c + 1, (c + 1) & 1);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float x0 = fmaf(__uint_as_float(v[i >> 1][4 * j + 2 * (i & 1)]), alpha, b2[c & 1][j].x) + bf16lo(rx[c & 1][i][j]);
        const float x1 = fmaf(__uint_as_float(v[i >> 1][4 * j + 2 * (i & 1) + 1]), alpha, b2[c & 1][j].y) + bf16hi(rx[c & 1][i][j]);
        const float y0 = fmaxf(fmaf((x0 - mean[i]) * rstd[i], g2[j].x, e2[j].x), relu_floor);
        const float y1 = fmaxf(fmaf((x1 - mean[i]) * rstd[i], g2[j].y, e2[j].y), relu_floor);
        w[j] = epi_pack_bf16x2(y0, y1);
        if (n32 != nullptr && rok[i])
          *reinterpret_cast<float2*>(n32 + (long long)(8 * i) * p.ld_norm_f32 + c * 32 + 8 * j) = make_float2(y0, y1);
      }
      if (n16 != nullptr) {  // (warp-uniform)
        __nv_bfloat16* dst = n16 + (long long)(8 * i) * p.ld_norm_bf16 + c * 32;
        if (p.vec16n) {
          quad_transpose4(w, lane & 3);
          if (rok[i]) *reinterpret_cast<uint4*>(dst - csub + 8 * (lane & 3)) = make_uint4(w[0], w[1], w[2], w[3]);
        } else if (rok[i]) {
#pragma unroll
          for (int j = 0; j < 4; ++j) *reinterpret_cast<uint32_t*>(dst + 8 * j) = w[j];
        }
      }
    }
  }
  // ---------------------------------------------------------------- zero halo rows around the sequence (group)
  if (p.norm_halo > 0 && q == 0 && p.norm_bf16 != nullptr) {
    __nv_bfloat16* base = p.norm_bf16 + g * p.gs_norm_bf16 + half * (BN / 2) + lane * 4;
    if (tc.m0 == 0)
      for (int h = 1; h <= p.norm_halo; ++h) *reinterpret_cast<uint2*>(base - (long long)h * p.ld_norm_bf16) = make_uint2(0u, 0u);
    if (tc.m0 + BM >= p.M)
      for (int h = 0; h < p.norm_halo; ++h)
        *reinterpret_cast<uint2*>(base + (long long)(p.M + h) * p.ld_norm_bf16) = make_uint2(0u, 0u);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// EPI_NORM_CL: the same row normalisation with the row SPLIT OVER A THREAD-BLOCK CLUSTER.  A 256- or 512-wide row is
// cut in CN = 4 column chunks of BN = 64 / 128; the CN CTAs of a cluster (cluster rank = chunk index) each run the
// main loop of one 128 x BN tile, reduce (sum x, sum x^2) over their own columns, push those partials into every
// peer's shared memory with st.async (data + mbarrier complete_tx in one DSMEM hop, as the recurrent kernels do),
// add up the 2*CN partials of a row and normalise their own columns.  Why: with one BN = 256 tile per row conv2 / 3 / 4
// of the encoder are 32 / 16 / 8 CTAs on 148 SMs, each running a 16-k-block main loop and a two-sweep epilogue alone
// (17 us); four times as many CTAs with a quarter of the MMA, load and epilogue work each are not bound by one SM's
// L2 link and TMEM read port (tools/gemm_phase_probe.py), and 512-wide rows (hidden 512) become fusable at all.
// One tile per CTA (grid = tiles, not persistent): the exchange barrier is used once, phase 0.
struct NormClSmem {
  float2* xstat;    // [2 * CN][BM]: slot 2 * rank + column half
  uint64_t* xbar;   // armed for (CN - 1) * 2 * BM * 8 bytes at kernel start
  int cn;
};

template <int BN>
__device__ __forceinline__ void epilogue_tile_norm_cl(const GemmKernelArgs& p, const TileCoord& tc, uint32_t taddr, int q,
                                                      int half, int lane, uint64_t* acc_full_bar, uint32_t parity,
                                                      const NormClSmem& x) {
  static_assert(BN == 64 || BN == 128, "cluster row norm: 64 or 128 columns per CTA");
  constexpr int NC = BN / 64;  // 32-column chunks per warp (this warp owns BN/2 columns)
  const int rsub = lane >> 2, csub = (lane & 3) * 2;
  const long long g = tc.grp;
  const int row0 = tc.m0 + q * 32 + rsub;
  const int col0 = tc.n0 + half * (BN / 2) + csub;  // global column of this lane's first pair
  const uint32_t rank = cluster_ctarank();
  bool rok[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) rok[i] = row0 + 8 * i < p.M;
  const float* bias = p.bias != nullptr ? p.bias + g * p.gs_bias + col0 : nullptr;
  const float* gam = p.norm_gamma + g * p.gs_norm_param + col0;
  const float* bet = p.norm_beta + g * p.gs_norm_param + col0;
  const __nv_bfloat16* res = p.res != nullptr ? p.res + g * p.gs_res + (long long)row0 * p.ld_res + col0 : nullptr;
  __nv_bfloat16* o16 = p.out_bf16 != nullptr ? p.out_bf16 + g * p.gs_bf16 + (long long)row0 * p.ldd_bf16 + col0 : nullptr;
  __nv_bfloat16* n16 = p.norm_bf16 != nullptr ? p.norm_bf16 + g * p.gs_norm_bf16 + (long long)row0 * p.ld_norm_bf16 + col0 : nullptr;
  float* n32 = p.norm_f32 != nullptr ? p.norm_f32 + g * p.gs_norm_f32 + (long long)row0 * p.ld_norm_f32 + col0 : nullptr;
  const float alpha = p.alpha;

  // everything the tile reads from global memory, requested before the accumulator is awaited
  float2 b2[NC][4];
  uint32_t rx[NC][4][4];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      b2[c][j] = bias != nullptr ? *reinterpret_cast<const float2*>(bias + c * 32 + 8 * j) : make_float2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        rx[c][i][j] = (res != nullptr && rok[i])
                          ? __ldg(reinterpret_cast<const unsigned int*>(res + (long long)(8 * i) * p.ld_res + c * 32 + 8 * j))
                          : 0u;
  }
  mbar_wait(acc_full_bar, parity);
#ifdef CPC_GEMM_TIMING
  if (threadIdx.x == 64) GEMM_TSTAMP(7);
#endif
  tc_fence_after();
  // ---------------------------------------------------------------- sweep 1: pre-norm values + partial row statistics
  float s[4] = {0.f, 0.f, 0.f, 0.f}, ss[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    uint32_t v[2][16];
    tmem_ld_16x256b_x4(taddr + c * 32, v[0]);
    tmem_ld_16x256b_x4(taddr + (16u << 16) + c * 32, v[1]);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float x0 = fmaf(__uint_as_float(v[i >> 1][4 * j + 2 * (i & 1)]), alpha, b2[c][j].x) + bf16lo(rx[c][i][j]);
        const float x1 = fmaf(__uint_as_float(v[i >> 1][4 * j + 2 * (i & 1) + 1]), alpha, b2[c][j].y) + bf16hi(rx[c][i][j]);
        s[i] += x0 + x1;
        ss[i] = fmaf(x0, x0, fmaf(x1, x1, ss[i]));
        w[j] = epi_pack_bf16x2(x0, x1);
      }
      if (o16 != nullptr) {  // (warp-uniform)
        __nv_bfloat16* dst = o16 + (long long)(8 * i) * p.ldd_bf16 + c * 32;
        if (p.vec16) {
          quad_transpose4(w, lane & 3);
          if (rok[i]) *reinterpret_cast<uint4*>(dst - csub + 8 * (lane & 3)) = make_uint4(w[0], w[1], w[2], w[3]);
        } else if (rok[i]) {
#pragma unroll
          for (int j = 0; j < 4; ++j) *reinterpret_cast<uint32_t*>(dst + 8 * j) = w[j];
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    s[i] += __shfl_xor_sync(0xffffffffu, s[i], 1);
    ss[i] += __shfl_xor_sync(0xffffffffu, ss[i], 1);
    s[i] += __shfl_xor_sync(0xffffffffu, s[i], 2);
    ss[i] += __shfl_xor_sync(0xffffffffu, ss[i], 2);
  }
  // ---------------------------------------------------------------- exchange: this warp's partials to every CTA of the cluster
  const int slot = 2 * (int)rank + half;
  if ((lane & 3) == 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2* mine = x.xstat + slot * BM + q * 32 + rsub + 8 * i;
      *mine = make_float2(s[i], ss[i]);
      const uint32_t a = smem_u32(mine), b = smem_u32(x.xbar);
      for (uint32_t r = 0; r < (uint32_t)x.cn; ++r)
        if (r != rank) st_async_v2_f32(mapa(a, r), s[i], ss[i], mapa(b, r));
    }
  }
  asm volatile("bar.sync 1, 256;" ::: "memory");  // the eight epilogue warps: local partials written
  mbar_wait(x.xbar, 0);                           // ... and the (CN - 1) * 2 KB of the peers have landed
  float mean[4], rstd[4];
  const int width = BN * x.cn;
  const float inv_n = 1.f / static_cast<float>(width);
  const float inv_d = 1.f / static_cast<float>(p.norm_unbiased ? width - 1 : width);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float tot = 0.f, tots = 0.f;
    for (int sl = 0; sl < 2 * x.cn; ++sl) {
      const float2 o = x.xstat[sl * BM + q * 32 + rsub + 8 * i];
      tot += o.x;
      tots += o.y;
    }
    mean[i] = tot * inv_n;
    const float var = fmaxf((tots - tot * mean[i]) * inv_d, 0.f);
    rstd[i] = rsqrtf(var + p.norm_eps);
  }
  if (rank == 0 && half == 0 && (lane & 3) == 0 && p.norm_mean != nullptr) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (rok[i]) {
        p.norm_mean[g * p.gs_norm_stat + row0 + 8 * i] = mean[i];
        p.norm_rstd[g * p.gs_norm_stat + row0 + 8 * i] = rstd[i];
      }
    }
  }
  // ---------------------------------------------------------------- sweep 2: normalise, affine, ReLU, store
  const float relu_floor = p.relu ? 0.f : -INFINITY;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    uint32_t v[2][16];
    tmem_ld_16x256b_x4(taddr + c * 32, v[0]);
    tmem_ld_16x256b_x4(taddr + (16u << 16) + c * 32, v[1]);
    float2 g2[4], e2[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      g2[j] = *reinterpret_cast<const float2*>(gam + c * 32 + 8 * j);
      e2[j] = *reinterpret_cast<const float2*>(bet + c * 32 + 8 * j);
    }
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t w[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float x0 = fmaf(__uint_as_float(v[i >> 1][4 * j + 2 * (i & 1)]), alpha, b2[c][j].x) + bf16lo(rx[c][i][j]);
        const float x1 = fmaf(__uint_as_float(v[i >> 1][4 * j + 2 * (i & 1) + 1]), alpha, b2[c][j].y) + bf16hi(rx[c][i][j]);
        const float y0 = fmaxf(fmaf((x0 - mean[i]) * rstd[i], g2[j].x, e2[j].x), relu_floor);
        const float y1 = fmaxf(fmaf((x1 - mean[i]) * rstd[i], g2[j].y, e2[j].y), relu_floor);
        w[j] = epi_pack_bf16x2(y0, y1);
        if (n32 != nullptr && rok[i])
          *reinterpret_cast<float2*>(n32 + (long long)(8 * i) * p.ld_norm_f32 + c * 32 + 8 * j) = make_float2(y0, y1);
      }
      if (n16 != nullptr) {  // (warp-uniform)
        __nv_bfloat16* dst = n16 + (long long)(8 * i) * p.ld_norm_bf16 + c * 32;
        if (p.vec16n) {
          quad_transpose4(w, lane & 3);
          if (rok[i]) *reinterpret_cast<uint4*>(dst - csub + 8 * (lane & 3)) = make_uint4(w[0], w[1], w[2], w[3]);
        } else if (rok[i]) {
#pragma unroll
          for (int j = 0; j < 4; ++j) *reinterpret_cast<uint32_t*>(dst + 8 * j) = w[j];
        }
      }
    }
  }
  // ---------------------------------------------------------------- zero halo rows around the sequence (own columns)
  if (p.norm_halo > 0 && q == 0 && p.norm_bf16 != nullptr && lane * 4 < BN / 2) {
    __nv_bfloat16* base = p.norm_bf16 + g * p.gs_norm_bf16 + tc.n0 + half * (BN / 2) + lane * 4;
    if (tc.m0 == 0)
      for (int h = 1; h <= p.norm_halo; ++h) *reinterpret_cast<uint2*>(base - (long long)h * p.ld_norm_bf16) = make_uint2(0u, 0u);
    if (tc.m0 + BM >= p.M)
      for (int h = 0; h < p.norm_halo; ++h)
        *reinterpret_cast<uint2*>(base + (long long)(p.M + h) * p.ld_norm_bf16) = make_uint2(0u, 0u);
  }
}

}  // namespace cpc
