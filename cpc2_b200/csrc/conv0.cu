// conv0 (C_in = 1, k = 10, stride 5, pad 3) + ChannelNorm + ReLU, forward and backward (cpc/model.py:85,103,52-60).
//
// With one input channel the layer moves 2 bytes per 10 MACs: it is HBM-bound (write of the bf16 activation in
// the forward, read of its gradient in the backward).  To get the instruction count out of the way the 10-tap
// filter runs on the warp-level tensor cores as a [16 frames x 32] . [32 x C] product per tile:
//     k index  0.. 9 : x_hi[tap]   * w_hi[tap]
//             10..19 : x_lo[tap]   * w_hi[tap]
//             20..29 : x_hi[tap]   * w_lo[tap]
//             30, 31 : 1           * bias_hi, bias_lo
// (hi = bf16(v), lo = bf16(v - hi): the split keeps ~16 mantissa bits of the waveform and of the weights, fp32
// accumulation).  A CTA owns a strip of 64 consecutive frames of one window; warp w owns channels [CW*w, CW*w+CW).
// Per 16-frame tile the warps exchange (mean, M2) partials through shared memory (Chan's combination: one barrier,
// no cancellation), normalise in registers and write bf16 rows through an stmatrix-staged, 16-byte-per-lane store.
// The backward recomputes the conv the same way (bit-identical ReLU mask), needs no input gradient (the input is
// audio) and reduces dW / db with a further MMA whose reduction index is the frame:
//     D[16 x C] += A1[16 x 16 frames] . (dy_hi + dy_lo),   A1 rows 0..9 = x_hi taps, row 10 = ones (-> db)
// (dy reaches the B-fragment layout through movmatrix.trans); dgamma / dbeta accumulate in fp32 registers.
#include <cstdlib>

#include "cpc_internal.h"
#include "ptx.cuh"

namespace cpc {

constexpr int K0 = 10, S0 = 5, P0 = 3;
constexpr int STRIP = 64;  // frames per CTA
constexpr int TILES = STRIP / 16;

struct Conv0Args {
  const float* x;  // [N, W] waveform, rows x_stride elements apart
  long long x_stride;
  const float* w;  // [C, 1, 10]
  const float* b;  // [C]
  const float* gamma;
  const float* beta;
  __nv_bfloat16* out;  // [N, L0 + 2*halo, C]
  float* mean;         // [N*L0]
  float* rstd;
  int N, W, L0, halo, C;
  float eps;
  // backward only
  const __nv_bfloat16* dout;  // [N, L0 + 2*halo, C] gradient w.r.t. the (padded) layer output
  float* dw;                  // [C, 10]
  float* db;
  float* dgamma;
  float* dbeta;
};

__device__ __forceinline__ void stsm_x4(uint32_t addr, uint32_t r0, uint32_t r1, uint32_t r2, uint32_t r3) {
  asm volatile("stmatrix.sync.aligned.m8n8.x4.shared.b16 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(r0), "r"(r1),
               "r"(r2), "r"(r3)
               : "memory");
}
__device__ __forceinline__ uint32_t movmatrix_t(uint32_t v) {
  uint32_t r;
  asm volatile("movmatrix.sync.aligned.m8n8.trans.b16 %0, %1;" : "=r"(r) : "r"(v));
  return r;
}
__device__ __forceinline__ float bf16_hi(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
__device__ __forceinline__ uint16_t bf16_bits(float v) {
  const __nv_bfloat16 h = __float2bfloat16_rn(v);
  return *reinterpret_cast<const uint16_t*>(&h);
}

// One 64-byte operand row (32 reduction slots, bf16): slots 0..9 = p0, 10..19 = p1, 20..29 = p2, 30 = e30, 31 = e31,
// written as four 16-byte chunks XOR-swizzled by (row >> 1) & 3 so that the 8 row addresses of an ldmatrix
// phase fall in 8 different bank groups.  Used for the A strip (rows = frames) and the B operand (rows = channels).
__device__ __forceinline__ uint32_t a_chunk_off(int row, int chunk) { return row * 64 + ((chunk ^ ((row >> 1) & 3)) << 4); }

__device__ __forceinline__ void store_operand_row(unsigned char* base, int row, const uint16_t (&p0)[K0],
                                                  const uint16_t (&p1)[K0], const uint16_t (&p2)[K0], uint16_t e30,
                                                  uint16_t e31) {
  uint16_t v[32];
#pragma unroll
  for (int k = 0; k < K0; ++k) {
    v[k] = p0[k];
    v[10 + k] = p1[k];
    v[20 + k] = p2[k];
  }
  v[30] = e30;
  v[31] = e31;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint32_t u[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) u[j] = (uint32_t)v[8 * c + 2 * j] | ((uint32_t)v[8 * c + 2 * j + 1] << 16);
    *reinterpret_cast<uint4*>(base + a_chunk_off(row, c)) = make_uint4(u[0], u[1], u[2], u[3]);
  }
}

__device__ __forceinline__ void split_bf16(float v, uint16_t& hi, uint16_t& lo) {
  const float h = bf16_hi(v);
  hi = bf16_bits(h);
  lo = bf16_bits(v - h);
}

// A strip: row r = frame t0 + r -> [x_hi taps | x_lo taps | x_hi taps | 1 1]; frames past the window are zero rows.
// When sA1 != null also the transposed-use operand of the backward: [frame][16] = x_hi taps, 1, zeros (32-byte rows).
__device__ __forceinline__ void build_a_rows(unsigned char* sA, unsigned char* sA1, int rows,
                                             const float* __restrict__ xs, int W, int L0, int t0) {
  for (int row = threadIdx.x; row < rows; row += blockDim.x) {
    const int t = t0 + row;
    uint16_t hi[K0], lo[K0];
    const bool valid = t < L0;
    const int i0 = t * S0 - P0;
#pragma unroll
    for (int k = 0; k < K0; ++k) {
      const int i = i0 + k;
      const float v = (valid && i >= 0 && i < W) ? __ldg(xs + i) : 0.f;
      split_bf16(v, hi[k], lo[k]);
    }
    const uint16_t one = valid ? 0x3F80 : 0;
    store_operand_row(sA, row, hi, lo, hi, one, one);
    if (sA1 != nullptr) {
      uint32_t u[8];
#pragma unroll
      for (int j = 0; j < 5; ++j) u[j] = (uint32_t)hi[2 * j] | ((uint32_t)hi[2 * j + 1] << 16);
      u[5] = one;
      u[6] = u[7] = 0u;
      uint4* d = reinterpret_cast<uint4*>(sA1 + row * 32);
      d[0] = make_uint4(u[0], u[1], u[2], u[3]);
      d[1] = make_uint4(u[4], u[5], u[6], u[7]);
    }
  }
}

// B operand: row n = channel -> [w_hi taps | w_hi taps | w_lo taps | bias_hi bias_lo]
__device__ __forceinline__ void build_b_rows(unsigned char* sB, const float* __restrict__ w,
                                             const float* __restrict__ bias, int C) {
  for (int n = threadIdx.x; n < C; n += blockDim.x) {
    uint16_t hi[K0], lo[K0], bh, bl;
#pragma unroll
    for (int k = 0; k < K0; ++k) split_bf16(__ldg(w + n * K0 + k), hi[k], lo[k]);
    split_bf16(__ldg(bias + n), bh, bl);
    store_operand_row(sB, n, hi, hi, lo, bh, bl);
  }
}

// y[16 frames x 8*NT channels] of one tile: acc[j][0..1] = frame g, channels 8j + 2q + {0,1}; acc[j][2..3] = frame g+8.
// B fragments come from shared memory per tile (one ldmatrix.x4 per 8 channels) instead of living in registers.
template <int NT>
__device__ __forceinline__ void conv_tile(float (&acc)[NT][4], uint32_t sB_addr, int c0, uint32_t sA_addr, int tile,
                                          int lane) {
  uint32_t a[2][4];
  const int m = lane >> 3, r = (lane & 7) + ((m & 1) << 3) + tile * 16;
#pragma unroll
  for (int s = 0; s < 2; ++s) ldsm_x4(a[s], sA_addr + a_chunk_off(r, 2 * s + (m >> 1)));
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    uint32_t b[4];  // k 0-7, 8-15, 16-23, 24-31 of channels c0 + 8j .. +7
    ldsm_x4(b, sB_addr + a_chunk_off(c0 + 8 * j + (lane & 7), m));
    acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f;
    mma_bf16_16816(acc[j], a[0], b[0], b[1]);
    mma_bf16_16816(acc[j], a[1], b[2], b[3]);
  }
}

template <int CW>
__device__ __forceinline__ uint32_t stage_off(int row, int chunk) {
  if constexpr (CW == 64) return row * 128 + ((chunk ^ (row & 7)) << 4);
  else return row * 64 + ((chunk ^ ((row >> 1) & 3)) << 4);
}

__device__ __forceinline__ float quad_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  return v;
}

// zero halo rows of one window's padded output (the next conv's implicit GEMM reads them as its zero padding)
__device__ __forceinline__ void zero_rows(__nv_bfloat16* base, int rows, int C) {
  uint4* p = reinterpret_cast<uint4*>(base);
  for (int i = threadIdx.x; i < rows * C / 8; i += blockDim.x) p[i] = make_uint4(0u, 0u, 0u, 0u);
}

template <int NT, int NW>
__global__ void __launch_bounds__(32 * NW) conv0_cn_relu_fwd_kernel(const Conv0Args a) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  constexpr int CW = 8 * NT, C = CW * NW;
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, q = lane & 3;
  unsigned char* sA = smem;
  unsigned char* sB = smem + STRIP * 64;                                    // [C][64 B]
  float2* sStat = reinterpret_cast<float2*>(sB + C * 64);                   // [2][NW][16]
  unsigned char* sStage = sB + C * 64 + 2 * NW * 16 * 8 + warp * 16 * CW * 2;
  const int spw = (a.L0 + STRIP - 1) / STRIP;
  const int n = blockIdx.x / spw, t0 = (blockIdx.x % spw) * STRIP;
  const int Lp = a.L0 + 2 * a.halo;
  build_a_rows(sA, nullptr, STRIP, a.x + (long long)n * a.x_stride, a.W, a.L0, t0);
  build_b_rows(sB, a.w, a.b, C);
  if (a.halo > 0) {
    if (t0 == 0) zero_rows(a.out + (long long)n * Lp * C, a.halo, C);
    if (t0 + STRIP >= a.L0) zero_rows(a.out + ((long long)n * Lp + a.halo + a.L0) * C, a.halo, C);
  }
  const int c0 = warp * CW;
  float2 ga[NT], be[NT];
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    ga[j] = *reinterpret_cast<const float2*>(a.gamma + c0 + 8 * j + 2 * q);
    be[j] = *reinterpret_cast<const float2*>(a.beta + c0 + 8 * j + 2 * q);
  }
  __syncthreads();
  const uint32_t sA_addr = smem_u32(sA), sB_addr = smem_u32(sB), stage_addr = smem_u32(sStage);
  const float invCm1 = 1.f / (float)(C - 1);
  constexpr int LPR = CW / 8, RPP = 32 / LPR;  // staged store: lanes per row, rows per pass
  __nv_bfloat16* out_lane = a.out + ((long long)n * Lp + a.halo + t0 + lane / LPR) * C + c0 + (lane % LPR) * 8;
  const int m = lane >> 3;
  const uint32_t stsm_lane = stage_addr + ((lane & 7) + ((m & 1) << 3)) * (CW * 2);
  const int stsm_row = (lane & 7) + ((m & 1) << 3);
#pragma unroll 1
  for (int tile = 0; tile < TILES; ++tile) {
    const int tt = t0 + tile * 16;
    if (tt >= a.L0) break;  // uniform over the CTA
    float acc[NT][4];
    conv_tile<NT>(acc, sB_addr, c0, sA_addr, tile, lane);
    // warp-local mean over this warp's CW channels for frames g and g+8; centre on it; M2 of the centred values
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      s0 += acc[j][0] + acc[j][1];
      s1 += acc[j][2] + acc[j][3];
    }
    const float m0 = quad_sum(s0) * (1.f / CW), m1 = quad_sum(s1) * (1.f / CW);
    float q0 = 0.f, q1 = 0.f;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      acc[j][0] -= m0; acc[j][1] -= m0; acc[j][2] -= m1; acc[j][3] -= m1;
      q0 = fmaf(acc[j][0], acc[j][0], q0); q0 = fmaf(acc[j][1], acc[j][1], q0);
      q1 = fmaf(acc[j][2], acc[j][2], q1); q1 = fmaf(acc[j][3], acc[j][3], q1);
    }
    q0 = quad_sum(q0);
    q1 = quad_sum(q1);
    float2* st = sStat + (tile & 1) * NW * 16;
    if (q == 0) {
      st[warp * 16 + g] = make_float2(m0, q0);
      st[warp * 16 + g + 8] = make_float2(m1, q1);
    }
    __syncthreads();
    // Chan's combination of the NW (mean, M2) partials of equal weight CW
    float2 u0[NW], u1[NW];
    float mean0 = 0.f, mean1 = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      u0[w] = st[w * 16 + g];
      u1[w] = st[w * 16 + g + 8];
      mean0 += u0[w].x;
      mean1 += u1[w].x;
    }
    mean0 *= 1.f / NW;
    mean1 *= 1.f / NW;
    float M0 = 0.f, M1 = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      M0 += fmaf(CW * (u0[w].x - mean0), u0[w].x - mean0, u0[w].y);
      M1 += fmaf(CW * (u1[w].x - mean1), u1[w].x - mean1, u1[w].y);
    }
    const float r0 = rsqrtf(M0 * invCm1 + a.eps), r1 = rsqrtf(M1 * invCm1 + a.eps);
    if (warp == 0 && q == 0) {
      const long long row = (long long)n * a.L0 + tt;
      if (tt + g < a.L0) { a.mean[row + g] = mean0; a.rstd[row + g] = r0; }
      if (tt + g + 8 < a.L0) { a.mean[row + g + 8] = mean1; a.rstd[row + g + 8] = r1; }
    }
    // xhat = (centred + (m_w - mean)) * rstd; y = relu(xhat * gamma + beta); pack to bf16, stage for 16-byte row stores
    const float k0 = (m0 - mean0) * r0, k1 = (m1 - mean1) * r1;
#pragma unroll
    for (int j = 0; j < NT; j += 2) {
      uint32_t pk[4];
#pragma unroll
      for (int jj = 0; jj < 2; ++jj) {
        const float2 gg = ga[j + jj], bb = be[j + jj];
        const float y00 = fmaxf(fmaf(fmaf(acc[j + jj][0], r0, k0), gg.x, bb.x), 0.f);
        const float y01 = fmaxf(fmaf(fmaf(acc[j + jj][1], r0, k0), gg.y, bb.y), 0.f);
        const float y10 = fmaxf(fmaf(fmaf(acc[j + jj][2], r1, k1), gg.x, bb.x), 0.f);
        const float y11 = fmaxf(fmaf(fmaf(acc[j + jj][3], r1, k1), gg.y, bb.y), 0.f);
        pk[2 * jj] = pack_bf16x2(y00, y01);
        pk[2 * jj + 1] = pack_bf16x2(y10, y11);
      }
      const int chunk = j + (m >> 1);
      const uint32_t swz = CW == 64 ? (uint32_t)((chunk ^ (stsm_row & 7)) << 4) : (uint32_t)((chunk ^ ((stsm_row >> 1) & 3)) << 4);
      stsm_x4(stsm_lane + swz, pk[0], pk[1], pk[2], pk[3]);
    }
    __syncwarp();
#pragma unroll
    for (int p = 0; p < 16 / RPP; ++p) {
      const int row = p * RPP + lane / LPR;
      const uint4 v = *reinterpret_cast<const uint4*>(sStage + stage_off<CW>(row, lane % LPR));
      if (tt + row < a.L0) *reinterpret_cast<uint4*>(out_lane + (long long)(tile * 16 + p * RPP) * C) = v;
    }
    __syncwarp();
  }
}

// ------------------------------------------------------------------------------------------------ backward
constexpr int STRIPB = 32;  // frames per work item of the backward (2 tiles); CTAs loop over items

template <int NT, int NW>
__global__ void __launch_bounds__(32 * NW) conv0_cn_relu_bwd_kernel(const Conv0Args a) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  constexpr int CW = 8 * NT;
  extern __shared__ __align__(128) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, q = lane & 3;
  unsigned char* sA = smem;                                           // [STRIPB][64 B]
  unsigned char* sA1 = smem + STRIPB * 64;                            // [STRIPB][32 B]
  unsigned char* sB = smem + STRIPB * 96;                             // [C][64 B]
  float2* sStat = reinterpret_cast<float2*>(sB + a.C * 64);           // [2][NW][16] (s1, s2) partials
  float2* sGB = sStat + 2 * NW * 16;                                  // [C] (gamma, beta)
  for (int c = threadIdx.x; c < a.C; c += blockDim.x) sGB[c] = make_float2(a.gamma[c], a.beta[c]);
  build_b_rows(sB, a.w, a.b, a.C);
  const int c0 = warp * CW;
  float D[NT][4];       // rows m = g (taps 0..7) and m = g + 8 (taps 8, 9, then db) of [16 x CW]
  float dgs[NT][2], dbs[NT][2];  // dgamma / dbeta partials of channels 8j + 2q + {0,1} (fp32, CUDA cores)
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    D[j][0] = D[j][1] = D[j][2] = D[j][3] = 0.f;
    dgs[j][0] = dgs[j][1] = dbs[j][0] = dbs[j][1] = 0.f;
  }
  const int ipw = (a.L0 + STRIPB - 1) / STRIPB;
  const int items = a.N * ipw;
  const int Lp = a.L0 + 2 * a.halo;
  const uint32_t sA_addr = smem_u32(sA), sA1_addr = smem_u32(sA1), sB_addr = smem_u32(sB);
  const float invC = 1.f / (float)a.C, invCm1 = 1.f / (float)(a.C - 1);
  int phase = 0;
  // Everything a tile reads from global memory (its 16 x CW slice of dout, mean and rstd of its rows) is requested ONE
  // TILE AHEAD (8 warps per SM and a block barrier in every tile leave nothing else to cover the round trip).  Measured:
  // 72.9 -> 72.3 us for the step's two launches -- the loads were not what the kernel waits for; a tile is ~3000
  // dependent instructions per warp (conv MMAs, 32 elements of fp32 arithmetic per thread, barrier, weight-gradient
  // MMAs) on two warps per scheduler.  Kept: it costs no registers (178 against 180).
  uint32_t dv[NT][2], dvn[NT][2];
  float rs0 = 0.f, rs1 = 0.f, mn0 = 0.f, mn1 = 0.f, rs0n = 0.f, rs1n = 0.f, mn0n = 0.f, mn1n = 0.f;
  auto load_tile = [&](int n, int tt, uint32_t (&d)[NT][2], float& r0, float& r1, float& m0, float& m1) {
    const bool v0 = tt + g < a.L0, v1 = tt + g + 8 < a.L0;
    const long long row = (long long)n * a.L0 + tt;
    r0 = v0 ? __ldg(a.rstd + row + g) : 0.f;
    r1 = v1 ? __ldg(a.rstd + row + g + 8) : 0.f;
    m0 = v0 ? __ldg(a.mean + row + g) : 0.f;
    m1 = v1 ? __ldg(a.mean + row + g + 8) : 0.f;
    const __nv_bfloat16* dr0 = a.dout + ((long long)n * Lp + a.halo + tt + g) * a.C + c0 + 2 * q;
    const __nv_bfloat16* dr1 = dr0 + 8LL * a.C;
#pragma unroll
    for (int j = 0; j < NT; ++j) {
      d[j][0] = v0 ? __ldg(reinterpret_cast<const unsigned int*>(dr0 + 8 * j)) : 0u;
      d[j][1] = v1 ? __ldg(reinterpret_cast<const unsigned int*>(dr1 + 8 * j)) : 0u;
    }
  };
  if ((int)blockIdx.x < items) load_tile(blockIdx.x / ipw, (blockIdx.x % ipw) * STRIPB, dv, rs0, rs1, mn0, mn1);
#pragma unroll 1
  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    const int n = item / ipw, t0 = (item % ipw) * STRIPB;
    __syncthreads();  // every warp is done with the previous item's operands
    build_a_rows(sA, sA1, STRIPB, a.x + (long long)n * a.x_stride, a.W, a.L0, t0);
    __syncthreads();
#pragma unroll 1
    for (int tile = 0; tile < STRIPB / 16; ++tile, phase ^= 1) {
      const int tt = t0 + tile * 16;
      if (tt >= a.L0) break;  // uniform over the CTA
      {  // the tile this CTA processes next: the following one of this item, else the first one of its next item
        int nn = n, ntt = tt + 16;
        bool more = tile + 1 < STRIPB / 16 && ntt < a.L0;
        if (!more) {
          const int ni = item + gridDim.x;
          more = ni < items;
          nn = ni / ipw;
          ntt = (ni % ipw) * STRIPB;
        }
        if (more) load_tile(nn, ntt, dvn, rs0n, rs1n, mn0n, mn1n);
      }
      const float rstd0 = rs0, rstd1 = rs1;
      const float nm0 = -mn0 * rstd0, nm1 = -mn1 * rstd1;
      float acc[NT][4];
      conv_tile<NT>(acc, sB_addr, c0, sA_addr, tile, lane);
      float s1a = 0.f, s2a = 0.f, s1b = 0.f, s2b = 0.f;
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        const float2 gb0 = sGB[c0 + 8 * j + 2 * q], gb1 = sGB[c0 + 8 * j + 2 * q + 1];
        float d[4] = {__uint_as_float(dv[j][0] << 16), __uint_as_float(dv[j][0] & 0xffff0000u),
                      __uint_as_float(dv[j][1] << 16), __uint_as_float(dv[j][1] & 0xffff0000u)};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 gb = (i & 1) ? gb1 : gb0;
          const float xh = fmaf(acc[j][i], i < 2 ? rstd0 : rstd1, i < 2 ? nm0 : nm1);
          if (fmaf(xh, gb.x, gb.y) <= 0.f) d[i] = 0.f;  // ReLU mask (same arithmetic as the forward)
          acc[j][i] = xh;
          dgs[j][i & 1] = fmaf(d[i], xh, dgs[j][i & 1]);
          dbs[j][i & 1] += d[i];
          const float dn = d[i] * gb.x;
          if (i < 2) { s1a += dn; s2a = fmaf(dn, xh, s2a); } else { s1b += dn; s2b = fmaf(dn, xh, s2b); }
        }
        dv[j][0] = pack_bf16x2(d[0], d[1]);  // masked gradient, exact in bf16
        dv[j][1] = pack_bf16x2(d[2], d[3]);
      }
      s1a = quad_sum(s1a); s2a = quad_sum(s2a); s1b = quad_sum(s1b); s2b = quad_sum(s2b);
      float2* st = sStat + phase * NW * 16;
      if (q == 0) {
        st[warp * 16 + g] = make_float2(s1a, s2a);
        st[warp * 16 + g + 8] = make_float2(s1b, s2b);
      }
      __syncthreads();
      float S1a = 0.f, S2a = 0.f, S1b = 0.f, S2b = 0.f;
#pragma unroll
      for (int w = 0; w < NW; ++w) {
        const float2 u0 = st[w * 16 + g], u1 = st[w * 16 + g + 8];
        S1a += u0.x; S2a += u0.y; S1b += u1.x; S2b += u1.y;
      }
      // dy = rstd * (dn - mean_c(dn) - xhat * sum_c(dn * xhat) / (C - 1)) = dn * rstd + (xhat * kb + ka)
      const float ka0 = -S1a * invC * rstd0, kb0 = -S2a * invCm1 * rstd0;
      const float ka1 = -S1b * invC * rstd1, kb1 = -S2b * invCm1 * rstd1;
      uint32_t a1[4];
      {
        const int mi = lane >> 3, r = lane & 7;
        ldsm_x4_t(a1, sA1_addr + (tile * 16 + r + 8 * (mi >> 1)) * 32 + (mi & 1) * 16);
      }
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        const float g0 = sGB[c0 + 8 * j + 2 * q].x, g1 = sGB[c0 + 8 * j + 2 * q + 1].x;
        const float dy0 = fmaf(__uint_as_float(dv[j][0] << 16) * g0, rstd0, fmaf(acc[j][0], kb0, ka0));
        const float dy1 = fmaf(__uint_as_float(dv[j][0] & 0xffff0000u) * g1, rstd0, fmaf(acc[j][1], kb0, ka0));
        const float dy2 = fmaf(__uint_as_float(dv[j][1] << 16) * g0, rstd1, fmaf(acc[j][2], kb1, ka1));
        const float dy3 = fmaf(__uint_as_float(dv[j][1] & 0xffff0000u) * g1, rstd1, fmaf(acc[j][3], kb1, ka1));
        // dW[tap] += sum_frames x[tap] * dy (rows 0..9), db += sum_frames dy (row 10)
        // dy enters the tensor core as bf16 hi + lo (~16 mantissa bits): the sum runs over every frame of the batch
        const uint32_t h01 = pack_bf16x2(dy0, dy1), h23 = pack_bf16x2(dy2, dy3);
        mma_bf16_16816(D[j], a1, movmatrix_t(h01), movmatrix_t(h23));
        const uint32_t l01 = pack_bf16x2(dy0 - __uint_as_float(h01 << 16), dy1 - __uint_as_float(h01 & 0xffff0000u));
        const uint32_t l23 = pack_bf16x2(dy2 - __uint_as_float(h23 << 16), dy3 - __uint_as_float(h23 & 0xffff0000u));
        mma_bf16_16816(D[j], a1, movmatrix_t(l01), movmatrix_t(l23));
      }
#pragma unroll
      for (int j = 0; j < NT; ++j) { dv[j][0] = dvn[j][0]; dv[j][1] = dvn[j][1]; }
      rs0 = rs0n; rs1 = rs1n; mn0 = mn0n; mn1 = mn1n;
    }
  }
  // D rows: thread holds m = g (taps 0..7) in D[j][0..1] and m = g + 8 (taps 8, 9, db) in D[j][2..3]
#pragma unroll
  for (int j = 0; j < NT; ++j) {
    const int c = c0 + 8 * j + 2 * q;
    red_add_f32(a.dw + (long long)c * K0 + g, D[j][0]);
    red_add_f32(a.dw + (long long)(c + 1) * K0 + g, D[j][1]);
    if (g < 2) {
      red_add_f32(a.dw + (long long)c * K0 + g + 8, D[j][2]);
      red_add_f32(a.dw + (long long)(c + 1) * K0 + g + 8, D[j][3]);
    } else if (g == 2) {
      red_add_v2_f32(a.db + c, D[j][2], D[j][3]);
    }
    // dgamma / dbeta: sum the 8 lanes that share q (different frames), lanes 0..3 publish
    float v[4] = {dgs[j][0], dgs[j][1], dbs[j][0], dbs[j][1]};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[i] += __shfl_xor_sync(0xffffffffu, v[i], 4);
      v[i] += __shfl_xor_sync(0xffffffffu, v[i], 8);
      v[i] += __shfl_xor_sync(0xffffffffu, v[i], 16);
    }
    if (g == 0) {
      red_add_v2_f32(a.dgamma + c, v[0], v[1]);
      red_add_v2_f32(a.dbeta + c, v[2], v[3]);
    }
  }
}

}  // namespace cpc

using namespace cpc;

static int fill_conv0(Conv0Args& a, const CpcConv0* d, int& NT, int& NW) {
  if (d == nullptr || d->N <= 0 || d->W <= 0 || d->C <= 0) return CPC_ERR_ARG;
  if (d->C % 64 == 0 && d->C / 64 <= 8) { NT = 8; NW = d->C / 64; }
  else if (d->C % 32 == 0 && d->C / 32 <= 16) { NT = 4; NW = d->C / 32; }
  else return CPC_ERR_UNSUPPORTED;
  a.x = static_cast<const float*>(d->x); a.w = static_cast<const float*>(d->w); a.b = static_cast<const float*>(d->b);
  a.gamma = static_cast<const float*>(d->gamma); a.beta = static_cast<const float*>(d->beta);
  a.out = static_cast<__nv_bfloat16*>(d->out); a.mean = static_cast<float*>(d->mean); a.rstd = static_cast<float*>(d->rstd);
  a.x_stride = d->x_stride > 0 ? d->x_stride : d->W;
  a.N = d->N; a.W = d->W; a.L0 = (d->W + 2 * P0 - K0) / S0 + 1; a.halo = d->halo; a.C = d->C; a.eps = d->eps;
  a.dout = static_cast<const __nv_bfloat16*>(d->dout);
  a.dw = static_cast<float*>(d->dw); a.db = static_cast<float*>(d->db);
  a.dgamma = static_cast<float*>(d->dgamma); a.dbeta = static_cast<float*>(d->dbeta);
  if (a.x == nullptr || a.w == nullptr || a.b == nullptr || a.gamma == nullptr || a.beta == nullptr) return CPC_ERR_ARG;
  return CPC_OK;
}

extern "C" int cpc_conv0_cn_relu_fwd(const CpcConv0* d, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  Conv0Args a;
  int NT, NW;
  int rc = fill_conv0(a, d, NT, NW);
  if (rc != CPC_OK) return rc;
  if (a.out == nullptr || a.mean == nullptr || a.rstd == nullptr) return CPC_ERR_ARG;
  const int spw = (a.L0 + STRIP - 1) / STRIP;
  const unsigned blocks = (unsigned)(a.N * spw);
  const size_t smem = STRIP * 64 + (size_t)a.C * 64 + 2 * NW * 16 * 8 + (size_t)NW * 16 * (8 * NT) * 2;
#define CPC_CONV0_FWD(NT_, NW_)                                                                                       \
  case NW_: {                                                                                                          \
    if (smem > 48 * 1024)                                                                                              \
      cudaFuncSetAttribute(conv0_cn_relu_fwd_kernel<NT_, NW_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    launch_k(conv0_cn_relu_fwd_kernel<NT_, NW_>, dim3(blocks), dim3(32 * NW_), smem, stream, a);                                         \
    break;                                                                                                             \
  }
  if (NT == 8) {
    switch (NW) {
      CPC_CONV0_FWD(8, 1) CPC_CONV0_FWD(8, 2) CPC_CONV0_FWD(8, 3) CPC_CONV0_FWD(8, 4) CPC_CONV0_FWD(8, 5)
      CPC_CONV0_FWD(8, 6) CPC_CONV0_FWD(8, 7) CPC_CONV0_FWD(8, 8)
      default: return CPC_ERR_UNSUPPORTED;
    }
  } else {
    switch (NW) {
      CPC_CONV0_FWD(4, 1) CPC_CONV0_FWD(4, 3) CPC_CONV0_FWD(4, 5) CPC_CONV0_FWD(4, 7) CPC_CONV0_FWD(4, 9)
      CPC_CONV0_FWD(4, 11) CPC_CONV0_FWD(4, 13) CPC_CONV0_FWD(4, 15)
      default: return CPC_ERR_UNSUPPORTED;
    }
  }
#undef CPC_CONV0_FWD
  return launch_status();
}

extern "C" int cpc_conv0_cn_relu_bwd(const CpcConv0* d, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  Conv0Args a;
  int NT, NW;
  int rc = fill_conv0(a, d, NT, NW);
  if (rc != CPC_OK) return rc;
  if (a.dout == nullptr || a.dw == nullptr || a.db == nullptr || a.dgamma == nullptr || a.dbeta == nullptr ||
      a.mean == nullptr || a.rstd == nullptr)
    return CPC_ERR_ARG;
  // 64 channels per warp (NT = 8, ~180 registers, 8 warps per SM).  32 channels per warp (NT = 4, 128 registers, 16 warps
  // per SM) was tried: per launch inside the step 36 -> 31 us, but the whole step got 0.4 % SLOWER (1.069 -> 1.074 ms,
  // same box, two repeats: twice the resident warps next to the side-stream weight-gradient GEMMs this kernel overlaps)
  // and alone with a cold L2 it is slower too (47 -> 53 us for 16 windows); profiles/r02_ab_conv0_bwd.txt.
  const int items = a.N * ((a.L0 + STRIPB - 1) / STRIPB);
  const size_t smem = STRIPB * 96 + (size_t)a.C * 64 + 2 * NW * 16 * 8 + (size_t)a.C * 8;
  int sms = 148, dev = 0;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
#define CPC_CONV0_BWD(NT_, NW_)                                                                                       \
  case NW_: {                                                                                                          \
    if (smem > 48 * 1024)                                                                                              \
      cudaFuncSetAttribute(conv0_cn_relu_bwd_kernel<NT_, NW_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    int occ = 1;                                                                                                       \
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, conv0_cn_relu_bwd_kernel<NT_, NW_>, 32 * NW_, smem);           \
    long long blocks = (long long)sms * (occ < 1 ? 1 : occ);                                                           \
    if (blocks > items) blocks = items;                                                                                \
    launch_k(conv0_cn_relu_bwd_kernel<NT_, NW_>, dim3((unsigned)blocks), dim3(32 * NW_), smem, stream, a);                               \
    break;                                                                                                             \
  }
  if (NT == 8) {
    switch (NW) {
      CPC_CONV0_BWD(8, 1) CPC_CONV0_BWD(8, 2) CPC_CONV0_BWD(8, 3) CPC_CONV0_BWD(8, 4) CPC_CONV0_BWD(8, 5)
      CPC_CONV0_BWD(8, 6) CPC_CONV0_BWD(8, 7) CPC_CONV0_BWD(8, 8)
      default: return CPC_ERR_UNSUPPORTED;
    }
  } else {
    switch (NW) {
      CPC_CONV0_BWD(4, 1) CPC_CONV0_BWD(4, 3) CPC_CONV0_BWD(4, 5) CPC_CONV0_BWD(4, 7) CPC_CONV0_BWD(4, 9)
      CPC_CONV0_BWD(4, 11) CPC_CONV0_BWD(4, 13) CPC_CONV0_BWD(4, 15)
      default: return CPC_ERR_UNSUPPORTED;
    }
  }
#undef CPC_CONV0_BWD
  return launch_status();
}
