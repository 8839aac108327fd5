// Persistent recurrent kernels for CPCAR's nn.LSTM (cpc/model.py:171-173,196), forward and backward.
//
// The recurrence is a chain of S dependent [16 x H] x [H x 4H] products: latency-bound, not FLOP- or
// HBM-bound.  The design removes every per-step trip to L2/HBM for weights:
//   * one thread-block CLUSTER of CL = H/32 CTAs per group of 16 sequences; CTA c owns hidden units
//     [32c, 32c+32) -> 128 gate rows of W_hh.  Those rows live in REGISTERS for the whole kernel as
//     mma.m16n8k16 A-fragments (64 regs/thread at H = 256), so a step reads no weights at all;
//   * per step each warp issues 2*H/16 HMMA (M = 16 gate rows = 4 units x {i,f,g,o}, N = 2x8 sequences,
//     K = H), the four gates of a (unit, sequence) meet in one thread after two lane shuffles, the cell
//     update runs in fp32 registers (c never leaves the register file), and the new h slice (16 x 32 bf16)
//     is pushed into the h buffer of every CTA of the cluster through distributed shared memory with
//     16-byte st.shared::cluster stores;
//   * steps are chained by remote mbarrier arrivals (release.cluster / acquire.cluster) on a double-buffered
//     h tile -- no cluster-wide barrier and no L1 flush inside the time loop.
// The input projection x.W_ih^T (+ both biases) is one large tcgen05 GEMM done beforehand (gx), and the
// weight / input gradients are large tcgen05 GEMMs over the saved gate gradients afterwards; only the
// sequential part lives here.  The backward kernel mirrors the forward: each CTA multiplies its own 128
// gate-gradient columns with its W_hh rows (K-split), and the partial dh tiles are reduce-scattered over
// DSMEM to the CTA that owns the unit.
//
// tcgen05 is deliberately NOT used for the recurrent product: with M = gate rows of one CTA (128) and
// N = 16 sequences a UMMA is ~8 cycles of tensor work behind a TMEM round trip per step, while HMMA from
// register-resident fragments has no accumulator round trip at all; the step time is set by the DSMEM
// exchange either way ("LSTM recurrence").
#include "cpc_internal.h"
#include "ptx.cuh"

namespace cpc {

__device__ __forceinline__ float sigmoid_f(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float tanh_f(float x) { return 2.f * __fdividef(1.f, 1.f + __expf(-2.f * x)) - 1.f; }
__device__ __forceinline__ uint32_t pack_bf16(__nv_bfloat16 lo, __nv_bfloat16 hi) {
  return static_cast<uint32_t>(__bfloat16_as_ushort(lo)) | (static_cast<uint32_t>(__bfloat16_as_ushort(hi)) << 16);
}

struct LstmArgs {
  int N, S;
  const float* gx;            // [N, S, 4H] fp32: x.W_ih^T + b_ih + b_hh
  const __nv_bfloat16* w_hh;  // [4H, H]
  const float* h0;            // [N, H] or null
  const float* c0;            // [N, H] or null
  __nv_bfloat16* h_pad;       // [N, S+1, H] bf16, row 0 = h0, row t+1 = h_t (operand of the dW_hh GEMM)
  float* out;                 // [N, S, H] fp32 module output (optional)
  __nv_bfloat16* gates;       // [N, S, 4H] bf16 post-activation i,f,g,o (saved for backward; optional)
  float* c_all;               // [N, S, H] fp32 cell states (saved for backward; optional)
  float* h_n;                 // [N, H] final hidden (optional)
  float* c_n;                 // [N, H]
  // backward
  const float* dout;          // [N, S, H] fp32 gradient w.r.t. out
  __nv_bfloat16* dgates;      // [N, S, 4H] bf16 gradient w.r.t. the gate pre-activations
};

constexpr int NB = 16;  // sequences per cluster

template <int H>
__global__ void __launch_bounds__(256, 1) lstm_fwd_kernel(const LstmArgs a) {
  constexpr int CL = H / 32, KT = H / 16, HP = H + 8;
  __shared__ __align__(16) __nv_bfloat16 h_s[2][NB][HP];
  __shared__ __align__(16) __nv_bfloat16 stage[NB][32];
  __shared__ __align__(8) uint64_t bar[2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t cta = cluster_ctarank();
  const int n_base = (blockIdx.x / CL) * NB;
  const int g0 = lane >> 4;          // 0: this thread's fragment rows are gates (i, g); 1: (f, o)
  const int u = (lane >> 2) & 3;     // unit inside the warp's 4
  const int unit_l = warp * 4 + u;   // unit inside the CTA's 32
  const int unit = cta * 32 + unit_l;
  const int H4 = 4 * H;

  // W_hh rows of this thread's two fragment rows, resident for the whole kernel
  uint32_t afr[KT][4];
  {
    const __nv_bfloat16* w0 = a.w_hh + ((long long)g0 * H + unit) * H;
    const __nv_bfloat16* w1 = a.w_hh + ((long long)(g0 + 2) * H + unit) * H;
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) {
      const int k = kt * 16 + (lane & 3) * 2;
      afr[kt][0] = *reinterpret_cast<const uint32_t*>(w0 + k);
      afr[kt][1] = *reinterpret_cast<const uint32_t*>(w1 + k);
      afr[kt][2] = *reinterpret_cast<const uint32_t*>(w0 + k + 8);
      afr[kt][3] = *reinterpret_cast<const uint32_t*>(w1 + k + 8);
    }
  }
  for (int i = tid; i < NB * H; i += 256) {
    const int b = i / H, k = i % H, n = n_base + b;
    const float v = (a.h0 != nullptr && n < a.N) ? a.h0[(long long)n * H + k] : 0.f;
    h_s[0][b][k] = __float2bfloat16(v);
    if (a.h_pad != nullptr && n < a.N && (k >> 5) == (int)cta) a.h_pad[(long long)n * (a.S + 1) * H + k] = __float2bfloat16(v);
  }
  if (tid == 0) {
    mbar_init(&bar[0], CL);
    mbar_init(&bar[1], CL);
    fence_barrier_init();
  }
  cluster_sync_all();

  int bsel[2], nidx[2];
  bool valid[2];
  float c_state[2], gxv[2][4];
#pragma unroll
  for (int nt = 0; nt < 2; ++nt) {
    bsel[nt] = nt * 8 + (lane & 3) * 2 + g0;
    nidx[nt] = n_base + bsel[nt];
    valid[nt] = nidx[nt] < a.N;
    c_state[nt] = (valid[nt] && a.c0 != nullptr) ? a.c0[(long long)nidx[nt] * H + unit] : 0.f;
#pragma unroll
    for (int g = 0; g < 4; ++g)
      gxv[nt][g] = valid[nt] ? a.gx[((long long)nidx[nt] * a.S + 0) * H4 + g * H + unit] : 0.f;
  }

  for (int t = 0; t < a.S; ++t) {
    const int p = t & 1;
    if (t > 0) mbar_wait_cluster(&bar[p], ((t - 1) >> 1) & 1);
    // prefetch next step's input projection while the MMAs run
    float gxn[2][4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int g = 0; g < 4; ++g)
        gxn[nt][g] = (valid[nt] && t + 1 < a.S) ? a.gx[((long long)nidx[nt] * a.S + t + 1) * H4 + g * H + unit] : 0.f;

    float acc[2][2][4];
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[i][j][e] = 0.f;
#pragma unroll
    for (int kt = 0; kt < KT; ++kt) {
      const int k = kt * 16 + (lane & 3) * 2;
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        const __nv_bfloat16* hb = &h_s[p][nt * 8 + (lane >> 2)][k];
        const uint32_t b0 = *reinterpret_cast<const uint32_t*>(hb);
        const uint32_t b1 = *reinterpret_cast<const uint32_t*>(hb + 8);
        mma_bf16_16816(acc[kt & 1][nt], afr[kt], b0, b1);
      }
    }
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      float c4[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) c4[e] = acc[0][nt][e] + acc[1][nt][e];
      // rows (lane/4) and (lane/4 + 8) are gates g0 and g0+2 of unit u; the partner lane (xor 16) holds the
      // other two gates.  After the exchange thread g0 owns sequence column (lane%4)*2 + g0.
      const float s0 = g0 ? c4[0] : c4[1];
      const float s1 = g0 ? c4[2] : c4[3];
      const float r0 = __shfl_xor_sync(0xffffffffu, s0, 16);
      const float r1 = __shfl_xor_sync(0xffffffffu, s1, 16);
      const float pi = (g0 ? r0 : c4[0]) + gxv[nt][0];
      const float pf = (g0 ? c4[1] : r0) + gxv[nt][1];
      const float pg = (g0 ? r1 : c4[2]) + gxv[nt][2];
      const float po = (g0 ? c4[3] : r1) + gxv[nt][3];
      const float ig = sigmoid_f(pi), fg = sigmoid_f(pf), gg = tanh_f(pg), og = sigmoid_f(po);
      const float cn = fg * c_state[nt] + ig * gg;
      const float hn = og * tanh_f(cn);
      c_state[nt] = cn;
      stage[bsel[nt]][unit_l] = __float2bfloat16(hn);
      if (valid[nt]) {
        const long long row = (long long)nidx[nt] * a.S + t;
        if (a.out != nullptr) a.out[row * H + unit] = hn;
        if (a.h_pad != nullptr) a.h_pad[((long long)nidx[nt] * (a.S + 1) + t + 1) * H + unit] = __float2bfloat16(hn);
        if (a.c_all != nullptr) a.c_all[row * H + unit] = cn;
        if (a.gates != nullptr) {
          __nv_bfloat16* gp = a.gates + row * H4 + unit;
          gp[0] = __float2bfloat16(ig);
          gp[H] = __float2bfloat16(fg);
          gp[2 * H] = __float2bfloat16(gg);
          gp[3 * H] = __float2bfloat16(og);
        }
        if (t == a.S - 1) {
          if (a.h_n != nullptr) a.h_n[(long long)nidx[nt] * H + unit] = hn;
          if (a.c_n != nullptr) a.c_n[(long long)nidx[nt] * H + unit] = cn;
        }
      }
#pragma unroll
      for (int g = 0; g < 4; ++g) gxv[nt][g] = gxn[nt][g];
    }
    __syncthreads();
    if (t + 1 < a.S) {
      // push this CTA's [16 x 32] slice of h_t into buffer p^1 of every CTA of the cluster
      const int chunk = tid & 63, b = chunk >> 2, part = chunk & 3;
      const uint4 v = *reinterpret_cast<const uint4*>(&stage[b][part * 8]);
      const uint32_t dst = smem_u32(&h_s[p ^ 1][b][cta * 32 + part * 8]);
      for (int d = tid >> 6; d < CL; d += 4) st_cluster_v4(mapa(dst, d), v.x, v.y, v.z, v.w);
      __syncthreads();
      if (tid < CL) mbar_arrive_remote(mapa(smem_u32(&bar[p ^ 1]), tid));
    }
  }
  cluster_sync_all();  // nobody exits while peers may still target its shared memory
}

template <int H>
__global__ void __launch_bounds__(256, 1) lstm_bwd_kernel(const LstmArgs a) {
  constexpr int CL = H / 32;
  constexpr int MTW = (H / 16 + 7) / 8;        // output (h) m-tiles per warp
  constexpr int ACTIVE = (H / 16) / MTW;       // warps that own m-tiles
  constexpr int GP = 128 + 8;                  // padded row of the local gate-gradient tile
  constexpr int SP = H + 4;                    // padded row of the fp32 partial-dh stage
  extern __shared__ __align__(16) uint8_t smem_dyn[];
  float* P = reinterpret_cast<float*>(smem_dyn);                         // [2][CL][NB][32]
  float* stage = P + 2 * CL * NB * 32;                                   // [NB][SP]
  __nv_bfloat16* dg_s = reinterpret_cast<__nv_bfloat16*>(stage + NB * SP);  // [NB][GP]
  __shared__ __align__(8) uint64_t bar[2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t cta = cluster_ctarank();
  const int n_base = (blockIdx.x / CL) * NB;
  const int g0 = lane >> 4;
  const int u = (lane >> 2) & 3;
  const int unit_l = warp * 4 + u;
  const int unit = cta * 32 + unit_l;
  const int H4 = 4 * H;

  // A fragments of W_hh^T restricted to this CTA's 128 gate rows: A[j][r] = W_hh[rg(r)][j],
  // local r = gate*32 + ul  ->  rg = gate*H + cta*32 + ul
  uint32_t afr[MTW][8][4];
  if (warp < ACTIVE) {
#pragma unroll
    for (int mt = 0; mt < MTW; ++mt) {
      const int j0 = (warp * MTW + mt) * 16 + (lane >> 2);
#pragma unroll
      for (int kt = 0; kt < 8; ++kt) {
        const int r = kt * 16 + (lane & 3) * 2;
        auto wv = [&](int rr, int j) {
          const long long rg = (long long)(rr >> 5) * H + cta * 32 + (rr & 31);
          return a.w_hh[rg * H + j];
        };
        afr[mt][kt][0] = pack_bf16(wv(r, j0), wv(r + 1, j0));
        afr[mt][kt][1] = pack_bf16(wv(r, j0 + 8), wv(r + 1, j0 + 8));
        afr[mt][kt][2] = pack_bf16(wv(r + 8, j0), wv(r + 9, j0));
        afr[mt][kt][3] = pack_bf16(wv(r + 8, j0 + 8), wv(r + 9, j0 + 8));
      }
    }
  }
  if (tid == 0) {
    mbar_init(&bar[0], CL);
    mbar_init(&bar[1], CL);
    fence_barrier_init();
  }
  cluster_sync_all();

  int bsel[2], nidx[2];
  bool valid[2];
  float dc_state[2] = {0.f, 0.f};
#pragma unroll
  for (int nt = 0; nt < 2; ++nt) {
    bsel[nt] = nt * 8 + (lane & 3) * 2 + g0;
    nidx[nt] = n_base + bsel[nt];
    valid[nt] = nidx[nt] < a.N;
  }

  for (int sidx = 0; sidx < a.S; ++sidx) {
    const int t = a.S - 1 - sidx;
    const int pp = sidx & 1;
    // loads that do not depend on the recurrence first
    float gi[2], gf[2], gg[2], go[2], ct[2], cprev[2], dh[2];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      if (valid[nt]) {
        const long long row = (long long)nidx[nt] * a.S + t;
        const __nv_bfloat16* gp = a.gates + row * H4 + unit;
        gi[nt] = __bfloat162float(gp[0]);
        gf[nt] = __bfloat162float(gp[H]);
        gg[nt] = __bfloat162float(gp[2 * H]);
        go[nt] = __bfloat162float(gp[3 * H]);
        ct[nt] = a.c_all[row * H + unit];
        cprev[nt] = t > 0 ? a.c_all[(row - 1) * H + unit]
                          : (a.c0 != nullptr ? a.c0[(long long)nidx[nt] * H + unit] : 0.f);
        dh[nt] = a.dout[row * H + unit];
      } else {
        gi[nt] = gf[nt] = gg[nt] = go[nt] = ct[nt] = cprev[nt] = dh[nt] = 0.f;
      }
    }
    if (sidx > 0) {
      mbar_wait_cluster(&bar[pp], ((sidx - 1) >> 1) & 1);
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        float s = 0.f;
#pragma unroll
        for (int src = 0; src < CL; ++src) s += P[((pp * CL + src) * NB + bsel[nt]) * 32 + unit_l];
        dh[nt] += s;
      }
    }
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      const float tc = tanh_f(ct[nt]);
      const float d_o = dh[nt] * tc * go[nt] * (1.f - go[nt]);
      const float dc = dh[nt] * go[nt] * (1.f - tc * tc) + dc_state[nt];
      const float d_i = dc * gg[nt] * gi[nt] * (1.f - gi[nt]);
      const float d_f = dc * cprev[nt] * gf[nt] * (1.f - gf[nt]);
      const float d_g = dc * gi[nt] * (1.f - gg[nt] * gg[nt]);
      dc_state[nt] = dc * gf[nt];
      const __nv_bfloat16 bi = __float2bfloat16(d_i), bf = __float2bfloat16(d_f), bg = __float2bfloat16(d_g),
                          bo = __float2bfloat16(d_o);
      __nv_bfloat16* ds = dg_s + bsel[nt] * GP + unit_l;
      ds[0] = bi; ds[32] = bf; ds[64] = bg; ds[96] = bo;
      if (valid[nt]) {
        __nv_bfloat16* gp = a.dgates + ((long long)nidx[nt] * a.S + t) * H4 + unit;
        gp[0] = bi; gp[H] = bf; gp[2 * H] = bg; gp[3 * H] = bo;
      }
    }
    __syncthreads();
    if (t > 0) {
      // partial dh[b][j] = sum_{r in my 128 gate rows} dgate[b][r] * W_hh[rg(r)][j]
      if (warp < ACTIVE) {
        float acc[MTW][2][4];
#pragma unroll
        for (int mt = 0; mt < MTW; ++mt)
#pragma unroll
          for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) acc[mt][nt][e] = 0.f;
#pragma unroll
        for (int kt = 0; kt < 8; ++kt) {
          const int r = kt * 16 + (lane & 3) * 2;
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) {
            const __nv_bfloat16* bp = dg_s + (nt * 8 + (lane >> 2)) * GP + r;
            const uint32_t b0 = *reinterpret_cast<const uint32_t*>(bp);
            const uint32_t b1 = *reinterpret_cast<const uint32_t*>(bp + 8);
#pragma unroll
            for (int mt = 0; mt < MTW; ++mt) mma_bf16_16816(acc[mt][nt], afr[mt][kt], b0, b1);
          }
        }
#pragma unroll
        for (int mt = 0; mt < MTW; ++mt) {
          const int j0 = (warp * MTW + mt) * 16 + (lane >> 2);
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) {
            const int bb = nt * 8 + (lane & 3) * 2;
            stage[bb * SP + j0] = acc[mt][nt][0];
            stage[(bb + 1) * SP + j0] = acc[mt][nt][1];
            stage[bb * SP + j0 + 8] = acc[mt][nt][2];
            stage[(bb + 1) * SP + j0 + 8] = acc[mt][nt][3];
          }
        }
      }
      __syncthreads();
      // reduce-scatter: block [16 x 32] of units owned by CTA d goes to P[pp^1][src = me] of CTA d
      for (int c = tid; c < CL * 128; c += 256) {
        const int d = c >> 7, b = (c & 127) >> 3, part = c & 7;
        const float4 v = *reinterpret_cast<const float4*>(stage + b * SP + d * 32 + part * 4);
        const uint32_t dst = smem_u32(P + (((pp ^ 1) * CL + cta) * NB + b) * 32 + part * 4);
        st_cluster_v4(mapa(dst, d), __float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z),
                      __float_as_uint(v.w));
      }
      __syncthreads();
      if (tid < CL) mbar_arrive_remote(mapa(smem_u32(&bar[pp ^ 1]), tid));
    }
  }
  cluster_sync_all();
}

template <int H>
static int launch_lstm(const LstmArgs& a, bool backward, cudaStream_t stream) {
  constexpr int CL = H / 32;
  const int clusters = (a.N + NB - 1) / NB;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(CL * clusters);
  cfg.blockDim = dim3(256);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e;
  if (!backward) {
    auto kern = lstm_fwd_kernel<H>;
    if (CL > 8) cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cfg.dynamicSmemBytes = 0;
    e = cudaLaunchKernelEx(&cfg, kern, a);
  } else {
    auto kern = lstm_bwd_kernel<H>;
    const size_t smem = (size_t)(2 * CL * NB * 32 + NB * (H + 4)) * sizeof(float) + (size_t)NB * (128 + 8) * 2;
    if (CL > 8) cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (smem > 48 * 1024) cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cfg.dynamicSmemBytes = smem;
    e = cudaLaunchKernelEx(&cfg, kern, a);
  }
  return e == cudaSuccess ? CPC_OK : CPC_ERR_LAUNCH;
}

}  // namespace cpc

using namespace cpc;

static int fill_lstm(LstmArgs& a, const CpcLstm* d) {
  if (d == nullptr || d->N <= 0 || d->S <= 0 || d->w_hh == nullptr) return CPC_ERR_ARG;
  a.N = d->N; a.S = d->S;
  a.gx = static_cast<const float*>(d->gx);
  a.w_hh = static_cast<const __nv_bfloat16*>(d->w_hh);
  a.h0 = static_cast<const float*>(d->h0); a.c0 = static_cast<const float*>(d->c0);
  a.h_pad = static_cast<__nv_bfloat16*>(d->h_pad); a.out = static_cast<float*>(d->out);
  a.gates = static_cast<__nv_bfloat16*>(d->gates); a.c_all = static_cast<float*>(d->c_all);
  a.h_n = static_cast<float*>(d->h_n); a.c_n = static_cast<float*>(d->c_n);
  a.dout = static_cast<const float*>(d->dout); a.dgates = static_cast<__nv_bfloat16*>(d->dgates);
  return CPC_OK;
}

static int dispatch_lstm(const LstmArgs& a, int H, bool backward, cudaStream_t stream) {
  switch (H) {
    case 32: return launch_lstm<32>(a, backward, stream);
    case 64: return launch_lstm<64>(a, backward, stream);
    case 128: return launch_lstm<128>(a, backward, stream);
    case 256: return launch_lstm<256>(a, backward, stream);
    case 512: return launch_lstm<512>(a, backward, stream);
    default: return CPC_ERR_UNSUPPORTED;
  }
}

extern "C" int cpc_lstm_fwd(const CpcLstm* d, void* stream_) {
  LstmArgs a;
  int rc = fill_lstm(a, d);
  if (rc != CPC_OK) return rc;
  if (a.gx == nullptr) return CPC_ERR_ARG;
  return dispatch_lstm(a, d->H, false, static_cast<cudaStream_t>(stream_));
}

extern "C" int cpc_lstm_bwd(const CpcLstm* d, void* stream_) {
  LstmArgs a;
  int rc = fill_lstm(a, d);
  if (rc != CPC_OK) return rc;
  if (a.gates == nullptr || a.c_all == nullptr || a.dout == nullptr || a.dgates == nullptr) return CPC_ERR_ARG;
  return dispatch_lstm(a, d->H, true, static_cast<cudaStream_t>(stream_));
}
