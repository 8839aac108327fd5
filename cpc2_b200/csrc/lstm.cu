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
// sigmoid(x) and its derivative s(1-s) = sigmoid(x) sigmoid(-x); tanh(x) and 1 - tanh^2 = 4 sigmoid(2x) sigmoid(-2x).
// Formed from exp(-|x|) so that the SMALL factor keeps its relative precision in saturation: the backward reads the
// gate PRE-activations (bf16) and recomputes the gates, because a post-activation gate rounded to bf16 has lost
// 1 - s (and 1 - tanh^2) entirely once |x| > 6 -- gradients of saturated units came out as exact zeros.
__device__ __forceinline__ void sigmoid_d(float x, float& s, float& sn) {
  const float e = __expf(-fabsf(x));
  const float a = __fdividef(1.f, 1.f + e), b = e * a;
  s = x >= 0.f ? a : b;
  sn = x >= 0.f ? b : a;  // 1 - s
}
__device__ __forceinline__ void tanh_d(float x, float& th, float& d) {
  const float e = __expf(-2.f * fabsf(x));
  const float a = __fdividef(1.f, 1.f + e);
  th = copysignf((1.f - e) * a, x);
  d = 4.f * e * a * a;  // 1 - th^2
}
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
  __nv_bfloat16* gates;       // [N, S, 4H] bf16 gate PRE-activations i,f,g,o (GRU: r, z, n, W_hn h; saved for backward; optional)
  float* c_all;               // [N, S, H] fp32 cell states (saved for backward; optional)
  float* h_n;                 // [N, H] final hidden (optional)
  float* c_n;                 // [N, H]
  // backward
  const float* dout;          // [N, S, H] fp32 gradient w.r.t. out
  __nv_bfloat16* dgates;      // [N, S, 4H] bf16 gradient w.r.t. the gate pre-activations
};

constexpr int CELL_LSTM = 0, CELL_GRU = 1, CELL_RNN = 2;
constexpr int NB = 8;          // sequences per cluster: one mma n-tile.  Small on purpose -- the recurrence is
                               // latency-bound and the rest of the chip is idle, so more clusters beat wider ones
constexpr int LSTM_THREADS = 256;
// Backward: the per-step inputs (saved gates, c_{t-1}, dout_t) are fetched PF steps ahead into a ring of PF + 1
// buffers.  Two steps (1.5 us) covered an idle memory system only: next to the fused Adam pass of the heads (500 MB, HBM
// saturated for ~90 us) or a gradient all-reduce, load latency grows past that and every step stalled on its inputs --
// the recurrence took 129 us instead of 98 (profiles/r02_timeline_dp_n1.txt).
constexpr int BWD_PF = 6, BWD_RING = BWD_PF + 1;

// Per-step input tile of one CTA: its 4 x 32 gate columns of gx for the cluster's 8 sequences, fp32, fetched two
// steps ahead with cp.async (16-byte chunks: every (sequence, gate) is one contiguous 128-byte segment).
constexpr int GX_PITCH = 132;  // floats per sequence row: bank = 4*b + unit -> conflict-free reads

template <int H, int CELL>
__global__ void __launch_bounds__(LSTM_THREADS, 1) lstm_fwd_kernel(const LstmArgs a) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  constexpr int CL = H / 32, KT = H / 16, HP = H + 8;
  __shared__ __align__(16) __nv_bfloat16 h_s[2][NB][HP];          // h_{t-1} of the whole cluster (double-buffered)
  __shared__ __align__(16) float gx_s[3][NB][GX_PITCH];           // prefetched input projections
  __shared__ __align__(16) __nv_bfloat16 st_h[2][NB][40];         // this CTA's new h slice (bf16): DSMEM push, h_pad
  __shared__ __align__(16) float st_o[2][NB][36];                 // ... fp32 (module output)
  __shared__ __align__(16) float st_c[2][NB][36];                 // new cell state
  __shared__ __align__(16) __nv_bfloat16 st_g[2][NB][136];        // gate pre-activations i|f|g|o
  __shared__ __align__(8) uint64_t bar[2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t cta = cluster_ctarank();
  const int n_base = (blockIdx.x / CL) * NB;
  const int g0 = lane >> 4;          // 0: this thread's fragment rows are gates (i, g); 1: (f, o)
  const int u = (lane >> 2) & 3;     // unit inside the warp's 4
  const int unit_l = warp * 4 + u;   // unit inside the CTA's 32
  const int unit = cta * 32 + unit_l;
  const int H4 = 4 * H;
  const int bsel = (lane & 3) * 2 + g0;   // the sequence this thread finishes after the gate exchange
  const int nidx = n_base + bsel;
  const bool valid = nidx < a.N;

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
  for (int i = tid; i < NB * H; i += LSTM_THREADS) {
    const int b = i / H, k = i % H, n = n_base + b;
    const float v = (a.h0 != nullptr && n < a.N) ? a.h0[(long long)n * H + k] : 0.f;
    h_s[0][b][k] = __float2bfloat16(v);
    if (a.h_pad != nullptr && n < a.N && (k >> 5) == (int)cta) a.h_pad[(long long)n * (a.S + 1) * H + k] = __float2bfloat16(v);
  }
  // bar[q] completes when the CL slices of h (CL x 512 bytes, st.async complete_tx) for a step of parity q have
  // landed in h_s[q]; its single arrival is this CTA arming it (expect_tx) once the previous phase is consumed
  constexpr uint32_t STEP_BYTES = CL * NB * 32 * 2;
  if (tid == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    fence_barrier_init();
    if (a.S > 1) mbar_expect_tx(&bar[1], STEP_BYTES);
  }
  // gx prefetch: thread -> (sequence, gate, 16-byte chunk)
  const int pf_b = tid >> 5, pf_g = (tid >> 3) & 3, pf_ch = tid & 7;
  const bool pf_ok = n_base + pf_b < a.N;
  const float* pf_src = a.gx + ((long long)(n_base + pf_b) * a.S) * H4 + pf_g * H + cta * 32 + pf_ch * 4;
  auto prefetch = [&](int t) {
    if (pf_ok && t < a.S) cp_async_16(&gx_s[t % 3][pf_b][pf_g * 32 + pf_ch * 4], pf_src + (long long)t * H4);
    cp_async_commit();
  };
  prefetch(0);
  prefetch(1);
  cp_async_wait<1>();
  cluster_sync_all();  // barriers initialised cluster-wide, h_s[0] and gx_s[0] visible

  const float* s0p = CELL == CELL_LSTM ? a.c0 : a.h0;  // the fp32 carried state: c (LSTM) or h (GRU, RNN)
  float c_state = (valid && s0p != nullptr) ? s0p[(long long)nidx * H + unit] : 0.f;

  for (int t = 0; t < a.S; ++t) {
    const int p = t & 1;
    prefetch(t + 2);
    if (t > 0) mbar_wait(&bar[p], ((t - 1) >> 1) & 1);
    if (tid == 0 && t + 2 < a.S) mbar_expect_tx(&bar[p], STEP_BYTES);  // arm for step t+2 (h_{t+1} lands here)
    // four independent accumulator chains: the step is latency-bound and a chain of KT dependent HMMAs is its
    // longest serial piece
    constexpr int CH = KT >= 8 ? 4 : 2;
    float acc[CH][4];
#pragma unroll
    for (int i = 0; i < CH; ++i)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[i][e] = 0.f;
    // B fragments: h_s[p][seq][k], two k-tiles per ldmatrix.x4
    const uint32_t hrow = smem_u32(&h_s[p][lane & 7][(lane >> 3) * 8]);
#pragma unroll
    for (int kt2 = 0; kt2 < KT / 2; ++kt2) {
      uint32_t bb[4];
      ldsm_x4(bb, hrow + kt2 * 64);
      mma_bf16_16816(acc[(2 * kt2) % CH], afr[2 * kt2], bb[0], bb[1]);
      mma_bf16_16816(acc[(2 * kt2 + 1) % CH], afr[2 * kt2 + 1], bb[2], bb[3]);
    }
    float c4[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      c4[e] = acc[0][e] + acc[1][e];
      if (CH == 4) c4[e] += acc[2][e] + acc[3][e];
    }
    // rows (lane/4) and (lane/4 + 8) are gates g0 and g0+2 of unit u for sequences 2*(lane%4), +1; the partner
    // lane (xor 16) holds the other two gates.  After the exchange this thread owns sequence 2*(lane%4) + g0.
    const float s0 = g0 ? c4[0] : c4[1];
    const float s1 = g0 ? c4[2] : c4[3];
    const float r0 = __shfl_xor_sync(0xffffffffu, s0, 16);
    const float r1 = __shfl_xor_sync(0xffffffffu, s1, 16);
    const float* gxr = &gx_s[t % 3][bsel][unit_l];
    const float pi = (g0 ? r0 : c4[0]) + gxr[0];
    const float pf = (g0 ? c4[1] : r0) + gxr[32];
    const float pg = (g0 ? r1 : c4[2]) + gxr[64];
    const float po = (g0 ? c4[3] : r1) + gxr[96];
    float ig, fg, gg, og, cn, hn;
    if constexpr (CELL == CELL_LSTM) {
      ig = sigmoid_f(pi); fg = sigmoid_f(pf); gg = tanh_f(pg); og = sigmoid_f(po);
      cn = valid ? fg * c_state + ig * gg : 0.f;
      hn = valid ? og * tanh_f(cn) : 0.f;
    } else if constexpr (CELL == CELL_GRU) {
      // slots (r, z, n_x, hn): pg = W_in x + b_in (its W_hh slot is zero), po = W_hn h + b_hn
      ig = sigmoid_f(pi); fg = sigmoid_f(pf); og = po; gg = tanh_f(pg + ig * po);
      hn = valid ? gg + fg * (c_state - gg) : 0.f;   // (1-z) n + z h_{t-1}; c_state carries h in fp32
      cn = hn;
    } else {
      // tanh RNN: slot 0 only
      ig = tanh_f(pi); fg = 0.f; gg = 0.f; og = 0.f;
      hn = valid ? ig : 0.f;
      cn = hn;
    }
    c_state = cn;
    st_h[p][bsel][unit_l] = __float2bfloat16(hn);
    st_o[p][bsel][unit_l] = hn;
    st_c[p][bsel][unit_l] = cn;
    st_g[p][bsel][unit_l] = __float2bfloat16(pi);
    st_g[p][bsel][32 + unit_l] = __float2bfloat16(pf);
    st_g[p][bsel][64 + unit_l] = __float2bfloat16(CELL == CELL_GRU ? pg + ig * po : pg);
    st_g[p][bsel][96 + unit_l] = __float2bfloat16(po);
    cp_async_wait<1>();  // gx of step t+1 has landed (only the newest group may be in flight)
    __syncthreads();
    if (t + 1 < a.S) {
      // warp d pushes this CTA's [8 x 32] bf16 slice of h_t into buffer p^1 of CTA d; the bytes count on d's barrier
      const int b = lane >> 2, part = lane & 3;
      const uint4 v = *reinterpret_cast<const uint4*>(&st_h[p][b][part * 8]);
      const uint32_t dst = smem_u32(&h_s[p ^ 1][b][cta * 32 + part * 8]);
      const uint32_t dbar = smem_u32(&bar[p ^ 1]);
      for (int d = warp; d < CL; d += LSTM_THREADS / 32) st_async_v4(mapa(dst, d), v.x, v.y, v.z, v.w, mapa(dbar, d));
    }
    // coalesced global writes of everything the backward (and the caller) needs, off the critical path
    {
      const int b = (tid & 63) >> 3, ch = tid & 7, n = n_base + b;
      const long long row = (long long)n * a.S + t;
      if (tid < 64) {
        if (a.out != nullptr && n < a.N)
          *reinterpret_cast<float4*>(a.out + row * H + cta * 32 + ch * 4) = *reinterpret_cast<const float4*>(&st_o[p][b][ch * 4]);
      } else if (tid < 128) {
        if (a.c_all != nullptr && n < a.N)
          *reinterpret_cast<float4*>(a.c_all + row * H + cta * 32 + ch * 4) = *reinterpret_cast<const float4*>(&st_c[p][b][ch * 4]);
      } else if (a.gates != nullptr) {
        // 128 threads: (sequence, gate, 16-byte chunk of the 64-byte segment)
        const int q = tid - 128, gb = q >> 4, gg_ = (q >> 2) & 3, gc = q & 3;
        if (n_base + gb < a.N)
          *reinterpret_cast<uint4*>(a.gates + ((long long)(n_base + gb) * a.S + t) * H4 + gg_ * H + cta * 32 + gc * 8) =
              *reinterpret_cast<const uint4*>(&st_g[p][gb][gg_ * 32 + gc * 8]);
      }
      if (tid < 32 && a.h_pad != nullptr) {
        const int hb = tid >> 2, hc = tid & 3;
        if (n_base + hb < a.N)
          *reinterpret_cast<uint4*>(a.h_pad + ((long long)(n_base + hb) * (a.S + 1) + t + 1) * H + cta * 32 + hc * 8) =
              *reinterpret_cast<const uint4*>(&st_h[p][hb][hc * 8]);
      }
      if (t == a.S - 1 && tid < 64 && n < a.N) {
        if (a.h_n != nullptr)
          *reinterpret_cast<float4*>(a.h_n + (long long)n * H + cta * 32 + ch * 4) = *reinterpret_cast<const float4*>(&st_o[p][b][ch * 4]);
        if (a.c_n != nullptr)
          *reinterpret_cast<float4*>(a.c_n + (long long)n * H + cta * 32 + ch * 4) = *reinterpret_cast<const float4*>(&st_c[p][b][ch * 4]);
      }
    }
  }
  cluster_sync_all();  // nobody exits while peers may still target its shared memory
}

// Backward recurrence.  CTA c owns hidden units [32c, 32c+32): it turns dh into the gate gradients of those units,
// multiplies them with its 128 rows of W_hh (register-resident A fragments of W_hh^T, K-split over gate rows) and
// sends each 32-unit block of the partial dh_{t-1} to the CTA that owns it (fp32, st.shared::cluster.v2 straight
// from the accumulator fragments; warp w's two m-tiles are exactly the block of CTA w when H = 256).
template <int H, int CELL>
__global__ void __launch_bounds__(LSTM_THREADS, 1) lstm_bwd_kernel(const LstmArgs a) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  constexpr int CL = H / 32;
  constexpr int MT = H / 16;                      // output (h) m-tiles
  constexpr int MTW = (MT + 7) / 8;               // m-tiles per warp
  constexpr int ACTIVE = MT / MTW;                // warps that own m-tiles
  constexpr int GP = 136;                         // bf16 pitch of the [seq][gate*32+unit] tiles
  constexpr int CP = 36;                          // fp32 pitch of the [seq][unit] tiles
  extern __shared__ __align__(16) uint8_t smem_dyn[];
  float* P = reinterpret_cast<float*>(smem_dyn);                                     // [2][CL][32 units][NB] partial dh
  __nv_bfloat16* in_g = reinterpret_cast<__nv_bfloat16*>(P + 2 * CL * 32 * NB);      // [RING][NB][GP] saved gates
  float* in_c = reinterpret_cast<float*>(in_g + BWD_RING * NB * GP);                 // [RING][NB][CP] c_{t-1}
  float* in_d = in_c + BWD_RING * NB * CP;                                           // [RING][NB][CP] dout_t
  __nv_bfloat16* dg_s = reinterpret_cast<__nv_bfloat16*>(in_d + BWD_RING * NB * CP); // [2][NB][GP] gate gradients
  __shared__ __align__(8) uint64_t bar[2];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t cta = cluster_ctarank();
  const int n_base = (blockIdx.x / CL) * NB;
  const int u = (lane >> 2) & 3;
  const int unit_l = warp * 4 + u;
  const int unit = cta * 32 + unit_l;
  const int bsel = (lane & 3) * 2 + (lane >> 4);
  const int nidx = n_base + bsel;
  const bool valid = nidx < a.N;
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
  // bar[q]: the CL partial dh tiles (CL x 32 units x 8 sequences fp32, st.async complete_tx) of a step of parity q
  constexpr uint32_t STEP_BYTES = CL * 32 * NB * 4;
  if (tid == 0) {
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    fence_barrier_init();
    if (a.S > 1) mbar_expect_tx(&bar[1], STEP_BYTES);
  }
  const float* s0p = CELL == CELL_LSTM ? a.c0 : a.h0;  // state before step 0: c0 (LSTM) or h0 (GRU / RNN, whose c_all holds h)
  // per-step input prefetch (cp.async, BWD_PF steps ahead): 128 threads fetch the gates tile, 64 c_{t-1}, 64 dout_t
  auto prefetch = [&](int sidx) {
    const int t = a.S - 1 - sidx;
    if (sidx < a.S) {
      const int buf = sidx % BWD_RING;
      if (tid < 128) {
        const int b = tid >> 4, g = (tid >> 2) & 3, ch = tid & 3;
        if (n_base + b < a.N)
          cp_async_16(in_g + (buf * NB + b) * GP + g * 32 + ch * 8,
                      a.gates + ((long long)(n_base + b) * a.S + t) * H4 + g * H + cta * 32 + ch * 8);
      } else if (tid < 192) {
        const int b = (tid - 128) >> 3, ch = tid & 7;
        if (n_base + b < a.N) {
          const float* src = t > 0 ? a.c_all + ((long long)(n_base + b) * a.S + t - 1) * H
                                   : (s0p != nullptr ? s0p + (long long)(n_base + b) * H : nullptr);
          float* dst = in_c + (buf * NB + b) * CP + ch * 4;
          if (src != nullptr) cp_async_16(dst, src + cta * 32 + ch * 4);
          else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      } else {
        const int b = (tid - 192) >> 3, ch = tid & 7;
        if (n_base + b < a.N)
          cp_async_16(in_d + (buf * NB + b) * CP + ch * 4, a.dout + ((long long)(n_base + b) * a.S + t) * H + cta * 32 + ch * 4);
      }
    }
    cp_async_commit();
  };
#pragma unroll
  for (int i = 0; i < BWD_PF; ++i) prefetch(i);
  cp_async_wait<BWD_PF - 1>();
  cluster_sync_all();

  float dc_state = 0.f;
  float ct = valid ? a.c_all[((long long)nidx * a.S + a.S - 1) * H + unit] : 0.f;  // c_{S-1}; later steps reuse c_{t-1}

  for (int sidx = 0; sidx < a.S; ++sidx) {
    const int t = a.S - 1 - sidx;
    const int pp = sidx & 1, buf = sidx % BWD_RING;
    prefetch(sidx + BWD_PF);
    // everything that does not depend on the recurrence first
    const __nv_bfloat16* gp = in_g + (buf * NB + bsel) * GP + unit_l;
    const float xi = __bfloat162float(gp[0]), xf = __bfloat162float(gp[32]), xg = __bfloat162float(gp[64]),
                xo = __bfloat162float(gp[96]);  // pre-activations
    const float cprev = in_c[(buf * NB + bsel) * CP + unit_l];
    float dh = in_d[(buf * NB + bsel) * CP + unit_l];
    float gi, gin, gf, gfn, gg, ggd, go, gon, tc, tcd;  // gate, complement (sigmoid) / 1 - gate^2 (tanh)
    if constexpr (CELL == CELL_LSTM) {
      sigmoid_d(xi, gi, gin); sigmoid_d(xf, gf, gfn); tanh_d(xg, gg, ggd); sigmoid_d(xo, go, gon);
      tanh_d(ct, tc, tcd);
    } else if constexpr (CELL == CELL_GRU) {
      sigmoid_d(xi, gi, gin); sigmoid_d(xf, gf, gfn); tanh_d(xg, gg, ggd);
      go = xo; gon = 0.f; tc = 0.f; tcd = 0.f;
    } else {
      tanh_d(xi, gi, gin);  // gin = 1 - h_t^2
      gf = gfn = gg = ggd = go = gon = tc = tcd = 0.f;
    }
    // pin the gate recomputation (ten MUFU ops) BEFORE the wait for the other CTAs' partial dh: volatile asm is not
    // reordered against the volatile try_wait loop, so these values exist when the barrier flips and the transcendental
    // latency hides behind the DSMEM exchange instead of extending the dependent chain of the step
    asm volatile("" ::"f"(gi), "f"(gin), "f"(gf), "f"(gfn), "f"(gg), "f"(ggd), "f"(go), "f"(gon), "f"(tc), "f"(tcd), "f"(dh));
    if (sidx > 0) {
      mbar_wait(&bar[pp], ((sidx - 1) >> 1) & 1);
      const float* pr = P + ((pp * CL) * 32 + unit_l) * NB + bsel;
      float s = 0.f;
#pragma unroll
      for (int src = 0; src < CL; ++src) s += pr[src * 32 * NB];
      dh += s;
    }
    if (tid == 0 && sidx + 2 < a.S) mbar_expect_tx(&bar[pp], STEP_BYTES);  // arm for step sidx+2
    float d_i = 0.f, d_f = 0.f, d_g = 0.f, d_o = 0.f;
    if (valid) {
      if constexpr (CELL == CELL_LSTM) {
        d_o = dh * tc * go * gon;
        const float dc = dh * go * tcd + dc_state;
        d_i = dc * gg * gi * gin;
        d_f = dc * cprev * gf * gfn;
        d_g = dc * gi * ggd;
        dc_state = dc * gf;
      } else if constexpr (CELL == CELL_GRU) {
        // saved slots: r, z, n pre-activations and hn = W_hn h + b_hn = go; cprev = h_{t-1}
        dh += dc_state;                                   // the z-gated path h_{t-1} -> h_t stays in this thread
        d_g = dh * gfn * ggd;                             // d n_pre (input side)
        d_f = dh * (cprev - gg) * gf * gfn;               // d z_pre
        d_i = d_g * go * gi * gin;                        // d r_pre
        d_o = d_g * gi;                                   // d (W_hn h + b_hn) (hidden side of the n gate)
        dc_state = dh * gf;
      } else {
        d_i = dh * gin;                                   // saved slot 0 = pre-activation of h_t
      }
    }
    ct = cprev;
    __nv_bfloat16* ds = dg_s + (pp * NB + bsel) * GP + unit_l;
    ds[0] = __float2bfloat16(d_i);
    ds[32] = __float2bfloat16(d_f);
    ds[64] = __float2bfloat16(d_g);
    ds[96] = __float2bfloat16(d_o);
    cp_async_wait<BWD_PF - 1>();  // the inputs of step sidx + 1 have landed
    __syncthreads();
    if (t > 0 && warp < ACTIVE) {
      // partial dh_{t-1}[j][b] = sum_{r in my 128 gate rows} W_hh[rg(r)][j] * dgate[b][r]
      float acc[MTW][4], acc2[MTW][4];  // two chains per m-tile: halves the dependent-HMMA latency of the step
#pragma unroll
      for (int mt = 0; mt < MTW; ++mt)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[mt][e] = acc2[mt][e] = 0.f;
      const uint32_t brow = smem_u32(dg_s + (pp * NB + (lane & 7)) * GP + (lane >> 3) * 8);
#pragma unroll
      for (int kt2 = 0; kt2 < 4; ++kt2) {
        uint32_t bb[4];
        ldsm_x4(bb, brow + kt2 * 64);
#pragma unroll
        for (int mt = 0; mt < MTW; ++mt) {
          mma_bf16_16816(acc[mt], afr[mt][2 * kt2], bb[0], bb[1]);
          mma_bf16_16816(acc2[mt], afr[mt][2 * kt2 + 1], bb[2], bb[3]);
        }
      }
#pragma unroll
      for (int mt = 0; mt < MTW; ++mt)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[mt][e] += acc2[mt][e];
      // accumulator (row j = m-tile*16 + lane/4 (+8), columns = sequences 2*(lane%4), +1) -> owner CTA of unit j
#pragma unroll
      for (int mt = 0; mt < MTW; ++mt) {
        const int m = warp * MTW + mt;          // m-tile index: units [16m, 16m+16)
        const int d = m >> 1;                   // owner CTA
        const int jl = (m & 1) * 16 + (lane >> 2);
        const uint32_t dst = smem_u32(P + (((pp ^ 1) * CL + cta) * 32 + jl) * NB + (lane & 3) * 2);
        const uint32_t dbar = mapa(smem_u32(&bar[pp ^ 1]), d);
        st_async_v2_f32(mapa(dst, d), acc[mt][0], acc[mt][1], dbar);
        st_async_v2_f32(mapa(dst + 8 * NB * 4, d), acc[mt][2], acc[mt][3], dbar);
      }
    }
    // coalesced write of this step's gate gradients (operand of the dW / dx GEMMs)
    if (tid < 128) {
      const int b = tid >> 4, g = (tid >> 2) & 3, ch = tid & 3;
      if (n_base + b < a.N)
        *reinterpret_cast<uint4*>(a.dgates + ((long long)(n_base + b) * a.S + t) * H4 + g * H + cta * 32 + ch * 8) =
            *reinterpret_cast<const uint4*>(dg_s + (pp * NB + b) * GP + g * 32 + ch * 8);
    }
  }
  cluster_sync_all();
}

template <int H, int CELL>
static int launch_lstm(const LstmArgs& a, bool backward, cudaStream_t stream) {
  constexpr int CL = H / 32;
  const int clusters = (a.N + NB - 1) / NB;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(CL * clusters);
  cfg.blockDim = dim3(LSTM_THREADS);
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // PDL (cpc_internal.h)
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  cudaError_t e;
  if (!backward) {
    auto kern = lstm_fwd_kernel<H, CELL>;
    if (CL > 8) cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    cfg.dynamicSmemBytes = 0;
    e = cudaLaunchKernelEx(&cfg, kern, a);
  } else {
    auto kern = lstm_bwd_kernel<H, CELL>;
    const size_t smem = (size_t)2 * CL * 32 * NB * 4 + (size_t)BWD_RING * NB * 136 * 2 + (size_t)2 * BWD_RING * NB * 36 * 4 + (size_t)2 * NB * 136 * 2;
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

template <int CELL>
static int dispatch_h(const LstmArgs& a, int H, bool backward, cudaStream_t stream) {
  switch (H) {
    case 32: return launch_lstm<32, CELL>(a, backward, stream);
    case 64: return launch_lstm<64, CELL>(a, backward, stream);
    case 128: return launch_lstm<128, CELL>(a, backward, stream);
    case 256: return launch_lstm<256, CELL>(a, backward, stream);
    case 512: return launch_lstm<512, CELL>(a, backward, stream);
    default: return CPC_ERR_UNSUPPORTED;
  }
}

static int dispatch_lstm(const LstmArgs& a, int H, int cell, bool backward, cudaStream_t stream) {
  switch (cell) {
    case CELL_LSTM: return dispatch_h<CELL_LSTM>(a, H, backward, stream);
    case CELL_GRU: return dispatch_h<CELL_GRU>(a, H, backward, stream);
    case CELL_RNN: return dispatch_h<CELL_RNN>(a, H, backward, stream);
    default: return CPC_ERR_UNSUPPORTED;
  }
}

extern "C" int cpc_lstm_fwd(const CpcLstm* d, void* stream_) {
  LstmArgs a;
  int rc = fill_lstm(a, d);
  if (rc != CPC_OK) return rc;
  if (a.gx == nullptr) return CPC_ERR_ARG;
  return dispatch_lstm(a, d->H, d->cell, false, static_cast<cudaStream_t>(stream_));
}

extern "C" int cpc_lstm_bwd(const CpcLstm* d, void* stream_) {
  LstmArgs a;
  int rc = fill_lstm(a, d);
  if (rc != CPC_OK) return rc;
  if (a.gates == nullptr || a.c_all == nullptr || a.dout == nullptr || a.dgates == nullptr) return CPC_ERR_ARG;
  return dispatch_lstm(a, d->H, d->cell, true, static_cast<cudaStream_t>(stream_));
}
