// Grouped bf16 GEMM on the 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM, operands staged by
// TMA into 128B-swizzled shared memory).  One kernel serves every dense contraction of the CPC step:
//   * conv1..4 forward / dgrad / wgrad as implicit GEMM over OVERLAPPING-ROW tensor maps of the
//     channels-last activations (row t of the im2col matrix is the contiguous k*C slice starting at
//     padded row t*stride, so no im2col buffer ever exists),
//   * LSTM/GRU input projections and their dX / dW,
//   * the K prediction heads / transformer-AR linears as one grouped launch (blockIdx.z = group),
//   * the all-pairs InfoNCE logits pred_k . z^T and their two backward products.
//
//   D[g][m,n] = alpha * sum_k A_g(m,k) * B_g(n,k)   (+ epilogue)
//
// Either operand can be K-major (reduction index contiguous) or MN-major (m / n index contiguous, i.e. the
// transposed products dY^T.X needed by every weight gradient) -- tcgen05 consumes both natively, only the
// shared-memory descriptor and the TMA box differ.  The reduction index may be split in "slabs"
// (k = slab*k_per_slab + kk) mapped to the outer tensor-map coordinate, which is how a reduction that runs
// over (batch, time) or over (head, feature) is expressed without copying.
//
// CTA = 128 x BN output tile, 6 warps: warp0 TMA producer, warp1 MMA issuer (one elected thread) + TMEM
// allocator, warps 2-5 epilogue (TMEM -> registers -> global).  K pipeline of STAGES smem stages guarded by
// full/empty mbarriers; split-K across blockIdx.z with fp32 red.global.add for the skinny weight-gradient
// shapes.
#include "cpc_internal.h"
#include "ptx.cuh"

namespace cpc {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 B = one swizzle row

struct GemmKernelArgs {
  int M, N;
  int kblocks_per_slab;  // number of BK blocks per slab (ceil)
  int k_per_slab;
  int total_kblocks;  // slabs * kblocks_per_slab
  int k_splits;
  int a_row_off, a_gmul;
  int b_row_off, b_gmul;
  float alpha;
  float* out_f32;
  long long ldd_f32, gs_f32;
  __nv_bfloat16* out_bf16;
  long long ldd_bf16, gs_bf16;
  const float* bias;
  long long gs_bias;
  const __nv_bfloat16* res;
  long long ld_res, gs_res;
  const __nv_bfloat16* gate;
  long long ld_gate, gs_gate;
  float gate_scale;
  const unsigned long long* rng;
  unsigned int rng_site;
  float drop_p;
  int relu;
  int accumulate;
};

template <int BN, int STAGES>
struct GemmSmem {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFF = STAGES * STAGE_BYTES;
  static constexpr int TOTAL = BAR_OFF + (2 * STAGES + 1) * 8 + 16 + 1024;  // + alignment slack
};

template <int BN, int STAGES, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(192, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const GemmKernelArgs p) {
  using L = GemmSmem<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* accum_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int grp = blockIdx.z / p.k_splits;
  const int split = blockIdx.z % p.k_splits;
  // contiguous range of k-blocks for this split
  const int kb_per = (p.total_kblocks + p.k_splits - 1) / p.k_splits;
  const int kb_begin = split * kb_per;
  const int kb_end = min(p.total_kblocks, kb_begin + kb_per);
  const int nkb = kb_end - kb_begin;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(accum_bar, 1);
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, BN);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (nkb > 0) {
    if (warp == 0) {
      if (lane == 0) {
        for (int it = 0; it < nkb; ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&empty_bar[s], ph ^ 1);
          mbar_expect_tx(&full_bar[s], L::STAGE_BYTES);
          const int kb = kb_begin + it;
          const int slab = kb / p.kblocks_per_slab;
          const int kk = (kb % p.kblocks_per_slab) * BK;
          uint8_t* sa = smem + s * L::STAGE_BYTES;
          uint8_t* sb = sa + L::A_BYTES;
          if (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_3d(sa + j * (BK * 128), &tmA, &full_bar[s], m0 + j * 64, kk + p.a_row_off,
                          grp * p.a_gmul + slab);
          } else {
            tma_load_3d(sa, &tmA, &full_bar[s], kk, m0 + p.a_row_off, grp * p.a_gmul + slab);
          }
          if (B_MN) {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_3d(sb + j * (BK * 128), &tmB, &full_bar[s], n0 + j * 64, kk + p.b_row_off,
                          grp * p.b_gmul + slab);
          } else {
            tma_load_3d(sb, &tmB, &full_bar[s], kk, n0 + p.b_row_off, grp * p.b_gmul + slab);
          }
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        constexpr uint32_t idesc = umma_idesc_bf16(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
        for (int it = 0; it < nkb; ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + s * L::STAGE_BYTES);
          const uint32_t sb = sa + L::A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // K-major, 128B swizzle: 8-row groups 1024 B apart, advance 32 B per 16-element k step.
            // MN-major, 128B swizzle: k rows of 128 B (64 mn elements), 8-row groups 1024 B apart,
            // 64-wide mn atoms BK*128 B apart, advance 16 rows = 2048 B per k step.
            const uint64_t ad = A_MN ? umma_smem_desc(sa + k * 2048, BK * 128, 1024, 2)
                                     : umma_smem_desc(sa + k * 32, 16, 1024, 2);
            const uint64_t bd = B_MN ? umma_smem_desc(sb + k * 2048, BK * 128, 1024, 2)
                                     : umma_smem_desc(sb + k * 32, 16, 1024, 2);
            umma_ss(tmem_base, ad, bd, idesc, (it > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);  // frees this smem stage when the MMAs above retire
        }
        umma_commit(accum_bar);  // accumulator complete
      }
    } else {
      // ------------------------------------------------------------------ epilogue (warps 2..5)
      const int q = warp & 3;  // TMEM lane quarter this warp may access
      const int row = m0 + q * 32 + lane;
      mbar_wait(accum_bar, 0);
      tc_fence_after();
      const bool row_ok = row < p.M;
      const long long g = grp;
      unsigned long long seed = 0;
      uint32_t drop_thresh = 0;
      float drop_scale = 1.f;
      if (p.rng != nullptr && p.drop_p > 0.f) {
        seed = p.rng[0] + p.rng[1] * 0x632BE59BD9B4E019ull;
        drop_thresh = static_cast<uint32_t>(fminf(p.drop_p, 0.999999f) * 4294967296.0f);
        drop_scale = 1.f / (1.f - p.drop_p);
      }
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32b_x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + c * 32, v);
        tmem_ld_wait();
        const int nb = n0 + c * 32;
        if (row_ok && nb < p.N) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            const int n = nb + j;
            if (n < p.N) {
              float x[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) x[e] = __uint_as_float(v[j + e]) * p.alpha;
              if (p.bias != nullptr) {
                const float4 b = *reinterpret_cast<const float4*>(p.bias + g * p.gs_bias + n);
                x[0] += b.x; x[1] += b.y; x[2] += b.z; x[3] += b.w;
              }
              if (p.res != nullptr) {
                const uint2 r = *reinterpret_cast<const uint2*>(p.res + g * p.gs_res + (long long)row * p.ld_res + n);
                const __nv_bfloat162 r0 = *reinterpret_cast<const __nv_bfloat162*>(&r.x);
                const __nv_bfloat162 r1 = *reinterpret_cast<const __nv_bfloat162*>(&r.y);
                x[0] += __low2float(r0); x[1] += __high2float(r0);
                x[2] += __low2float(r1); x[3] += __high2float(r1);
              }
              if (p.relu) {
#pragma unroll
                for (int e = 0; e < 4; ++e) x[e] = fmaxf(x[e], 0.f);
              }
              if (p.gate != nullptr) {
                const uint2 r = *reinterpret_cast<const uint2*>(p.gate + g * p.gs_gate + (long long)row * p.ld_gate + n);
                const __nv_bfloat162 r0 = *reinterpret_cast<const __nv_bfloat162*>(&r.x);
                const __nv_bfloat162 r1 = *reinterpret_cast<const __nv_bfloat162*>(&r.y);
                x[0] = __low2float(r0) > 0.f ? x[0] * p.gate_scale : 0.f;
                x[1] = __high2float(r0) > 0.f ? x[1] * p.gate_scale : 0.f;
                x[2] = __low2float(r1) > 0.f ? x[2] * p.gate_scale : 0.f;
                x[3] = __high2float(r1) > 0.f ? x[3] * p.gate_scale : 0.f;
              }
              if (drop_thresh != 0) {
                const unsigned long long base = ((unsigned long long)g * p.M + row) * (unsigned long long)p.N + n;
#pragma unroll
                for (int e = 0; e < 4; ++e)
                  x[e] = rng_u32(seed, p.rng_site, base + e) >= drop_thresh ? x[e] * drop_scale : 0.f;
              }
              if (p.out_f32 != nullptr) {
                float* o = p.out_f32 + g * p.gs_f32 + (long long)row * p.ldd_f32 + n;
                if (p.k_splits > 1) {
                  red_add_v4_f32(o, x[0], x[1], x[2], x[3]);
                } else if (p.accumulate) {
                  float4 old = *reinterpret_cast<float4*>(o);
                  old.x += x[0]; old.y += x[1]; old.z += x[2]; old.w += x[3];
                  *reinterpret_cast<float4*>(o) = old;
                } else {
                  *reinterpret_cast<float4*>(o) = make_float4(x[0], x[1], x[2], x[3]);
                }
              }
              if (p.out_bf16 != nullptr) {
                __nv_bfloat162 o0 = __floats2bfloat162_rn(x[0], x[1]);
                __nv_bfloat162 o1 = __floats2bfloat162_rn(x[2], x[3]);
                uint2 pk;
                pk.x = *reinterpret_cast<uint32_t*>(&o0);
                pk.y = *reinterpret_cast<uint32_t*>(&o1);
                *reinterpret_cast<uint2*>(p.out_bf16 + g * p.gs_bf16 + (long long)row * p.ldd_bf16 + n) = pk;
              }
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, BN);
}

// ----------------------------------------------------------------------------------------------- host
static int encode_operand_map(CUtensorMap* map, const void* ptr, long long d0, long long d1, long long d2,
                              long long s1, long long s2, int mn_major, int tile_mn) {
  PFN_encodeTiled enc = get_encode_tiled();
  if (enc == nullptr) return CPC_ERR_DRIVER;
  cuuint64_t dims[3] = {(cuuint64_t)d0, (cuuint64_t)d1, (cuuint64_t)(d2 > 0 ? d2 : 1)};
  cuuint64_t strides[2] = {(cuuint64_t)s1 * 2, (cuuint64_t)(d2 > 1 ? s2 : (s1 * d1)) * 2};
  if (strides[1] == 0) strides[1] = strides[0];
  cuuint32_t box[3];
  if (mn_major) {
    box[0] = 64; box[1] = BK; box[2] = 1;
  } else {
    box[0] = BK; box[1] = (cuuint32_t)tile_mn; box[2] = 1;
  }
  cuuint32_t estr[3] = {1, 1, 1};
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (strides[0] & 15) || (strides[1] & 15)) return CPC_ERR_ALIGN;
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? CPC_OK : CPC_ERR_TENSORMAP;
}

template <int BN, int STAGES, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmKernelArgs& ka, int groups,
                       cudaStream_t stream) {
  using L = GemmSmem<BN, STAGES>;
  auto kern = gemm_bf16_tcgen05_kernel<BN, STAGES, A_MN, B_MN>;
  static bool attr_done = false;  // per instantiation
  if (!attr_done) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL) != cudaSuccess)
      return CPC_ERR_LAUNCH;
    attr_done = true;
  }
  dim3 grid((ka.M + BM - 1) / BM, (ka.N + BN - 1) / BN, groups * ka.k_splits);
  kern<<<grid, 192, L::TOTAL, stream>>>(ta, tb, ka);
  return cudaGetLastError() == cudaSuccess ? CPC_OK : CPC_ERR_LAUNCH;
}

template <int BN, int STAGES>
static int dispatch_major(const CUtensorMap& ta, const CUtensorMap& tb, const GemmKernelArgs& ka, int groups,
                          int a_mn, int b_mn, cudaStream_t stream) {
  if (a_mn) {
    return b_mn ? launch_gemm<BN, STAGES, true, true>(ta, tb, ka, groups, stream)
                : launch_gemm<BN, STAGES, true, false>(ta, tb, ka, groups, stream);
  }
  return b_mn ? launch_gemm<BN, STAGES, false, true>(ta, tb, ka, groups, stream)
              : launch_gemm<BN, STAGES, false, false>(ta, tb, ka, groups, stream);
}

}  // namespace cpc

using namespace cpc;

extern "C" int cpc_gemm_bf16(const CpcGemmDesc* d, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (d == nullptr || d->M <= 0 || d->N <= 0 || d->K <= 0 || d->groups <= 0) return CPC_ERR_ARG;
  if (d->N % 8 != 0) return CPC_ERR_ARG;
  if (d->out_f32 == nullptr && d->out_bf16 == nullptr) return CPC_ERR_ARG;
  const int k_per_slab = d->k_per_slab > 0 ? d->k_per_slab : d->K;
  if (d->K % k_per_slab != 0) return CPC_ERR_ARG;
  const int slabs = d->K / k_per_slab;
  int k_splits = d->k_splits > 0 ? d->k_splits : 1;
  if (k_splits > 1 && (d->out_f32 == nullptr || d->out_bf16 != nullptr || d->bias || d->res || d->relu ||
                       d->gate || d->drop_p > 0.f))
    return CPC_ERR_ARG;

  GemmKernelArgs ka;
  ka.M = d->M; ka.N = d->N;
  ka.k_per_slab = k_per_slab;
  ka.kblocks_per_slab = (k_per_slab + BK - 1) / BK;
  ka.total_kblocks = ka.kblocks_per_slab * slabs;
  if (k_splits > ka.total_kblocks) k_splits = ka.total_kblocks;
  // make sure no split is empty
  while (k_splits > 1 && ((ka.total_kblocks + k_splits - 1) / k_splits) * (k_splits - 1) >= ka.total_kblocks) --k_splits;
  ka.k_splits = k_splits;
  ka.a_row_off = d->a_row_off; ka.a_gmul = d->a_gmul;
  ka.b_row_off = d->b_row_off; ka.b_gmul = d->b_gmul;
  ka.alpha = d->alpha;
  ka.out_f32 = static_cast<float*>(d->out_f32); ka.ldd_f32 = d->ldd_f32; ka.gs_f32 = d->gs_f32;
  ka.out_bf16 = static_cast<__nv_bfloat16*>(d->out_bf16); ka.ldd_bf16 = d->ldd_bf16; ka.gs_bf16 = d->gs_bf16;
  ka.bias = static_cast<const float*>(d->bias); ka.gs_bias = d->gs_bias;
  ka.res = static_cast<const __nv_bfloat16*>(d->res); ka.ld_res = d->ld_res; ka.gs_res = d->gs_res;
  ka.gate = static_cast<const __nv_bfloat16*>(d->gate); ka.ld_gate = d->ld_gate; ka.gs_gate = d->gs_gate;
  ka.gate_scale = d->gate_scale;
  ka.rng = static_cast<const unsigned long long*>(d->rng); ka.rng_site = d->rng_site; ka.drop_p = d->drop_p;
  ka.relu = d->relu; ka.accumulate = d->accumulate;

  // tile-N selection: widest tile unless that leaves most of the 148 SMs without a CTA
  int bn = 256;
  if (d->N <= 64) bn = 64;
  else if (d->N <= 128) bn = 128;
  else {
    const long long tiles256 = (long long)((d->M + BM - 1) / BM) * ((d->N + 255) / 256) * d->groups * k_splits;
    if (tiles256 < 120) bn = 128;
  }
  if (d->force_bn == 64 || d->force_bn == 128 || d->force_bn == 256) bn = d->force_bn;

  CUtensorMap ta, tb;
  int rc = encode_operand_map(&ta, d->a, d->a_dim0, d->a_dim1, d->a_dim2, d->a_stride1, d->a_stride2, d->a_major, BM);
  if (rc != CPC_OK) return rc;
  rc = encode_operand_map(&tb, d->b, d->b_dim0, d->b_dim1, d->b_dim2, d->b_stride1, d->b_stride2, d->b_major, bn);
  if (rc != CPC_OK) return rc;

  switch (bn) {
    case 64: return dispatch_major<64, 6>(ta, tb, ka, d->groups, d->a_major, d->b_major, stream);
    case 128: return dispatch_major<128, 6>(ta, tb, ka, d->groups, d->a_major, d->b_major, stream);
    default: return dispatch_major<256, 4>(ta, tb, ka, d->groups, d->a_major, d->b_major, stream);
  }
}
