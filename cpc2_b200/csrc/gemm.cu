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
// PERSISTENT CTAs (grid = min(tiles, SMs)) walk 128 x BN output tiles; 10 warps: warp0 TMA producer, warp1 MMA
// issuer (one elected thread) + TMEM allocator, warps 2-9 epilogue.  K pipeline of STAGES smem stages guarded by
// full/empty mbarriers; the accumulator is DOUBLE-BUFFERED in TMEM (2*BN columns) so the epilogue of one tile
// overlaps the main loop of the next.  The epilogue reads TMEM with tcgen05.ld.16x256b (mma C-fragment layout:
// four lanes cover 32 contiguous bytes of an output row) so every global access (outputs, residual, gate,
// split-K reductions) is sector-coalesced: the thread-per-row 32x32b layout costs 32 sectors per store and
// was 5-10x slower than the MMAs of the K=256 shapes that dominate this workload.  Split-K uses fp32
// red.global.add.v2.
#include <cstdlib>

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
  int tiles_m, tiles_n, total_tiles;
  int a_row_off, a_gmul, a_smul;
  int b_row_off, b_gmul, b_smul;  // outer coordinate = grp * gmul + slab * smul
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
  // fused row normalisation (EPI_NORM), see CpcGemmDesc
  const float* norm_gamma;
  const float* norm_beta;
  long long gs_norm_param;
  __nv_bfloat16* norm_bf16;
  long long ld_norm_bf16, gs_norm_bf16;
  float* norm_f32;
  long long ld_norm_f32, gs_norm_f32;
  float* norm_mean;
  float* norm_rstd;
  long long gs_norm_stat;
  float norm_eps;
  int norm_unbiased, norm_halo;
  int vec16, vec16n;  // out_bf16 / norm_bf16 rows take 16-byte stores (base, row stride and group stride 16-byte aligned)
  int cl_n;           // EPI_NORM_CL: CTAs per cluster = column chunks per row (tile index: chunk fastest); 0 otherwise
};

constexpr int EPI_WARPS = 8;                 // two warps per TMEM lane quarter, each owns half of the tile's columns
constexpr int GEMM_THREADS = (2 + EPI_WARPS) * 32;

template <int BN, int STAGES>
struct GemmSmem {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFF = STAGES * STAGE_BYTES;
  static constexpr int STAT_OFF = BAR_OFF + (2 * STAGES + 4) * 8 + 16;  // EPI_NORM: [tile parity][column half][BM] float2
  static constexpr int STAT_BYTES = 2 * 2 * BM * 8;
  static constexpr int XCL_MAX = 4;                                  // EPI_NORM_CL: cluster size
  static constexpr int XSTAT_OFF = STAT_OFF + STAT_BYTES;            // [2 * XCL_MAX][BM] float2 partial row statistics
  static constexpr int XSTAT_BYTES = BN <= 128 ? 2 * XCL_MAX * BM * 8 : 0;
  static constexpr int XBAR_OFF = XSTAT_OFF + XSTAT_BYTES;           // the exchange mbarrier
  static constexpr int TOTAL = XBAR_OFF + 16 + 1024;  // + alignment slack
  static constexpr int TMEM_COLS = 2 * BN;  // double-buffered accumulator (128 / 256 / 512 columns)
};

struct TileCoord {
  int m0, n0, grp, kb_begin, nkb;
};
__device__ __forceinline__ TileCoord decode_tile(const GemmKernelArgs& p, int tile, int bn) {
  TileCoord t;
  int mi, ni, z;
  if (p.cl_n > 0) {  // the column chunks of a row are consecutive tiles = the CTAs of one cluster
    ni = tile % p.tiles_n;
    const int r = tile / p.tiles_n;
    mi = r % p.tiles_m;
    z = r / p.tiles_m;
  } else {
    mi = tile % p.tiles_m;
    const int r = tile / p.tiles_m;
    ni = r % p.tiles_n;
    z = r / p.tiles_n;
  }
  t.m0 = mi * BM;
  t.n0 = ni * bn;
  t.grp = z / p.k_splits;
  const int split = z % p.k_splits;
  const int kb_per = (p.total_kblocks + p.k_splits - 1) / p.k_splits;
  t.kb_begin = split * kb_per;
  t.nkb = min(p.total_kblocks, t.kb_begin + kb_per) - t.kb_begin;
  return t;
}

// Phase timestamps of the first and last tile of every CTA (tools/gemm_phase_probe.py builds a second library with
// -DCPC_GEMM_TIMING; the shipped library compiles none of this).
#ifdef CPC_GEMM_TIMING
__device__ long long* g_gemm_timing = nullptr;
#define GEMM_TSTAMP(slot) do { if (g_gemm_timing != nullptr) g_gemm_timing[blockIdx.x * 16 + (slot)] = clock64(); } while (0)
#else
#define GEMM_TSTAMP(slot) do { } while (0)
#endif

}  // namespace cpc
#include "gemm_epilogue.cuh"
namespace cpc {

// PERSISTENT kernel: grid = min(tiles, SMs); CTA walks tiles blockIdx.x, +gridDim.x, ...  The accumulator is
// double-buffered in TMEM so the epilogue of tile i (TMEM -> registers -> smem transpose -> coalesced global)
// overlaps the TMA/MMA main loop of tile i+1.
template <int BN, int STAGES, bool A_MN, bool B_MN, int MODE>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const GemmKernelArgs p) {
  using L = GemmSmem<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* acc_full = empty_bar + STAGES;  // [2]
  uint64_t* acc_empty = acc_full + 2;       // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (threadIdx.x == 0) GEMM_TSTAMP(0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], EPI_WARPS);
    }
    if constexpr (MODE == EPI_NORM_CL) {
      uint64_t* xbar = reinterpret_cast<uint64_t*>(smem + L::XBAR_OFF);
      mbar_init(xbar, 1);
      fence_barrier_init();
      mbar_expect_tx(xbar, static_cast<uint32_t>((p.cl_n - 1) * 2 * BM * 8));  // the peers' partial row statistics
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, L::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if constexpr (MODE == EPI_NORM_CL) cluster_sync_all();  // every CTA's exchange barrier is armed before anyone sends
  const uint32_t tmem_base = *tmem_slot;
  // PDL: everything above (tensor-map fetch, barrier init, TMEM allocation) overlapped the tail of the preceding grid;
  // nothing below reads or writes global memory before that grid has completed (cpc_internal.h)
  pdl_sync();
  if (threadIdx.x == 0) GEMM_TSTAMP(1);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const TileCoord tc = decode_tile(p, tile, BN);
        for (int i = 0; i < tc.nkb; ++i, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&empty_bar[s], ph ^ 1);
          mbar_expect_tx(&full_bar[s], L::STAGE_BYTES);
          const int kb = tc.kb_begin + i;
          const int slab = kb / p.kblocks_per_slab;
          const int kk = (kb % p.kblocks_per_slab) * BK;
          uint8_t* sa = smem + s * L::STAGE_BYTES;
          uint8_t* sb = sa + L::A_BYTES;
          if (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j)
              tma_load_3d(sa + j * (BK * 128), &tmA, &full_bar[s], tc.m0 + j * 64, kk + p.a_row_off,
                          tc.grp * p.a_gmul + slab * p.a_smul);
          } else {
            tma_load_3d(sa, &tmA, &full_bar[s], kk, tc.m0 + p.a_row_off, tc.grp * p.a_gmul + slab * p.a_smul);
          }
          if (B_MN) {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_3d(sb + j * (BK * 128), &tmB, &full_bar[s], tc.n0 + j * 64, kk + p.b_row_off,
                          tc.grp * p.b_gmul + slab * p.b_smul);
          } else {
            tma_load_3d(sb, &tmB, &full_bar[s], kk, tc.n0 + p.b_row_off, tc.grp * p.b_gmul + slab * p.b_smul);
          }
          if (it == 0) GEMM_TSTAMP(2);
        }
      }
      GEMM_TSTAMP(3);  // every load of this CTA issued
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      int it = 0, lt = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++lt) {
        const TileCoord tc = decode_tile(p, tile, BN);
        const int buf = lt & 1;
        mbar_wait(&acc_empty[buf], ((lt >> 1) & 1) ^ 1);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t dcol = tmem_base + buf * BN;
        for (int i = 0; i < tc.nkb; ++i, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          if (it == 0) GEMM_TSTAMP(4);
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
            umma_ss(dcol, ad, bd, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);  // frees this smem stage when the MMAs above retire
        }
        umma_commit(&acc_full[buf]);  // accumulator of this tile complete
        if (lt == 0) GEMM_TSTAMP(5);
      }
      GEMM_TSTAMP(6);  // every MMA of this CTA issued
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9), gemm_epilogue.cuh
    const int e = warp - 2;
    const int q = warp & 3;   // TMEM lane quarter this warp may access
    const int half = e >> 2;  // which half of the tile's columns
    EpiRng rng = {0u, 0u, 0u, 1.f};
    if (MODE == EPI_DROP) {
      const unsigned long long seed = p.rng[0] + p.rng[1] * 0x632BE59BD9B4E019ull;
      rng.seed_lo = static_cast<uint32_t>(seed);
      rng.seed_hi = static_cast<uint32_t>(seed >> 32) + p.rng_site * 0x85EBCA77u;
      rng.thresh = static_cast<uint32_t>(fminf(p.drop_p, 0.9999f) * 65536.0f);
      rng.scale = 1.f / (1.f - p.drop_p);
    }
    int lt = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++lt) {
      const TileCoord tc = decode_tile(p, tile, BN);
      const int buf = lt & 1;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN + half * (BN / 2);
      if constexpr (MODE == EPI_NORM)
        epilogue_tile_norm<BN>(p, tc, taddr, q, half, lane, &acc_full[buf], (lt >> 1) & 1,
                               reinterpret_cast<float2*>(smem + L::STAT_OFF) + (lt & 1) * 2 * BM);
      else if constexpr (MODE == EPI_NORM_CL)
        epilogue_tile_norm_cl<BN>(p, tc, taddr, q, half, lane, &acc_full[buf], (lt >> 1) & 1,
                                  NormClSmem{reinterpret_cast<float2*>(smem + L::XSTAT_OFF),
                                             reinterpret_cast<uint64_t*>(smem + L::XBAR_OFF), p.cl_n});
      else
        epilogue_tile<MODE, BN>(p, tc, taddr, q, half, lane, rng, &acc_full[buf], (lt >> 1) & 1);
      // this warp's TMEM reads of the buffer are complete (tcgen05.wait::ld inside): hand it back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
      if (warp == 2 && lane == 0 && lt == 0) GEMM_TSTAMP(8);
    }
    if (warp == 2 && lane == 0) GEMM_TSTAMP(9);
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) GEMM_TSTAMP(10);
  if (warp == 1) tmem_dealloc(tmem_base, L::TMEM_COLS);
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

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0)
      n = v;
    else
      n = 148;
  }
  return n;
}

template <int BN, int STAGES, bool A_MN, bool B_MN, int MODE>
static int launch_gemm_mode(const CUtensorMap& ta, const CUtensorMap& tb, GemmKernelArgs ka, int groups,
                            cudaStream_t stream) {
  using L = GemmSmem<BN, STAGES>;
  auto kern = gemm_bf16_tcgen05_kernel<BN, STAGES, A_MN, B_MN, MODE>;
  static bool attr_done = false;  // per instantiation
  if (!attr_done) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL) != cudaSuccess)
      return CPC_ERR_LAUNCH;
    attr_done = true;
  }
  ka.tiles_m = (ka.M + BM - 1) / BM;
  ka.tiles_n = (ka.N + BN - 1) / BN;
  ka.total_tiles = ka.tiles_m * ka.tiles_n * groups * ka.k_splits;
  if constexpr (MODE == EPI_NORM_CL) {
    // one tile per CTA, the tiles_n column chunks of a row form one cluster (their exchange barrier is single-use)
    if (ka.tiles_n != ka.cl_n || ka.cl_n < 2 || ka.cl_n > L::XCL_MAX || ka.k_splits != 1) return CPC_ERR_UNSUPPORTED;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ka.total_tiles);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = L::TOTAL;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = ka.cl_n;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    return cudaLaunchKernelEx(&cfg, kern, ta, tb, ka) == cudaSuccess ? CPC_OK : CPC_ERR_LAUNCH;
  }
  int cap = sm_count();
  if constexpr (MODE == EPI_SPLITK || MODE == EPI_ACCUM) {
    // weight-gradient products run on side streams next to the dependent chain; their persistent CTAs hold an SM each
    // (200 KB of shared memory: nothing of the chain co-resides), so they may be capped (CPC_WGRAD_CTAS) to leave SMs free
    static const int wcap = [] {
      const char* e = getenv("CPC_WGRAD_CTAS");
      return e != nullptr && atoi(e) > 0 ? atoi(e) : 0;
    }();
    if (wcap > 0 && wcap < cap) cap = wcap;
  }
  const int grid = ka.total_tiles < cap ? ka.total_tiles : cap;
  launch_k(kern, dim3(grid), dim3(GEMM_THREADS), L::TOTAL, stream, ta, tb, ka);
  return cudaGetLastError() == cudaSuccess ? CPC_OK : CPC_ERR_LAUNCH;
}

template <int BN, int STAGES, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmKernelArgs& ka, int groups,
                       cudaStream_t stream) {
  // epilogue mode (exclusive; the entry point has rejected combinations no mode covers)
  if (ka.norm_gamma != nullptr) {
    if constexpr (BN == 256 && !A_MN && !B_MN) return launch_gemm_mode<BN, STAGES, A_MN, B_MN, EPI_NORM>(ta, tb, ka, groups, stream);
    else if constexpr (!A_MN && !B_MN) return launch_gemm_mode<BN, STAGES, A_MN, B_MN, EPI_NORM_CL>(ta, tb, ka, groups, stream);
    else return CPC_ERR_UNSUPPORTED;
  }
  if (ka.k_splits > 1) return launch_gemm_mode<BN, STAGES, A_MN, B_MN, EPI_SPLITK>(ta, tb, ka, groups, stream);
  if (ka.accumulate) return launch_gemm_mode<BN, STAGES, A_MN, B_MN, EPI_ACCUM>(ta, tb, ka, groups, stream);
  if (ka.rng != nullptr && ka.drop_p > 0.f) return launch_gemm_mode<BN, STAGES, A_MN, B_MN, EPI_DROP>(ta, tb, ka, groups, stream);
  if (ka.gate != nullptr) return launch_gemm_mode<BN, STAGES, A_MN, B_MN, EPI_GATE>(ta, tb, ka, groups, stream);
  if (ka.res != nullptr) return launch_gemm_mode<BN, STAGES, A_MN, B_MN, EPI_RES>(ta, tb, ka, groups, stream);
  return launch_gemm_mode<BN, STAGES, A_MN, B_MN, EPI_PLAIN>(ta, tb, ka, groups, stream);
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

#ifdef CPC_GEMM_TIMING
extern "C" int cpc_gemm_set_timing(void* buf) {
  long long* p = static_cast<long long*>(buf);
  return cudaMemcpyToSymbol(g_gemm_timing, &p, sizeof(p)) == cudaSuccess ? CPC_OK : CPC_ERR_LAUNCH;
}
#endif

extern "C" int cpc_gemm_bf16(const CpcGemmDesc* d, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (d == nullptr || d->M <= 0 || d->N <= 0 || d->K <= 0 || d->groups <= 0) return CPC_ERR_ARG;
  if (d->N % 8 != 0) return CPC_ERR_ARG;
  if (d->out_f32 == nullptr && d->out_bf16 == nullptr && d->norm_gamma == nullptr) return CPC_ERR_ARG;
  const int k_per_slab = d->k_per_slab > 0 ? d->k_per_slab : d->K;
  if (d->K % k_per_slab != 0) return CPC_ERR_ARG;
  const int slabs = d->K / k_per_slab;
  int k_splits = d->k_splits > 0 ? d->k_splits : 1;
  if (k_splits > 1 && (d->out_f32 == nullptr || d->out_bf16 != nullptr || d->bias || d->res || d->relu ||
                       d->gate || d->drop_p > 0.f))
    return CPC_ERR_ARG;

  {  // one epilogue feature that needs extra global loads per launch (see gemm_epilogue.cuh)
    const bool drop = d->rng != nullptr && d->drop_p > 0.f;
    const int extra = (d->accumulate ? 1 : 0) + (drop ? 1 : 0) + (d->gate ? 1 : 0) + (d->res ? 1 : 0);
    if (extra > 1) return CPC_ERR_UNSUPPORTED;
    if (d->accumulate && (d->out_bf16 != nullptr || d->relu)) return CPC_ERR_UNSUPPORTED;
  }
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
  ka.a_smul = d->a_slab_mul == 0 ? 1 : (d->a_slab_mul < 0 ? 0 : d->a_slab_mul);
  ka.b_smul = d->b_slab_mul == 0 ? 1 : (d->b_slab_mul < 0 ? 0 : d->b_slab_mul);
  ka.alpha = d->alpha;
  ka.out_f32 = static_cast<float*>(d->out_f32); ka.ldd_f32 = d->ldd_f32; ka.gs_f32 = d->gs_f32;
  ka.out_bf16 = static_cast<__nv_bfloat16*>(d->out_bf16); ka.ldd_bf16 = d->ldd_bf16; ka.gs_bf16 = d->gs_bf16;
  ka.bias = static_cast<const float*>(d->bias); ka.gs_bias = d->gs_bias;
  ka.res = static_cast<const __nv_bfloat16*>(d->res); ka.ld_res = d->ld_res; ka.gs_res = d->gs_res;
  ka.gate = static_cast<const __nv_bfloat16*>(d->gate); ka.ld_gate = d->ld_gate; ka.gs_gate = d->gs_gate;
  ka.gate_scale = d->gate_scale;
  ka.rng = static_cast<const unsigned long long*>(d->rng); ka.rng_site = d->rng_site; ka.drop_p = d->drop_p;
  ka.relu = d->relu; ka.accumulate = d->accumulate;
  ka.norm_gamma = static_cast<const float*>(d->norm_gamma); ka.norm_beta = static_cast<const float*>(d->norm_beta);
  ka.gs_norm_param = d->gs_norm_param;
  ka.norm_bf16 = static_cast<__nv_bfloat16*>(d->norm_bf16); ka.ld_norm_bf16 = d->ld_norm_bf16; ka.gs_norm_bf16 = d->gs_norm_bf16;
  ka.norm_f32 = static_cast<float*>(d->norm_f32); ka.ld_norm_f32 = d->ld_norm_f32; ka.gs_norm_f32 = d->gs_norm_f32;
  ka.norm_mean = static_cast<float*>(d->norm_mean); ka.norm_rstd = static_cast<float*>(d->norm_rstd);
  ka.gs_norm_stat = d->gs_norm_stat;
  ka.norm_eps = d->norm_eps; ka.norm_unbiased = d->norm_unbiased; ka.norm_halo = d->norm_halo;
  static const bool vec_ok = [] {  // CPC_EPI_VEC16=0: 4-byte bf16x2 stores (A/B switch of the quad-transposed epilogue)
    const char* e = getenv("CPC_EPI_VEC16");
    return e == nullptr || e[0] != '0';
  }();
  ka.vec16 = vec_ok && ka.out_bf16 != nullptr && (reinterpret_cast<uintptr_t>(ka.out_bf16) & 15) == 0 && ka.ldd_bf16 % 8 == 0 &&
             ka.gs_bf16 % 8 == 0;
  ka.vec16n = vec_ok && ka.norm_bf16 != nullptr && (reinterpret_cast<uintptr_t>(ka.norm_bf16) & 15) == 0 &&
              ka.ld_norm_bf16 % 8 == 0 && ka.gs_norm_bf16 % 8 == 0;
  const bool fused_norm = d->norm_gamma != nullptr;
  if (fused_norm) {
    if ((d->N != 256 && d->N != 512) || d->a_major != 0 || d->b_major != 0 || k_splits > 1 || d->accumulate || d->gate != nullptr ||
        (d->rng != nullptr && d->drop_p > 0.f) || d->norm_beta == nullptr || d->out_f32 != nullptr ||
        (d->norm_bf16 == nullptr && d->norm_f32 == nullptr) || d->norm_halo < 0 || d->norm_halo > 4)
      return CPC_ERR_UNSUPPORTED;
  }

  // tile-N selection: minimise (waves over the SMs) x (cost of one 128 x bn tile), in SM cycles.  Per 64-wide
  // k-block the tensor pipe needs max(2*bn, shared-memory operand reads at 128 B/cycle) cycles, and TMA has to
  // deliver 16 KB + bn*128 B at the SM's share of the ~6300 B/cycle the L2 sustains chip-wide (<= ~128 B/cycle
  // per SM): narrow tiles re-read A more often, wide tiles leave SMs idle on small problems.  Draining the
  // accumulator costs ~8*bn cycles (TMEM read 64 B/cycle) and overlaps the next tile's main loop.
  int bn = 64;
  {
    const int sms = sm_count();
    const int nkb = (ka.total_kblocks + k_splits - 1) / k_splits;
    double best = 1e30;
    for (int cand = 64; cand <= 256; cand *= 2) {
      if (cand > 64 && cand / 2 >= d->N) break;  // a wider tile would only add padding
      const long long tiles = (long long)((d->M + BM - 1) / BM) * ((d->N + cand - 1) / cand) * d->groups * k_splits;
      const long long waves = (tiles + sms - 1) / sms;
      const double active = tiles < sms ? (double)tiles : (double)sms;
      double rate = 6300.0 / active;
      if (rate > 128.0) rate = 128.0;
      const double load = (16384.0 + 128.0 * cand) / rate;
      double mma = 2.0 * cand;
      const double smem_rd = (4096.0 + 32.0 * cand) * 4.0 / 128.0;
      if (smem_rd > mma) mma = smem_rd;
      const double kb = load > mma ? load : mma;
      const double cost = waves * nkb * kb + 8.0 * cand + 3000.0;
      if (cost < best) { best = cost; bn = cand; }
    }
  }
  if (d->force_bn == 64 || d->force_bn == 128 || d->force_bn == 256) bn = d->force_bn;
  ka.cl_n = 0;
  if (fused_norm) {
    // one BN = 256 accumulator tile = one whole row -- unless that leaves most SMs idle (conv3 / conv4 of the encoder:
    // 16 / 8 row tiles per half batch) or the row is 512 wide: then the row is split over a 4-CTA cluster (EPI_NORM_CL).
    // The split multiplies the main-loop work by four (every chunk streams the whole A tile, and the loop is bound by
    // TMA latency / pipeline depth, not by bytes), so it only pays while 4 x row tiles x 2 concurrent half batches still
    // fit the 148 SMs in one wave: measured, conv1 / conv2 (64 / 32 row tiles) got slower in the step (1.24 vs 1.22 ms).
    static const int cl_below = [] {  // CPC_NORM_CLUSTER_BELOW=<row tiles>: 0 disables the cluster form for 256-wide rows
      const char* e = getenv("CPC_NORM_CLUSTER_BELOW");
      return e != nullptr ? atoi(e) : 17;
    }();
    const long long row_tiles = (long long)((d->M + BM - 1) / BM) * d->groups;
    if (d->N == 512) { bn = 128; ka.cl_n = 4; }
    else if (d->force_bn != 256 && row_tiles < cl_below) { bn = 64; ka.cl_n = 4; }  // force_bn = 256: one tile per row
    else bn = 256;
  }

  CUtensorMap ta, tb;
  int rc = encode_operand_map(&ta, d->a, d->a_dim0, d->a_dim1, d->a_dim2, d->a_stride1, d->a_stride2, d->a_major, BM);
  if (rc != CPC_OK) return rc;
  rc = encode_operand_map(&tb, d->b, d->b_dim0, d->b_dim1, d->b_dim2, d->b_stride1, d->b_stride2, d->b_major, bn);
  if (rc != CPC_OK) return rc;

  switch (bn) {
    case 64: return dispatch_major<64, 8>(ta, tb, ka, d->groups, d->a_major, d->b_major, stream);  // 8 x 24 KB in flight
    case 128: return dispatch_major<128, 6>(ta, tb, ka, d->groups, d->a_major, d->b_major, stream);
    default: return dispatch_major<256, 4>(ta, tb, ka, d->groups, d->a_major, d->b_major, stream);
  }
}
