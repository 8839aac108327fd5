/* cpc_b200.h -- C ABI of libcpc_b200.so: hand-written sm_100a kernels for the CPC2 training step.
 *
 * The reference (MarvinLvn/CPC2) has no FFI: its hot path is stock PyTorch modules
 * (cpc/model.py, cpc/transformers.py, cpc/criterion/criterion.py) driven by cpc/train.py:72-142.
 * This library is the native layer those modules are re-implemented on; every entry point names the
 * reference expression it replaces.  The Python side (cpc2_b200/*.py) binds it with ctypes and keeps the
 * reference's module interface; INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - all pointers are DEVICE pointers borrowed from the caller (torch tensors); nothing is allocated or
 *     freed here; `stream` is a cudaStream_t passed as void*;
 *   - activations are bf16, CHANNELS-LAST ([batch, time, channel]); statistics, losses and parameter
 *     gradients are fp32; indices are int64 like torch's;
 *   - every function returns 0 (CPC_OK) or a negative CpcStatus; cpc_strerror() names it; launches are
 *     asynchronous on `stream` (CUDA-graph capturable, no host sync, no global state except the cached
 *     driver entry point for tensor-map encoding).
 */
#ifndef CPC_B200_H
#define CPC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  CPC_OK = 0,
  CPC_ERR_ARG = -1,       /* bad size / null pointer / unsupported combination */
  CPC_ERR_ALIGN = -2,     /* pointer or stride not 16-byte aligned */
  CPC_ERR_DRIVER = -3,    /* cuTensorMapEncodeTiled entry point not found */
  CPC_ERR_TENSORMAP = -4, /* cuTensorMapEncodeTiled rejected the operand view */
  CPC_ERR_LAUNCH = -5,    /* kernel launch failed (see cudaGetLastError) */
  CPC_ERR_UNSUPPORTED = -6
} CpcStatus;

const char* cpc_strerror(int status);
int cpc_version(void);
/* compute capability major*10+minor of the current device, or a negative status */
int cpc_device_arch(void);

/* ------------------------------------------------------------------------------------------------
 * Grouped bf16 GEMM on tcgen05 tensor cores:  D[g][m,n] = alpha * sum_k A_g(m,k) * B_g(n,k) (+epilogue)
 * Replaces every dense contraction of the step: nn.Conv1d (cpc/model.py:85-94) as implicit GEMM,
 * nn.Linear (cpc/transformers.py:81-84,110-111,129; criterion.py:146), the LSTM/GRU input projections
 * (cpc/model.py:171-179) and the broadcast-multiply-mean logits (criterion.py:170-171).
 *
 * Each operand is described as a 3-D view (dim0 contiguous, dim1, dim2) in ELEMENTS:
 *   major = 0 (K-major):  dim0 = reduction index k, dim1 = m (or n) row,   dim2 = group / slab
 *   major = 1 (MN-major): dim0 = m (or n),          dim1 = reduction row,  dim2 = group / slab
 * The reduction index is k = slab*k_per_slab + kk; coordinates used for tile (m0|n0, kk, slab, group g):
 *   K-major : (kk,      m0 + row_off, g*gmul + slab)
 *   MN-major: (m0,      kk + row_off, g*gmul + slab)
 * Out-of-range coordinates read as zero (TMA fill), so M, N, k_per_slab need not be tile multiples.
 * Rows may overlap (stride1 < dim0): that is how the im2col matrix of a strided conv is expressed.
 */
typedef struct {
  int M, N, K;     /* per group; K = slabs * k_per_slab */
  int groups;      /* blockIdx.z groups */
  int k_per_slab;  /* 0 => K */
  int k_splits;    /* >1: split-K, fp32 atomic accumulation into out_f32 (must be pre-initialised) */
  int force_bn;    /* 0 = auto, else 64/128/256 */
  /* operand A */
  const void* a; int a_major;
  int64_t a_dim0, a_dim1, a_dim2, a_stride1, a_stride2;
  int a_row_off, a_gmul;
  /* operand B */
  const void* b; int b_major;
  int64_t b_dim0, b_dim1, b_dim2, b_stride1, b_stride2;
  int b_row_off, b_gmul;
  /* epilogue, applied in this order: *alpha, +bias[n], +res, relu, gate, dropout */
  float alpha;
  void* out_f32;  int64_t ldd_f32,  gs_f32;   /* fp32 output (optional), row stride / group stride */
  void* out_bf16; int64_t ldd_bf16, gs_bf16;  /* bf16 output (optional) */
  const void* bias; int64_t gs_bias;          /* fp32 [groups][N] */
  const void* res;  int64_t ld_res, gs_res;   /* bf16 residual added to the result */
  const void* gate; int64_t ld_gate, gs_gate; float gate_scale; /* x = gate>0 ? x*gate_scale : 0 (ReLU/dropout bwd) */
  const void* rng; uint32_t rng_site; float drop_p; /* rng: device uint64[2] {seed, step}; inverted dropout */
  int relu;
  int accumulate;  /* out_f32 += result instead of = */
} CpcGemmDesc;

int cpc_gemm_bf16(const CpcGemmDesc* desc, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CPC_B200_H */
