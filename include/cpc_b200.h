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
  CPC_ERR_UNSUPPORTED = -6,
  CPC_ERR_TOO_LARGE = -7  /* problem exceeds the built-in scratch of a no-allocation entry point */
} CpcStatus;

const char* cpc_strerror(int status);
int cpc_version(void);
/* compute capability major*10+minor of the current device, or a negative status */
int cpc_device_arch(void);
/* Programmatic dependent launch of every kernel of the library (default off -- measured neutral inside the step's
 * CUDA graph; environment CPC_PDL=1 turns it on).
 * Returns the previous setting.  No reference counterpart: launch plumbing of the sm_100a path. */
int cpc_set_pdl(int on);

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
 *   K-major : (kk,      m0 + row_off, g*gmul + slab*slab_mul)
 *   MN-major: (m0,      kk + row_off, g*gmul + slab*slab_mul)        (slab_mul: see the end of the struct; default 1)
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
  /* Fused row normalisation (norm_gamma != NULL): ChannelNorm (cpc/model.py:52-60, unbiased variance) / nn.LayerNorm
   * (cpc/transformers.py:124,130) of every output row, applied to x = alpha*acc + bias (+ res) inside the epilogue --
   * a 256-wide row is exactly one 128 x 256 accumulator tile, so its statistics never leave the SM.  Requires
   * N == 256, K-major operands, no split-K / accumulate / gate / dropout.  out_bf16 (optional) receives the
   * PRE-norm x (what the backward needs); the normalised rows y = (x - mean) * rstd * gamma + beta (ReLU if relu) go
   * to norm_bf16 / norm_f32; mean / rstd fp32 [groups][M] (gs_stat apart).  norm_halo > 0: the norm_bf16 buffer of
   * every group carries that many extra rows in front of row 0 and behind row M-1 (the zero padding of the next
   * strided conv); the kernel clears them. */
  const void* norm_gamma; const void* norm_beta; int64_t gs_norm_param; /* fp32 [groups][N] */
  void* norm_bf16; int64_t ld_norm_bf16, gs_norm_bf16;
  void* norm_f32;  int64_t ld_norm_f32,  gs_norm_f32;
  void* norm_mean; void* norm_rstd; int64_t gs_norm_stat;
  float norm_eps; int norm_unbiased; int norm_halo;
  /* Outer-coordinate step per slab: coordinate = g*gmul + slab*slab_mul.  0 selects the default 1 (consecutive slabs),
   * -1 a step of zero (the operand is read again for every slab: a reduction over [hi | lo] parts of A against ONE B),
   * any other value is the step itself (slabs that are `groups` planes apart: A laid out [part][group][m][k]). */
  int a_slab_mul, b_slab_mul;
} CpcGemmDesc;

int cpc_gemm_bf16(const CpcGemmDesc* desc, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Row normalisation: ChannelNorm(+ReLU) of the encoder (cpc/model.py:52-60,103-107: mean / UNBIASED
 * variance over channels, eps inside the sqrt, affine [1,C,1]) and nn.LayerNorm of the transformer
 * blocks (cpc/transformers.py:124,130).  Rows are [C] bf16 vectors; a CpcRowMap places logical row r
 * at buffer row (r / seq_len) * seq_stride + row_off + r % seq_len (seq_len == 0: identity), which is
 * how the zero-halo channels-last conv buffers are addressed; the halo rows of an OUTPUT map (rows [0, row_off)
 * and [row_off + seq_len, seq_stride) of every sequence) are cleared by the kernel.  C must be 32 * {1,2,4,8,16}.
 */
typedef struct { int64_t seq_len, seq_stride, row_off; } CpcRowMap;

typedef struct {
  int64_t rows; int C;
  int64_t rows_per_group;            /* gamma/beta are [groups][C]; 0 => one group */
  const void* in; CpcRowMap in_map;  /* bf16 */
  const void* gamma; const void* beta; /* fp32, may be null */
  void* out_bf16; void* out_f32; CpcRowMap out_map; /* either may be null */
  void* mean; void* rstd;            /* fp32 [rows], saved for backward (may be null) */
  float eps; int unbiased; int relu;
} CpcRowNormFwd;

typedef struct {
  int64_t rows; int C;
  int64_t rows_per_group;
  const void* dout; int dout_f32; CpcRowMap dout_map; /* gradient w.r.t. the output, bf16 or fp32 */
  const void* dout2; CpcRowMap dout2_map;             /* optional second fp32 gradient, added */
  const void* in; CpcRowMap in_map;                   /* forward input (bf16) */
  const void* mean; const void* rstd; const void* gamma; const void* beta;
  void* dx_bf16; void* dx_f32; CpcRowMap dx_map; int dx_accumulate;
  void* dgamma; void* dbeta;  /* fp32 [groups][C], atomically accumulated (pre-zeroed by caller) */
  void* dxsum;                /* fp32 [groups][C] column sums of dx = bias gradient of the producer */
  int unbiased; int relu;
} CpcRowNormBwd;

int cpc_rownorm_fwd(const CpcRowNormFwd* desc, void* stream);
int cpc_rownorm_bwd(const CpcRowNormBwd* desc, void* stream);

/* ------------------------------------------------------------------------------------------------
 * conv0 (1 -> C channels, k 10, stride 5, pad 3) + ChannelNorm + ReLU fused (cpc/model.py:85,103).
 * x fp32 [N, W]; out bf16 [N, L0 + 2*halo, C] channels-last with `halo` zero rows either side
 * (L0 = (W - 4) / 5 + 1); mean/rstd fp32 [N*L0].  Backward recomputes the conv from x and accumulates
 * dw [C,10], db, dgamma, dbeta (fp32, pre-zeroed) from dout (same layout as out).
 */
typedef struct {
  int N, W, C, halo; float eps;
  const void* x; const void* w; const void* b; const void* gamma; const void* beta;
  void* out; void* mean; void* rstd;
  const void* dout; void* dw; void* db; void* dgamma; void* dbeta;
  int64_t x_stride; /* elements between the rows of x; 0 => W (lets the past / future windows of a [B,2,1,W] batch,
                       cpc/train.py:95-96, be read in place) */
} CpcConv0;

int cpc_conv0_cn_relu_fwd(const CpcConv0* desc, void* stream);
int cpc_conv0_cn_relu_bwd(const CpcConv0* desc, void* stream);

/* ------------------------------------------------------------------------------------------------
 * The encoder's other normalisation modes (cpc/model.py:71-82: normMode "batchNorm" -> nn.BatchNorm1d,
 * "instanceNorm" -> nn.InstanceNorm1d(affine=True), "ID" -> IDModule), each followed by ReLU (cpc/model.py:102-107).
 * `in` is the convolution output, bf16 [rows, C] dense; statistics run down the columns over groups of
 * `rows_per_group` consecutive rows (batchNorm: one group of N*L rows; instanceNorm: N groups of L rows), biased
 * variance, fp32.  C in {32, 64, 128, 256, 512}.
 *   mode CPC_COLNORM_ID    : out = relu(in)
 *   mode CPC_COLNORM_STATS : mean / rstd [groups, C] are COMPUTED (and saved for the backward); running_mean / running_var
 *                            [C] (may be null; one group only) are updated as nn.BatchNorm1d does in training mode
 *                            (momentum, unbiased variance)
 *   mode CPC_COLNORM_GIVEN : mean / rstd [groups, C] are READ (nn.BatchNorm1d in eval mode: running_mean and
 *                            1 / sqrt(running_var + eps)); the backward treats them as constants
 * `partial` is caller-provided scratch of 2 * groups * ceil(rows_per_group / CPC_COLNORM_CHUNK_ROWS) * C floats
 * (chunk partials, combined in a fixed order: results do not depend on the launch shape).
 * Forward: out_bf16 through out_map (halo rows cleared, see CpcRowMap) and / or out_f32 dense [rows, C].
 * Backward: dout (bf16 or fp32, through dout_map) -> dx_bf16 through dx_map (halo rows cleared) and / or dx_f32 dense;
 * dgamma / dbeta [C] and dxsum [C] (column sums of dx = bias gradient of the convolution in front; exactly zero under
 * computed statistics) are ADDED atomically; `sums` is scratch of 2 * groups * C floats.
 */
#define CPC_COLNORM_CHUNK_ROWS 256
enum { CPC_COLNORM_ID = 0, CPC_COLNORM_STATS = 1, CPC_COLNORM_GIVEN = 2 };

typedef struct {
  int64_t rows; int C;
  int64_t rows_per_group;              /* 0 => one group */
  int mode; int relu;
  const void* in;                      /* bf16 [rows, C] */
  const void* gamma; const void* beta; /* fp32 [C], both or neither */
  void* mean; void* rstd;              /* fp32 [groups, C] */
  void* running_mean; void* running_var; float momentum; float eps;
  void* partial;
  void* out_bf16; CpcRowMap out_map; void* out_f32;
  const void* dout; int dout_f32; CpcRowMap dout_map;
  void* sums;
  void* dx_bf16; CpcRowMap dx_map; void* dx_f32;
  void* dgamma; void* dbeta; void* dxsum;
} CpcColNorm;

int cpc_colnorm_fwd(const CpcColNorm* desc, void* stream);
int cpc_colnorm_bwd(const CpcColNorm* desc, void* stream);

/* conv0 (cpc/model.py:85) WITHOUT a fused norm, for the modes above: x fp32 [N, W] (x_stride as in CpcConv0) ->
 * out bf16 [N * L0, C] dense = conv + bias (gamma, beta, mean, rstd, halo unused).  Backward: dout = fp32 [N * L0, C]
 * dense gradient of that output -> dw [C, 10] ADDED atomically (the bias gradient is cpc_colnorm_bwd's dxsum). */
int cpc_conv0_plain_fwd(const CpcConv0* desc, void* stream);
int cpc_conv0_plain_bwd(const CpcConv0* desc, void* stream);

/* conv weight fp32 [Co,Ci,k] (k == 2*s) -> bf16 GEMM operands: wf [Co, k*Ci] (forward / wgrad) and
 * wd [s*Ci, 2*Co] (dgrad; may be null).  cpc_conv_wgrad_unpack is the inverse for the fp32 gradient and ADDS
 * (red.global) into dw, so the two half-batches of a step can reduce into one zero-initialised gradient. */
int cpc_conv_weight_pack(const void* w, void* wf, void* wd, int Co, int Ci, int k, int s, void* stream);
int cpc_conv_wgrad_unpack(const void* dwf, void* dw, int Co, int Ci, int k, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Gradient all-reduce of the data-parallel step (DistributedDataParallel, cpc/train.py:523-527;
 * cpc/distributed_training/distributed_mode.py:139-142) as ONE kernel over NVSwitch multicast memory, in place:
 * every rank calls it with the multicast address of the same n-float span of a symmetric buffer (n % 4 == 0,
 * 16-byte aligned) and a device array of `world` pointers to the ranks' signal pads (each at least
 * cpc_allreduce_pad_bytes(world) bytes, zero-initialised once).  Afterwards every rank's copy holds
 * scale * sum over ranks.  The caller maps the buffers (CUDA multicast objects / torch symmetric memory); the
 * kernel only needs the addresses.  max_blocks <= 0: default CTA count. */
int cpc_allreduce_mc(void* mc_ptr, int64_t n, void* const* signal_pads_dev, int rank, int world, float scale,
                     int max_blocks, void* stream);
int cpc_allreduce_pad_bytes(int world);

/* ------------------------------------------------------------------------------------------------
 * Streaming helpers. */
int cpc_cast_f32_to_bf16(const void* in, void* out, int64_t n, void* stream);
int cpc_cast_bf16_to_f32(const void* in, void* out, int64_t n, void* stream);
/* nn.Dropout on the predictions (cpc/criterion/criterion.py:113,168-169, `--dropout`), in place on a contiguous
 * bf16 / fp32 tensor: x = keep ? x / (1 - p) : 0 with a counter-hash mask of (rng = device uint64[2] {seed, step},
 * site, element index); calling it again with the same (rng, site) on the gradient is the backward. */
int cpc_dropout(void* x, int is_bf16, int64_t n, const void* rng, uint32_t site, float p, void* stream);
/* out[g][c] += sum_r in[g][r][c] (fp32 out, pre-initialised): bias gradients */
int cpc_colsum(const void* in, int in_is_bf16, void* out, int64_t rows, int C, int64_t ld, int64_t gs_in,
               int groups, void* stream);
/* out[i] (+)= sum_g in[g][i] */
int cpc_groupsum(const void* in, int in_is_bf16, void* out_f32, void* out_bf16, int64_t n, int groups,
                 int accumulate, void* stream);
/* ------------------------------------------------------------------------------------------------
 * Multi-tensor Adam: the optimizer.step() of cpc/train.py:112 with torch.optim.Adam(g_params, lr, betas, eps)
 * (cpc/train.py:472-479; no weight decay, no amsgrad) as ONE launch over every parameter.
 * spans: DEVICE array; span i covers ceil(n / cpc_adam_chunk()) CTAs starting at first_block (prefix sum,
 * total_blocks = the sum).  state: DEVICE double[8] = {step, lr, beta1, beta2, eps, -, -, -}; with advance != 0
 * the call first advances step by one and derives the bias corrections on the device (CUDA-graph safe) -- an
 * optimizer step issued in several launches (one per gradient bucket, as each is reduced) advances once.
 * p16 (optional): bf16 copy of the updated parameter, written in the same pass.
 * max_ctas > 0: the update runs on at most that many persistent CTAs (a THROTTLED pass: issued next to the latency-bound
 * recurrent backward it would otherwise starve of HBM bandwidth, it takes a fraction of the bandwidth for longer);
 * 0: one CTA per chunk. */
typedef struct {
  void* p; const void* g; void* m; void* v; void* p16;
  int64_t n; int64_t first_block;
} CpcAdamSpan;
int cpc_adam_step(const CpcAdamSpan* spans_dev, int n_spans, int64_t total_blocks, double* state_dev, int advance,
                  int max_ctas, void* stream);
int cpc_adam_chunk(void);

/* extIdx = ((seqIdx + t) mod S) + batchIdx*S, int64, flat (b,n,t) order -- criterion.py:256-266 */
int cpc_neg_ext_index(const void* batch_idx, const void* seq_idx, void* ext, int B, int NEG, int W, int S,
                      void* stream);

/* ------------------------------------------------------------------------------------------------
 * InfoNCE gather + cross entropy + argmax (criterion.py:345-357) on the all-pairs logits
 * logits fp32 [K][B*W][ld] (ld >= B*S), ext int64 [B][NEG][W], weights fp32 [B*W] or null.
 * fwd: lse [K][B*W]; loss_sum[k] += sum_rows w*(lse - l_pos); hits[k] += #rows with argmax == positive.
 * bwd: dlogits bf16 [K][B*W][ld] = gscale[k] * w/(B*W) * (softmax - onehot0), dense, zeros elsewhere;
 *      dlogits_lo (optional) receives the bf16 rounding residual so hi + lo ~ fp32 accuracy.
 */
typedef struct {
  int K, B, W, S, NEG; int64_t ld;
  const void* logits; const void* ext; const void* weights;
  void* lse; void* loss_sum; void* hits;
  const void* gscale; void* dlogits;
  void* dlogits_lo; /* optional bf16 residual (dlogits_f32 - dlogits), same shape; null to skip */
  float out_scale;  /* fwd: loss_sum / hits receive out_scale * (their sums); 0 => 1 (1/(B*W) yields the means directly) */
} CpcNce;

int cpc_nce_fwd(const CpcNce* desc, void* stream);
int cpc_nce_bwd(const CpcNce* desc, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Recurrence of CPCAR's nn.LSTM / nn.GRU / nn.RNN (cpc/model.py:171-179,196), H in {32,64,128,256,512}.
 * cell = 0 LSTM (gate slots i,f,g,o), 1 GRU, 2 tanh RNN.  Every cell uses FOUR gate slots of width H:
 *   LSTM: gx fp32 [N,S,4H] = x.W_ih^T + b_ih + b_hh, w_hh bf16 [4H,H] as torch stores them;
 *   GRU : slots (r, z, n_x, hn): gx = (W_ir x + b_ir + b_hr | W_iz x + b_iz + b_hz | W_in x + b_in | b_hn),
 *         w_hh = (W_hr | W_hz | 0 | W_hn); n = tanh(n_x + r * hn), h = (1-z) n + z h_prev;
 *   RNN : slot 0 only (the other three slots of gx / w_hh are zero).
 * fwd writes out fp32 [N,S,H], h_pad bf16 [N,S+1,H] (row 0 = h0), gates bf16 [N,S,4H] (gate PRE-activations, recomputed by bwd;
 * GRU slot 3 = W_hn h + b_hn), c_all fp32 [N,S,H] (LSTM: cell states; GRU/RNN: h in fp32), h_n / c_n fp32 [N,H]
 * (any may be null).  bwd consumes dout fp32 [N,S,H] and writes dgates bf16 [N,S,4H], the gradient w.r.t. the
 * four slot pre-activations (GRU: d r, d z, d n_x, d hn = d n_x * r); dW / dx follow as GEMMs.
 */
typedef struct {
  int N, S, H, cell;
  const void* gx; const void* w_hh; const void* h0; const void* c0;
  void* h_pad; void* out; void* gates; void* c_all; void* h_n; void* c_n;
  const void* dout; void* dgates;
} CpcLstm;

int cpc_lstm_fwd(const CpcLstm* desc, void* stream);
int cpc_lstm_bwd(const CpcLstm* desc, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Causal relative-position multi-head attention (cpc/transformers.py:52-70, :97-104).
 * qkv bf16 [G][Bt*S][3*d] (q|k|v, head h at columns h*dk), krel fp32 [G][dk][S] or null,
 * out bf16 [G][Bt*S][d], probs bf16 [G][Bt][Hh][S][S] (saved), S <= 256, dk <= 64.
 * bwd: dout bf16 like out -> dqkv bf16 like qkv, dkrel fp32 [G][dk][S] (accumulated), dscore scratch.
 */
typedef struct {
  int G, Bt, Hh, S, dk; float scale;
  const void* qkv; int64_t gs_qkv; const void* krel;
  void* out; int64_t gs_out; void* probs;
  const void* rng; uint32_t rng_site; float drop_p;
  const void* dout; void* dqkv; void* dscore; void* dkrel;
  /* Row statistic of the backward, D_i = sum_j P_ij dP_ij (the softmax Jacobian's common term).  dS = P o (dP - D) must
   * sum to zero along every row to fp32 rounding: K and V of a trained model carry a large component common to all
   * keys and any row-coherent error of D is multiplied by it in dQ.  Two exact ways to get it:
   *   (a) out_f32 + rowsum given (fwd writes both): fwd forms O from the SAME bf16-rounded probabilities it stores,
   *       O = dropout_scale * sum_j keep_ij P^_ij V_j, and rowsum_i = sum_j P^_ij; bwd takes D_i = dO_i . O_i / rowsum_i
   *       -- algebraically sum_j P^_ij dP~_ij / sum_j P^_ij with the very P^ and V the backward multiplies;
   *   (b) either null: bwd recomputes dP once more in a pre-pass and reduces P^ o dP~ itself (~20 % more time).
   * out_f32: [G][Bt*S][d] fp32, group stride gs_out.  rowsum: [G][Bt][Hh][S] fp32. */
  void* out_f32;
  void* rowsum;
} CpcAttn;

int cpc_attn_fwd(const CpcAttn* desc, void* stream);
int cpc_attn_bwd(const CpcAttn* desc, void* stream);

/* cpc_attn_bwd for callers that cannot allocate device memory (code under CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and results as cpc_attn_bwd on its general (CUDA-core, any S <= 256 / dk <= 64) kernel -- the one that
 * needs the `dscore` scratch -- but desc->dscore is ignored: the scratch is a file-scope device array inside the
 * library of CPC_ATTN_NOALLOC_MAX_ELEMS bf16 elements (8 MiB, held by every GPU that loads the library until exit).
 * G*Bt*Hh*S*S above that limit returns CPC_ERR_TOO_LARGE.  The call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine).  Every call shares the one array and it keeps the previous call's contents (the kernel
 * writes every element before reading it, nothing needs clearing): calls that OVERLAP on the GPU -- two streams, two
 * branches of a graph -- race; issue them on one stream or order them with events. */
#define CPC_ATTN_NOALLOC_MAX_ELEMS (4 * 1024 * 1024)
int cpc_attn_bwd_noalloc(const CpcAttn* desc, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CPC_B200_H */
