// Argument block shared by the attention kernels (attention.cu: general CUDA-core path; attention_tc.cu:
// tensor-core path for S <= 128).
#pragma once
#include <cuda_bf16.h>

namespace cpc {

struct AttnArgs {
  const __nv_bfloat16* qkv;  // [G][T][3d]: q | k | v, head h at columns h*dk
  long long gs_qkv;
  const float* krel;  // [G][dk][S] or null
  __nv_bfloat16* out;  // [G][T][d]
  long long gs_out;
  __nv_bfloat16* probs;  // [G][Bt][Hh][S][S] softmax probabilities before dropout
  int S, dk, d, Bt, Hh;
  float scale;
  const unsigned long long* rng;
  unsigned int site;
  float drop_p;
  // backward
  const __nv_bfloat16* dout;  // [G][T][d]
  __nv_bfloat16* dqkv;        // [G][T][3d]
  __nv_bfloat16* dscore;      // scratch, same shape as probs
  float* dkrel;               // [G][dk][S], atomically accumulated
  float* out_f32;             // optional fp32 copy of out (fwd writes; with rowsum: bwd reads it for D_i = dO_i . O_i / rowsum_i)
  float* rowsum;              // optional [G][Bt][Hh][S]: sum_j of the bf16-rounded probabilities of row i (fwd writes)
};

// tensor-core path (attention_tc.cu); returns CPC_ERR_UNSUPPORTED when the shape is outside its range
int attn_tc_supported(const AttnArgs& a);
int attn_fwd_tc(const AttnArgs& a, int G, cudaStream_t stream);
int attn_bwd_tc(const AttnArgs& a, int G, cudaStream_t stream);

}  // namespace cpc
