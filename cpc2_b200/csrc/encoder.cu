// Encoder-specific kernels (cpc/model.py:63-108).
//   * conv0 + ChannelNorm + ReLU lives in conv0.cu.
//   * packing of conv1..4 weights into the two bf16 GEMM operand layouts (forward / wgrad and dgrad).
#include "cpc_internal.h"
#include "ptx.cuh"

namespace cpc {

// conv weight [Co, Ci, k] fp32 -> Wf bf16 [Co, k*Ci] (index kk*Ci + ci; forward B operand, K-major) and
// Wd bf16 [s*Ci, 2*Co] (row r*Ci + ci, column h*Co + co = W[co, ci, r + (1-h)*s]; dgrad B operand).
__global__ void conv_weight_pack_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wf,
                                        __nv_bfloat16* __restrict__ wd, int Co, int Ci, int k, int s) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  // 32-bit index arithmetic (the launcher checks Co*Ci*k < 2^31): three 64-bit divisions per element made this
  // 2 MB repack as slow as a 30 MB one, and the first conv GEMM of the step waits for it
  const unsigned total = (unsigned)Co * Ci * k, uk = k, uci = Ci;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned q = i / uk;
    const int kk = (int)(i - q * uk);
    const int co = (int)(q / uci);
    const int ci = (int)(q - co * uci);
    const __nv_bfloat16 v = __float2bfloat16(w[i]);
    wf[(long long)co * k * Ci + (long long)kk * Ci + ci] = v;
    if (wd != nullptr) {
      const int h = kk >= s ? 0 : 1;
      const int r = kk >= s ? kk - s : kk;
      wd[((long long)r * Ci + ci) * (2 * Co) + (long long)h * Co + co] = v;
    }
  }
}
// inverse of the forward packing for the weight gradient: dWf fp32 [Co, k*Ci] -> dW fp32 [Co, Ci, k]
__global__ void conv_wgrad_unpack_kernel(const float* __restrict__ dwf, float* __restrict__ dw, int Co, int Ci, int k) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  const unsigned total = (unsigned)Co * Ci * k, uk = k, uci = Ci;
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned q = i / uk;
    const int kk = (int)(i - q * uk);
    const int co = (int)(q / uci);
    const int ci = (int)(q - co * uci);
    red_add_f32(dw + i, dwf[(long long)co * k * Ci + (long long)kk * Ci + ci]);
  }
}

}  // namespace cpc

using namespace cpc;

extern "C" int cpc_conv_weight_pack(const void* w, void* wf, void* wd, int Co, int Ci, int k, int s, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (w == nullptr || wf == nullptr || Co <= 0 || Ci <= 0 || k != 2 * s) return CPC_ERR_ARG;
  const long long total = (long long)Co * Ci * k;
  if (total >= 0x7fffffffLL) return CPC_ERR_UNSUPPORTED;
  const int blocks = (int)((total + 255) / 256 > 1184 ? 1184 : (total + 255) / 256);
  launch_k(conv_weight_pack_kernel, dim3(blocks), dim3(256), 0, stream, static_cast<const float*>(w), static_cast<__nv_bfloat16*>(wf),
                                                      static_cast<__nv_bfloat16*>(wd), Co, Ci, k, s);
  return launch_status();
}

extern "C" int cpc_conv_wgrad_unpack(const void* dwf, void* dw, int Co, int Ci, int k, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (dwf == nullptr || dw == nullptr) return CPC_ERR_ARG;
  const long long total = (long long)Co * Ci * k;
  if (total <= 0 || total >= 0x7fffffffLL) return CPC_ERR_UNSUPPORTED;
  const int blocks = (int)((total + 255) / 256 > 1184 ? 1184 : (total + 255) / 256);
  launch_k(conv_wgrad_unpack_kernel, dim3(blocks), dim3(256), 0, stream, static_cast<const float*>(dwf), static_cast<float*>(dw), Co, Ci, k);
  return launch_status();
}
