// Multi-tensor Adam for the optimizer step of cpc/train.py:112-113,472-479 (torch.optim.Adam(g_params, lr,
// betas, eps), no weight decay, no amsgrad): ONE launch updates every parameter of the model and the criterion
// (~110 tensors, 18.4 M fp32 at the default config) instead of torch's per-tensor / foreach kernel lists.
// HBM-bound: per parameter element it reads p, g, m, v and writes p, m, v (28 bytes), optionally also the bf16
// shadow of p that the next step's GEMMs consume (+2 bytes, replaces that step's cast kernel).
//
// CUDA-graph safe: the step counter and the learning rate live in device memory (`state`), a one-thread
// prologue kernel advances the counter and derives the two bias corrections in double precision exactly like
// the Python scalar arithmetic of torch/optim/adam.py (`1 - beta ** step`), so a captured step replays correctly.
#include "cpc_internal.h"
#include "ptx.cuh"

namespace cpc {

constexpr int ADAM_THREADS = 256;
constexpr int ADAM_CHUNK = ADAM_THREADS * 16;  // elements per CTA

struct AdamSpan {  // mirrors CpcAdamSpan
  float* p;
  const float* g;
  float* m;
  float* v;
  __nv_bfloat16* p16;  // optional bf16 shadow of p
  long long n;
  long long first_block;  // prefix sum of ceil(n / ADAM_CHUNK) over the spans before this one
};

// state: [0] step (as double), [1] lr, [2] beta1, [3] beta2, [4] eps, [5] step_size = lr / (1 - beta1^step),
// [6] sqrt(1 - beta2^step)
__global__ void adam_prologue_kernel(double* state) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  const double step = state[0] + 1.0;
  state[0] = step;
  const double bc1 = 1.0 - pow(state[2], step);
  const double bc2 = 1.0 - pow(state[3], step);
  state[5] = state[1] / bc1;
  state[6] = sqrt(bc2);
}

// b1 here is (float)(1 - beta1) and omb2 (float)(1 - beta2): the scalars torch hands to its kernels
__device__ __forceinline__ void adam_elem(float& p, float g, float& m, float& v, float omb1, float b2, float omb2,
                                          float eps, float step_size, float bc2_sqrt) {
  m = m + (g - m) * omb1;                           // exp_avg.lerp_(grad, 1 - beta1)
  v = v * b2 + omb2 * g * g;                        // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
  const float denom = sqrtf(v) / bc2_sqrt + eps;
  p = p - step_size * (m / denom);                  // param.addcdiv_(exp_avg, denom, value=-step_size)
}

__global__ void __launch_bounds__(ADAM_THREADS) adam_kernel(const AdamSpan* __restrict__ spans, int n_spans,
                                                            const double* __restrict__ state, long long total_blocks) {
  pdl_sync();  // PDL: nothing below touches global memory before the preceding grid has completed (cpc_internal.h)
  const float omb1 = (float)(1.0 - state[2]), b2 = (float)state[3], omb2 = (float)(1.0 - state[3]), eps = (float)state[4];
  const float step_size = (float)state[5], bc2_sqrt = (float)state[6];
  // grid == total_blocks: one chunk per CTA; a smaller (throttled) grid walks the chunks persistently
  for (long long blk = blockIdx.x; blk < total_blocks; blk += gridDim.x) {
  // span of this chunk: binary search over the block prefix
  int lo = 0, hi = n_spans - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (spans[mid].first_block <= blk) lo = mid;
    else hi = mid - 1;
  }
  const AdamSpan s = spans[lo];
  const long long base = (blk - s.first_block) * ADAM_CHUNK;
  const bool vec = ((reinterpret_cast<uintptr_t>(s.p) | reinterpret_cast<uintptr_t>(s.g) | reinterpret_cast<uintptr_t>(s.m) |
                     reinterpret_cast<uintptr_t>(s.v)) & 15) == 0 &&
                   (s.p16 == nullptr || (reinterpret_cast<uintptr_t>(s.p16) & 7) == 0);
  if (vec) {
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const long long i = base + ((long long)it * ADAM_THREADS + threadIdx.x) * 4;
      if (i + 3 < s.n) {
        float4 p = *reinterpret_cast<const float4*>(s.p + i);
        const float4 g = *reinterpret_cast<const float4*>(s.g + i);
        float4 m = *reinterpret_cast<const float4*>(s.m + i);
        float4 v = *reinterpret_cast<const float4*>(s.v + i);
        adam_elem(p.x, g.x, m.x, v.x, omb1, b2, omb2, eps, step_size, bc2_sqrt);
        adam_elem(p.y, g.y, m.y, v.y, omb1, b2, omb2, eps, step_size, bc2_sqrt);
        adam_elem(p.z, g.z, m.z, v.z, omb1, b2, omb2, eps, step_size, bc2_sqrt);
        adam_elem(p.w, g.w, m.w, v.w, omb1, b2, omb2, eps, step_size, bc2_sqrt);
        *reinterpret_cast<float4*>(s.p + i) = p;
        *reinterpret_cast<float4*>(s.m + i) = m;
        *reinterpret_cast<float4*>(s.v + i) = v;
        if (s.p16 != nullptr) {
          uint2 o;
          o.x = pack_bf16x2(p.x, p.y);
          o.y = pack_bf16x2(p.z, p.w);
          *reinterpret_cast<uint2*>(s.p16 + i) = o;
        }
      } else {
        for (long long j = i; j < s.n && j < i + 4; ++j) {
          float p = s.p[j], m = s.m[j], v = s.v[j];
          adam_elem(p, s.g[j], m, v, omb1, b2, omb2, eps, step_size, bc2_sqrt);
          s.p[j] = p; s.m[j] = m; s.v[j] = v;
          if (s.p16 != nullptr) s.p16[j] = __float2bfloat16(p);
        }
      }
    }
  } else {
    for (int it = 0; it < 16; ++it) {
      const long long j = base + (long long)it * ADAM_THREADS + threadIdx.x;
      if (j < s.n) {
        float p = s.p[j], m = s.m[j], v = s.v[j];
        adam_elem(p, s.g[j], m, v, omb1, b2, omb2, eps, step_size, bc2_sqrt);
        s.p[j] = p; s.m[j] = m; s.v[j] = v;
        if (s.p16 != nullptr) s.p16[j] = __float2bfloat16(p);
      }
    }
  }
  }
}

}  // namespace cpc

using namespace cpc;

extern "C" int cpc_adam_step(const CpcAdamSpan* spans_dev, int n_spans, int64_t total_blocks, double* state_dev,
                             int advance, int max_ctas, void* stream_) {
  static_assert(sizeof(CpcAdamSpan) == sizeof(AdamSpan), "CpcAdamSpan layout");
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (spans_dev == nullptr || state_dev == nullptr || n_spans <= 0 || total_blocks <= 0 || total_blocks > 0x7fffffffLL)
    return CPC_ERR_ARG;
  if (advance) launch_k(adam_prologue_kernel, dim3(1), dim3(1), 0, stream, state_dev);
  const long long grid = max_ctas > 0 && max_ctas < total_blocks ? max_ctas : total_blocks;
  launch_k(adam_kernel, dim3((unsigned)grid), dim3(ADAM_THREADS), 0, stream, reinterpret_cast<const AdamSpan*>(spans_dev), n_spans,
           static_cast<const double*>(state_dev), (long long)total_blocks);
  return launch_status();
}

extern "C" int cpc_adam_chunk(void) { return ADAM_CHUNK; }
