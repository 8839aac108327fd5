// Internal helpers shared by the .cu files of libcpc_b200.so.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/cpc_b200.h"

namespace cpc {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
// cuTensorMapEncodeTiled resolved through the runtime (no link-time dependency on libcuda)
PFN_encodeTiled get_encode_tiled();

inline int launch_status() { return cudaGetLastError() == cudaSuccess ? CPC_OK : CPC_ERR_LAUNCH; }

// ---- programmatic dependent launch (PDL) ------------------------------------------------------------------------
// The CPC step is a chain of ~75 dependent launches of 5-60 us; between two of them the GPU otherwise drains the
// first grid, flushes, and only then starts scheduling the second (2-4 us of launch latency + the second kernel's
// prologue: barrier init, TMEM allocation, tensor-map fetch).  Every kernel of this library is therefore launched
// with cudaLaunchAttributeProgrammaticStreamSerialization and begins with
//     <prologue that touches no global memory>;  pdl_wait();  pdl_trigger();
// pdl_wait() (griddepcontrol.wait) blocks until the preceding grid of the stream has COMPLETED and its writes are
// visible -- so no kernel reads or writes global memory before its predecessor is done (RAW, WAR and WAW hazards are
// all covered); pdl_trigger() (griddepcontrol.launch_dependents) then lets the NEXT launch of the stream bring its CTAs
// up next to this kernel's.  Triggering only after the own wait bounds the chain to one early-launched kernel at a
// time (its idle CTAs do not crowd out side-stream work).  Both instructions are no-ops in a grid launched without the
// attribute, and the attribute degrades to a normal dependency after a memset / copy / foreign kernel; inside a
// stream capture the edges become programmatic graph edges.
// MEASURED (profiles/r02_bench_pdl.txt): inside the step's CUDA graph consecutive kernel nodes already start back to
// back (gaps of ~0.1 us in the replay timeline), so PDL has only the prologues left to hide and the early CTAs it
// parks on the SMs take room from the side-stream kernels: 1.165 ms/step with PDL against 1.153 ms without.  The
// attribute is therefore OFF by default; cpc_set_pdl(1) / CPC_PDL=1 turns it on (every GPU test passes either way).
bool pdl_enabled();

#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() { pdl_wait(); pdl_trigger(); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                            Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

}  // namespace cpc
