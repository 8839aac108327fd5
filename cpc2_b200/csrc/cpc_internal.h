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

}  // namespace cpc
