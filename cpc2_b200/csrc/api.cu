// Library-level entry points: status strings, version, driver entry point cache.
#include <stdlib.h>

#include <mutex>

#include "cpc_internal.h"

namespace cpc {

PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

static int g_pdl = -1;  // -1: not decided yet (environment CPC_PDL; default off, see cpc_internal.h)

bool pdl_enabled() {
  if (g_pdl < 0) {
    const char* e = getenv("CPC_PDL");
    g_pdl = (e != nullptr && e[0] == '1') ? 1 : 0;
  }
  return g_pdl != 0;
}

}  // namespace cpc

extern "C" int cpc_set_pdl(int on) {
  const int prev = cpc::pdl_enabled() ? 1 : 0;
  cpc::g_pdl = on ? 1 : 0;
  return prev;
}

extern "C" const char* cpc_strerror(int s) {
  switch (s) {
    case CPC_OK: return "ok";
    case CPC_ERR_ARG: return "invalid argument";
    case CPC_ERR_ALIGN: return "pointer or stride not 16-byte aligned";
    case CPC_ERR_DRIVER: return "cuTensorMapEncodeTiled not available from the driver";
    case CPC_ERR_TENSORMAP: return "tensor map encoding failed";
    case CPC_ERR_LAUNCH: return "kernel launch failed";
    case CPC_ERR_UNSUPPORTED: return "unsupported configuration";
    case CPC_ERR_TOO_LARGE: return "problem exceeds the built-in scratch of the no-allocation entry point";
    default: return "unknown status";
  }
}

extern "C" int cpc_version(void) { return 100; }

extern "C" int cpc_device_arch(void) {
  int dev = 0, maj = 0, min = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return CPC_ERR_LAUNCH;
  if (cudaDeviceGetAttribute(&maj, cudaDevAttrComputeCapabilityMajor, dev) != cudaSuccess) return CPC_ERR_LAUNCH;
  if (cudaDeviceGetAttribute(&min, cudaDevAttrComputeCapabilityMinor, dev) != cudaSuccess) return CPC_ERR_LAUNCH;
  return maj * 10 + min;
}
