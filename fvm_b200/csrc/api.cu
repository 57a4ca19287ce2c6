// Status strings and version of the C-ABI (include/fvm_b200.h).
#include <cuda_runtime.h>
#include "../../include/fvm_b200.h"

extern "C" int fvm_version(void) { return 200; }  // 0.2.0

extern "C" const char* fvm_status_string(int status) {
  switch (status) {
    case FVM_OK: return "ok";
    case FVM_ERR_INVALID_ARG: return "invalid argument (NULL pointer, J < 3, P < 1, theta outside (0, 1] or bad enum)";
    case FVM_ERR_UNSUPPORTED: return "size not supported by this entry point";
    case FVM_ERR_SCRATCH_LIMIT: return "P * n_tiles exceeds the library's built-in scratch (fvm_spike_static_max_tiles); "
                                       "use fvm_spike_local with a caller-provided rs array";
    case FVM_ERR_COLLECTIVE: return "the caller's all-gather (fvm_spike_steps) reported an error";
    default: break;
  }
  if (status > 0) return cudaGetErrorString((cudaError_t)status);
  return "unknown fvm_b200 status";
}
