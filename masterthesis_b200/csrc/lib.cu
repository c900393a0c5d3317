// Library-level entry points: version and error strings.
#include <cuda_runtime.h>
#include "../../include/mt_b200.h"

extern "C" int mt_version(void) { return 100; }

extern "C" const char* mt_error_string(int code) {
  if (code == 0) return "ok";
  if (code <= -3000) return "cudaFuncSetAttribute(MaxDynamicSharedMemorySize) failed (code = -(3000+cudaError))";
  if (code <= -2000) return "cuTensorMapEncodeTiled rejected the tensor map (code = -(2000+CUresult))";
  switch (code) {
    case -1000: return "launch grid out of range";
    case -1001: return "cuTensorMapEncodeTiled driver entry point unavailable (no CUDA driver?)";
    case -1002: return "null descriptor";
    case -1003: return "channel counts must be stored as multiples of 8 / unsupported channel count";
    case -1004: return "unsupported stride";
    case -1005: return "bad dimensions";
    case -1006: return "stride-2 phase without taps";
    case -1007: return "reflection pad must be smaller than the spatial size";
    case -1010: return "unsupported (cg, BN) implicit-GEMM tile";
    case -1013: return "stream fork/join of the sub-pixel phases failed";
    case -1014: return "merged-parity / paired-pixel conv modes need even extents: set MtConv2d.single = 1 for odd sizes (and pack the plain operand layouts)";
    case -1015: return "weight gradient: `splits` does not match mt_conv2d_wgrad_splits() / mt_conv_transpose2d_wgrad_splits()";
    case -1016: return "mt_loss_l1_static: n exceeds mt_loss_l1_static_max_n() (the built-in scratch array), or the array's device address is unavailable";
    case -1012: return "`parts` does not match the *_parts() query of this entry point for this shape";
    case -1011: return "implicit-GEMM shared-memory ring is not at the expected offset";
    case -1009: return "row-window mode requested for an ineligible convolution";
    case -1008: return "missing parameter tensor for this norm mode";
    default: break;
  }
  if (code < 0 && code > -1000) return cudaGetErrorString((cudaError_t)(-code));
  return "unknown error";
}
