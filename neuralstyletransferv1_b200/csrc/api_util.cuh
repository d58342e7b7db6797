// Helpers shared by the translation units that export C entry points (nst_api.cu, masks.cu, temporal.cu).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include "../../include/b200nst.h"

namespace nst {
// record the message nst_last_error() returns (thread-local) and hand back `code`
int api_fail(int code, const char* fmt, ...);
// mask_r = GaussianBlur_reflect(label == r), sigma = feather / 3 (nst_api.cu); label: H*W bytes, tmp: R*H*W floats, all [dev]
int feather_labels(const uint8_t* label, int R, int feather, float* masks, float* tmp, int H, int W, cudaStream_t st, const char* who);
}  // namespace nst

#define NST_CUDA_TRY(expr)                                                                                                    \
  do {                                                                                                                        \
    cudaError_t _e = (expr);                                                                                                  \
    if (_e != cudaSuccess) return nst::api_fail(NST_ERR_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
  } while (0)
