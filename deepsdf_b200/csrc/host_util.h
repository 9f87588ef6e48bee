// Host-side helpers shared by the C-ABI translation units: thread-local error string, CUDA error
// checks, TMA tensor-map construction through the driver entry point (no link-time libcuda dependency,
// so the library also loads on a GPU-less build box).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

namespace dsdf {

extern thread_local char g_last_error[512];

inline int set_error(const char* fmt, const char* a = "", const char* b = "") {
  snprintf(g_last_error, sizeof(g_last_error), fmt, a, b);
  return 1;
}

#define DSDF_CHECK_CUDA(expr)                                                                   \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      snprintf(::dsdf::g_last_error, sizeof(::dsdf::g_last_error), "%s failed: %s (%s:%d)", #expr, \
               cudaGetErrorString(_e), __FILE__, __LINE__);                                     \
      return 2;                                                                                 \
    }                                                                                           \
  } while (0)

#define DSDF_REQUIRE(cond, msg)                                                                  \
  do {                                                                                          \
    if (!(cond)) {                                                                              \
      snprintf(::dsdf::g_last_error, sizeof(::dsdf::g_last_error), "%s (%s:%d)", msg, __FILE__, __LINE__); \
      return 1;                                                                                 \
    }                                                                                           \
  } while (0)

// Number of kernels this library has launched since it was loaded (host-side counter next to every launch statement of the
// fused / small-network / sampler / consumer paths; the generic fp32 path is not counted).  bench.py reports the
// difference over its timed region as `gpu_launches`.
extern long long g_kernel_launches;
#define DSDF_LAUNCHED (++::dsdf::g_kernel_launches)

// sm_100 only: anything else is an error, there is no fallback path.
int require_sm100();
int sm_count();

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
PFN_encodeTiled get_encode_tiled();

// 2-D bf16 tensor [outer][inner] with row pitch `pitch_bytes`; box {box_inner, box_outer}; SWIZZLE_128B.
int make_tmap_bf16_2d(CUtensorMap* out, const void* gptr, uint64_t inner, uint64_t outer, uint64_t pitch_bytes,
                      uint32_t box_inner, uint32_t box_outer);
// Same tensor, no swizzle: a box {8, rows} lands as `rows` contiguous 16-byte pieces = one k-core column of the
// SWIZZLE_NONE K-major canonical layout (the aux operands of the forward chain).
int make_tmap_bf16_2d_noswz(CUtensorMap* out, const void* gptr, uint64_t inner, uint64_t outer, uint64_t pitch_bytes,
                            uint32_t box_inner, uint32_t box_outer);

}  // namespace dsdf
