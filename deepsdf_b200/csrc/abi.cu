// Small C-ABI entry points: version, error string, device check, dropout-mask export.
#include "common.cuh"
#include "host_util.h"
#include "../../include/dsdf.h"

namespace dsdf {

__global__ void dropout_mask_kernel(uint64_t seed, uint32_t layer, uint64_t row0, int64_t n_rows, int width, uint32_t thr,
                                    uint8_t* out) {
  const int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= n_rows * width) return;
  const int64_t r = idx / width;
  const int c = int(idx % width);
  out[idx] = dropout_keep(seed, layer, row0 + uint64_t(r), uint32_t(c), thr) ? 1 : 0;
}

}  // namespace dsdf

using namespace dsdf;

extern "C" int dsdf_abi_version(void) { return DSDF_ABI_VERSION; }
extern "C" const char* dsdf_last_error(void) { return g_last_error; }
extern "C" int dsdf_check_device(void) { return require_sm100(); }
extern "C" int64_t dsdf_launch_count(void) { return int64_t(g_kernel_launches); }
extern "C" float dsdf_dropout_effective_p(float p) { return float(dropout_threshold(p)) / 256.0f; }
extern "C" float dsdf_dropout_scale(float p) { return dropout_scale(p); }

extern "C" int dsdf_dropout_mask(uint64_t seed, uint32_t layer, uint64_t row0, int64_t n_rows, int32_t width, float p,
                                 uint8_t* mask_out, void* stream) {
  if (int e = require_sm100()) return e;
  const int64_t total = n_rows * width;
  if (total <= 0) return 0;
  DSDF_LAUNCHED;
  dropout_mask_kernel<<<unsigned((total + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      seed, layer, row0, n_rows, width, dropout_threshold(p), mask_out);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}
