// dsdf_clip_coef_static: the gradient-clipping coefficient of dsdf_clip_coef (train_host.cu) for callers that cannot
// provide - or allocate - the 4 KB workspace: code under CUDA-graph capture, tight serving loops, harnesses that forbid
// cudaMalloc.  No counterpart in the reference.
//
// The workspace (the 1024 partial sums of ||g||^2) lives in a file-scope `static __device__` array: every GPU that loads
// the library holds these 4 KB until the process exits.  In HOST code the array's name is only a host-side stand-in for
// the symbol, never a device address, so the address is resolved once per device with cudaGetSymbolAddress (that device
// current) and cached; the kernels receive the resolved pointer.  The entry point calls no cudaMalloc / cudaFree and no
// allocating library routine.
//
// Sharing: every call on a device uses the same array.  It keeps the previous call's partial sums, but nothing has to
// start at zero - all 1024 blocks of stage 1 overwrite their slot on every call (an empty element range writes 0), and stage 2
// reads only what stage 1 of the same call wrote, ordered by the stream.  Two calls that overlap on the GPU (different
// streams, or concurrent graph branches) race on the array: serialise them.
//
// Same arithmetic, block ranges and summation order as dsdf_clip_coef: the two entry points agree bit for bit.
#include <mutex>

#include "host_util.h"
#include "../../include/dsdf.h"

namespace dsdf {

namespace {

constexpr int kClipParts = 1024;                  // partial sums = stage-1 blocks (== dsdf_clip_coef's workspace)
constexpr int kMaxDevices = 64;

static __device__ __align__(256) float g_clip_static_scratch[kClipParts];

// stage 1: block b sums the squares of the fixed element range [b * per, (b + 1) * per)
__global__ void __launch_bounds__(256) sumsq_partial_static_kernel(const float* __restrict__ g, int64_t n, float* __restrict__ part) {
  __shared__ float red[256];
  const int64_t per = (n + gridDim.x - 1) / gridDim.x;
  const int64_t b = per * blockIdx.x;
  int64_t e = b + per;
  if (e > n) e = n;
  float acc = 0.f;
  for (int64_t i = b + threadIdx.x; i < e; i += 256) { const float t = g[i]; acc += t * t; }
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if (threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) part[blockIdx.x] = red[0];
}

// stage 2: fixed-order tree over the 1024 partials; out[0] = ||g||, out[1] = min(1, max_norm / (||g|| + 1e-6))
__global__ void __launch_bounds__(kClipParts) clip_coef_static_kernel(const float* __restrict__ part, float max_norm, float* __restrict__ out) {
  __shared__ float red[kClipParts];
  red[threadIdx.x] = part[threadIdx.x];
  __syncthreads();
  for (int w = kClipParts / 2; w > 0; w >>= 1) {
    if (threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float nrm = sqrtf(red[0]);
    out[0] = nrm;
    const float c = max_norm / (nrm + 1e-6f);
    out[1] = c < 1.0f ? c : 1.0f;
  }
}

std::mutex g_scratch_mutex;
float* g_scratch_addr[kMaxDevices] = {};          // device address of g_clip_static_scratch, per device ordinal

// Device address of the scratch array on the CURRENT device: resolved once per device, with that device current.
int scratch_for_current_device(float** out) {
  int dev = 0;
  DSDF_CHECK_CUDA(cudaGetDevice(&dev));
  DSDF_REQUIRE(dev >= 0 && dev < kMaxDevices, "dsdf_clip_coef_static: device ordinal beyond 63");
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (g_scratch_addr[dev] == nullptr) {
    float* p = nullptr;
    DSDF_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_clip_static_scratch));
    DSDF_REQUIRE(p != nullptr, "dsdf_clip_coef_static: cudaGetSymbolAddress returned a null address");
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return 0;
}

}  // namespace

}  // namespace dsdf

extern "C" int dsdf_clip_coef_static(const float* g, int64_t n, float max_norm, float* out, void* stream) {
  using namespace dsdf;
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(g != nullptr && out != nullptr && n > 0, "dsdf_clip_coef_static: bad arguments");
  if (n > int64_t(DSDF_CLIP_STATIC_MAX_N)) {
    snprintf(g_last_error, sizeof(g_last_error),
             "dsdf_clip_coef_static: n = %lld exceeds DSDF_CLIP_STATIC_MAX_N = %lld (use dsdf_clip_coef with a caller workspace)",
             static_cast<long long>(n), static_cast<long long>(DSDF_CLIP_STATIC_MAX_N));
    return 1;
  }
  float* scratch = nullptr;
  if (int e = scratch_for_current_device(&scratch)) return e;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DSDF_LAUNCHED;
  sumsq_partial_static_kernel<<<kClipParts, 256, 0, st>>>(g, n, scratch);
  DSDF_LAUNCHED;
  clip_coef_static_kernel<<<1, kClipParts, 0, st>>>(scratch, max_norm, out);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}
