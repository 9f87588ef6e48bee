// Pre/post kernels of the microstructure evaluation (deep_sdf/mesh.py:457-514): the queries of `evaluate_network` are the
// voxel-grid vertices folded into one unit cell by the TILING TRANSFORM (`prepare_samples`, mesh.py:470-477), joined with a
// per-vertex latent code; after the decoder the SDF is CAPPED at the requested domain borders and outside the unit cube
// (mesh.py:499-514).  Element-wise, HBM-bound; fp32 with explicit roundings so that a torch restatement is bit-exact.
#include "common.cuh"
#include "host_util.h"

namespace dsdf {
namespace {

// torch.remainder for floats (sign of the divisor): c10's  fmod + fix-up
__device__ __forceinline__ float torch_remainder(float a, float b) {
  float m = fmodf(a, b);
  if (m != 0.0f && ((b < 0.0f) != (m < 0.0f))) m = __fadd_rn(m, b);
  return m;
}

struct TileArgs {
  float shift[3], period2[3], half[3], gain[3];      // t % 2, 2p, p, 2/p   with p = 2/t   (mesh.py:470-472)
};

// queries[r] = [lat[r, 0:L], transform(x), transform(y), transform(z)]; rows [n, n_pad) are zero-filled
__global__ void tile_queries_kernel(const float* __restrict__ samples_orig, int32_t ld_s, const float* __restrict__ lat, int32_t ld_l,
                                    int32_t L, int64_t n, int64_t n_pad, TileArgs t, float* __restrict__ queries) {
  const int64_t r = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (r >= n_pad) return;
  float* q = queries + r * (L + 3);
  if (r >= n) {
    for (int k = 0; k < L + 3; ++k) q[k] = 0.0f;
    return;
  }
  for (int k = 0; k < L; ++k) q[k] = __ldg(lat + r * ld_l + k);
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const float x = __ldg(samples_orig + r * ld_s + d);
    const float m = torch_remainder(__fsub_rn(x, t.shift[d]), t.period2[d]);
    q[L + d] = __fsub_rn(__fmul_rn(t.gain[d], fabsf(__fsub_rn(m, t.half[d]))), 1.0f);
  }
}

struct CapArgs {
  int32_t n_caps;
  int32_t dim[6];
  float offset[6], neg_mult[6];                      // multiplier * (1 - measure) (formed in double like Python), -multiplier
  int32_t cap[6];                                    // -1 or +1 (mesh.py:503-508)
};

// sdf <- caps(sdf); where a cap decides the value the latent derivative is zero: G[r, 0:L] = 0
__global__ void cap_sdf_kernel(float* __restrict__ sdf, float* __restrict__ G, int64_t ldg, int32_t L, const float* __restrict__ samples_orig,
                               int32_t ld_s, int64_t n, CapArgs c) {
  const int64_t r = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (r >= n) return;
  const float p[3] = {__ldg(samples_orig + r * ld_s), __ldg(samples_orig + r * ld_s + 1), __ldg(samples_orig + r * ld_s + 2)};
  float s = sdf[r];
  bool capped = false;
  for (int i = 0; i < c.n_caps; ++i) {
    // border_sdf = (x - multiplier * (1 - measure)) * -multiplier            (mesh.py:502)
    const float b = __fmul_rn(__fsub_rn(p[c.dim[i]], c.offset[i]), c.neg_mult[i]);
    if (c.cap[i] < 0) { if (-b > s) { s = -b; capped = true; } }            // torch.maximum(sdf, -border)
    else              { if (b < s) { s = b; capped = true; } }              // torch.minimum(sdf, border)
  }
#pragma unroll
  for (int d = 0; d < 3; ++d)                                               // everything outside the unit cube (mesh.py:512-514)
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const float m = k ? 1.0f : -1.0f;
      const float nb = -__fmul_rn(__fsub_rn(p[d], m), -m);
      if (nb > s) { s = nb; capped = true; }
    }
  sdf[r] = s;
  if (capped && G != nullptr)
    for (int k = 0; k < L; ++k) G[r * ldg + k] = 0.0f;
}

}  // namespace
}  // namespace dsdf

using namespace dsdf;

extern "C" int dsdf_tile_queries(const float* samples_orig, int32_t ld_s, const float* lat, int32_t ld_l, int32_t latent_size,
                                 int64_t n, int64_t n_pad, const float* tiling3, float* queries, void* stream) {
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(samples_orig && lat && queries && tiling3 && n >= 0 && n_pad >= n && ld_s >= 3 && ld_l >= latent_size,
               "tile_queries: bad arguments");
  if (n_pad == 0) return 0;
  TileArgs t;
  for (int d = 0; d < 3; ++d) {
    const double tt = double(tiling3[d]);
    DSDF_REQUIRE(tt > 0.0, "tiling must be positive");
    const double p = 2.0 / tt;                       // Python doubles, each rounded once when it meets the fp32 tensor
    t.shift[d] = float(fmod(tt, 2.0));
    t.period2[d] = float(p * 2.0);
    t.half[d] = float(p);
    t.gain[d] = float(2.0 / p);
  }
  DSDF_LAUNCHED;
  tile_queries_kernel<<<unsigned((n_pad + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(samples_orig, ld_s, lat, ld_l,
                                                                                                   latent_size, n, n_pad, t, queries);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_cap_sdf(float* sdf, float* G, int64_t ldg, int32_t latent_size, const float* samples_orig, int32_t ld_s, int64_t n,
                            int32_t n_caps, const int32_t* dims, const double* multipliers, const double* measures,
                            const int32_t* caps, void* stream) {
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(sdf && samples_orig && n >= 0 && n_caps >= 0 && n_caps <= 6 && ld_s >= 3, "cap_sdf: bad arguments");
  if (n == 0) return 0;
  CapArgs c;
  c.n_caps = n_caps;
  for (int i = 0; i < n_caps; ++i) {
    DSDF_REQUIRE(dims[i] >= 0 && dims[i] < 3 && (caps[i] == 1 || caps[i] == -1), "Cap must be -1 or 1");   // mesh.py:509
    c.dim[i] = dims[i]; c.offset[i] = float(multipliers[i] * (1.0 - measures[i])); c.neg_mult[i] = float(-multipliers[i]);
    c.cap[i] = caps[i];
  }
  DSDF_LAUNCHED;
  cap_sdf_kernel<<<unsigned((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(sdf, G, ldg, latent_size, samples_orig, ld_s, n, c);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}
