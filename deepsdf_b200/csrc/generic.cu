// Generic path: Decoder forward / backward for ANY NetworkSpecs (small nets such as 4x32, 4x64, 6x128; LayerNorm;
// xyz_in_all; latent_dropout; geom_dimension != 3) as plain fp32 CUDA-core kernels on sm_100a.  These networks are
// launch/latency bound, not tensor bound, so the code favours exactness (fp32 everywhere, fixed
// reduction orders) over speed.  It is also the fp32 on-device cross-check of the bf16 tcgen05 path at full size.
// Replaces deep_sdf/networks/deep_sdf_decoder.py:76-111 and its autograd.
#include "common.cuh"
#include "host_util.h"
#include "../../include/dsdf.h"

namespace dsdf {

namespace {

constexpr uint32_t kLatentDropLayer = 0xFFFFu;   // Philox "layer" id of the latent_dropout stream
constexpr float kLnEps = 1e-5f;

inline size_t al(size_t x) { return (x + 63) / 64 * 64; }   // in floats

struct GenericLayout {
  size_t weff[DSDF_MAX_LAYERS], invn[DSDF_MAX_LAYERS], dweff;   // folded weights, 1/||v||, wgrad scratch (max layer)
  size_t x[DSDF_MAX_LAYERS];        // input of layer l: [N][in_dim[l]]
  size_t h[DSDF_MAX_LAYERS];        // pre-LayerNorm output of layer l (only if LN): [N][out]
  size_t mean[DSDF_MAX_LAYERS], rstd[DSDF_MAX_LAYERS];
  size_t ylast, t1;                 // pre-tanh output, first tanh
  size_t da, db;                    // backward ping-pong [N][max_w]
  size_t partial;                   // wgrad split partials [splits][max_wsize]
  size_t colpart;                   // column-sum partials [CS_SPLITS][max_w]
  size_t total;
  int splits;
  int max_w;
  size_t max_wsize;
};

constexpr int CS_SPLITS = 64;

GenericLayout make_layout(const dsdf_net_desc* d, int64_t N) {
  GenericLayout g;
  memset(&g, 0, sizeof(g));
  size_t off = 0;
  const int nl = d->num_linear;
  g.max_w = 1; g.max_wsize = 1;
  for (int l = 0; l < nl; ++l) {
    g.weff[l] = off; off += al(size_t(d->out_dim[l]) * d->in_dim[l]);
    g.invn[l] = off; off += al(d->out_dim[l]);
    if (d->in_dim[l] > g.max_w) g.max_w = d->in_dim[l];
    if (d->out_dim[l] > g.max_w) g.max_w = d->out_dim[l];
    size_t ws = size_t(d->out_dim[l]) * d->in_dim[l];
    if (ws > g.max_wsize) g.max_wsize = ws;
  }
  g.dweff = off; off += al(g.max_wsize);
  for (int l = 0; l < nl; ++l) {
    g.x[l] = off; off += al(size_t(N) * d->in_dim[l]);
    if ((d->layernorm_mask >> l) & 1) {
      g.h[l] = off; off += al(size_t(N) * d->out_dim[l]);
      g.mean[l] = off; off += al(N);
      g.rstd[l] = off; off += al(N);
    }
  }
  g.ylast = off; off += al(N);
  g.t1 = off; off += al(N);
  g.da = off; off += al(size_t(N) * g.max_w);
  g.db = off; off += al(size_t(N) * g.max_w);
  int splits = int((N + 4095) / 4096);
  if (splits < 1) splits = 1;
  if (splits > 32) splits = 32;
  g.splits = splits;
  g.partial = off; off += al(size_t(splits) * g.max_wsize);
  g.colpart = off; off += al(size_t(CS_SPLITS) * g.max_w);
  g.total = off;
  return g;
}

// ------------------------------------------------------------------------------------------ kernels
// W = g * v / ||v|| per output row (torch._weight_norm, dim=0).  One block (128 threads) per row.
__global__ void fold_kernel(const float* __restrict__ v, const float* __restrict__ g, int out, int in, float* __restrict__ weff,
                            float* __restrict__ invn) {
  const int o = blockIdx.x;
  __shared__ float red[4];
  __shared__ float s_scale;
  const float* row = v + size_t(o) * in;
  float ss = 0.f;
  for (int k = threadIdx.x; k < in; k += blockDim.x) ss += row[k] * row[k];
  for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = (red[0] + red[1]) + (red[2] + red[3]);
    float inv = 1.f, sc = 1.f;
    if (g != nullptr) { inv = 1.f / sqrtf(tot); sc = g[o] * inv; }
    invn[o] = inv;
    s_scale = sc;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < in; k += blockDim.x) weff[size_t(o) * in + k] = row[k] * s_scale;
}

// C[i][j] = sum_{p in split} A(i,p) * B(p,j) (+ bias[j]);  A(i,p) = a[i*sai + p*sap], B(p,j) = b[p*sbp + j*sbj].
// blockIdx.z selects the split of the p range; its output goes to c + z*c_split.  64x64 tile, 16 deep, 256 threads.
__global__ void __launch_bounds__(256) gemm_kernel(const float* __restrict__ a, int64_t sai, int64_t sap,
                                                   const float* __restrict__ b, int64_t sbp, int64_t sbj,
                                                   float* __restrict__ c, int64_t ldc, int64_t c_split,
                                                   const float* __restrict__ bias, int64_t I, int J, int64_t P, int64_t p_per_split) {
  __shared__ float As[16][65];
  __shared__ float Bs[16][65];
  const int64_t i0 = int64_t(blockIdx.y) * 64;
  const int j0 = blockIdx.x * 64;
  const int64_t pbeg = int64_t(blockIdx.z) * p_per_split;
  int64_t pend = pbeg + p_per_split;
  if (pend > P) pend = P;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4] = {};
  for (int64_t p0 = pbeg; p0 < pend; p0 += 16) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      int e = threadIdx.x + t * 256;
      {  // A tile: 64 (i) x 16 (p).  Map so that the faster-varying global index is contiguous across threads.
        int ii, pp;
        if (sap == 1) { pp = e & 15; ii = e >> 4; } else { ii = e & 63; pp = e >> 6; }
        int64_t gi = i0 + ii, gp = p0 + pp;
        As[pp][ii] = (gi < I && gp < pend) ? a[gi * sai + gp * sap] : 0.f;
      }
      {  // B tile: 16 (p) x 64 (j)
        int jj, pp;
        if (sbj == 1) { jj = e & 63; pp = e >> 6; } else { pp = e & 15; jj = e >> 4; }
        int gj = j0 + jj; int64_t gp = p0 + pp;
        Bs[pp][jj] = (gj < J && gp < pend) ? b[gp * sbp + int64_t(gj) * sbj] : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int pp = 0; pp < 16; ++pp) {
      float av[4], bv[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) { av[t] = As[pp][ty * 4 + t]; bv[t] = Bs[pp][tx * 4 + t]; }
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int s = 0; s < 4; ++s) acc[r][s] += av[r] * bv[s];
    }
    __syncthreads();
  }
  float* cz = c + int64_t(blockIdx.z) * c_split;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    int64_t gi = i0 + ty * 4 + r;
    if (gi >= I) continue;
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      int gj = j0 + tx * 4 + s;
      if (gj >= J) continue;
      float vv = acc[r][s];
      if (bias != nullptr) vv += bias[gj];
      cz[gi * ldc + gj] = vv;
    }
  }
}

// out[e] = sum_s part[s][e] in fixed order (deterministic)
__global__ void reduce_partials_kernel(const float* __restrict__ part, int splits, size_t stride, size_t n, float* __restrict__ out) {
  size_t e = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= n) return;
  float s = 0.f;
  for (int k = 0; k < splits; ++k) s += part[size_t(k) * stride + e];
  out[e] = s;
}

// x[r][c] (ld) = relu(x) * dropout, c < width, in place
__global__ void relu_dropout_kernel(float* __restrict__ x, int64_t N, int width, int64_t ld, int do_drop, uint64_t seed,
                                    uint32_t layer, uint64_t row0, uint32_t thr, float scale) {
  int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= N * width) return;
  int64_t r = idx / width; int c = int(idx % width);
  float v = fmaxf(x[r * ld + c], 0.f);
  if (do_drop) v = dropout_keep(seed, layer, row0 + uint64_t(r), uint32_t(c), thr) ? v * scale : 0.f;
  x[r * ld + c] = v;
}

// dst[r][dcol0 + c] = src[r][scol0 + c] (optionally latent-dropout on the first `ndrop` columns)
__global__ void copy_cols_kernel(const float* __restrict__ src, int64_t lds, int scol0, float* __restrict__ dst, int64_t ldd,
                                 int dcol0, int64_t N, int width, int add, int ndrop, uint64_t seed, uint64_t row0, uint32_t thr,
                                 float scale) {
  int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= N * width) return;
  int64_t r = idx / width; int c = int(idx % width);
  float v = src[r * lds + scol0 + c];
  if (c < ndrop) v = dropout_keep(seed, kLatentDropLayer, row0 + uint64_t(r), uint32_t(c), thr) ? v * scale : 0.f;
  float* d = dst + r * ldd + dcol0 + c;
  *d = add ? (*d + v) : v;
}

// LayerNorm forward: one warp per row.  y written to dst (ld ldd), stats saved.
__global__ void layernorm_fwd_kernel(const float* __restrict__ h, int64_t N, int width, const float* __restrict__ w,
                                     const float* __restrict__ b, float* __restrict__ dst, int64_t ldd, float* __restrict__ mean,
                                     float* __restrict__ rstd) {
  int64_t r = int64_t(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= N) return;
  const int lane = threadIdx.x & 31;
  const float* row = h + r * width;
  float s = 0.f;
  for (int c = lane; c < width; c += 32) s += row[c];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mu = s / width;
  float v = 0.f;
  for (int c = lane; c < width; c += 32) { float t = row[c] - mu; v += t * t; }
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const float rs = rsqrtf(v / width + kLnEps);
  if (lane == 0) { mean[r] = mu; rstd[r] = rs; }
  for (int c = lane; c < width; c += 32) dst[r * ldd + c] = (row[c] - mu) * rs * w[c] + b[c];
}

// LayerNorm backward (input grad), in place on dy (ld): dy <- dh.  One warp per row.
__global__ void layernorm_bwd_kernel(float* __restrict__ dy, int64_t ld, const float* __restrict__ h, int64_t N, int width,
                                     const float* __restrict__ w, const float* __restrict__ mean, const float* __restrict__ rstd) {
  int64_t r = int64_t(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= N) return;
  const int lane = threadIdx.x & 31;
  const float mu = mean[r], rs = rstd[r];
  float s1 = 0.f, s2 = 0.f;
  for (int c = lane; c < width; c += 32) {
    float xh = (h[r * width + c] - mu) * rs;
    float dxh = dy[r * ld + c] * w[c];
    s1 += dxh; s2 += dxh * xh;
  }
  for (int o = 16; o > 0; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
  s1 /= width; s2 /= width;
  for (int c = lane; c < width; c += 32) {
    float xh = (h[r * width + c] - mu) * rs;
    float dxh = dy[r * ld + c] * w[c];
    dy[r * ld + c] = rs * (dxh - s1 - xh * s2);
  }
}

// partial column sums: part[split][c] = sum_{r in split} a[r][c] * (mode 1: xhat(r,c); mode 0: 1).  block (32, 8)
__global__ void colsum_partial_kernel(const float* __restrict__ a, int64_t lda, const float* __restrict__ h,
                                      const float* __restrict__ mean, const float* __restrict__ rstd, int64_t N, int width,
                                      float* __restrict__ part, int64_t rows_per_split) {
  const int c = blockIdx.x * 32 + threadIdx.x;
  const int64_t rbeg = int64_t(blockIdx.y) * rows_per_split;
  int64_t rend = rbeg + rows_per_split;
  if (rend > N) rend = N;
  float s = 0.f;
  if (c < width) {
    for (int64_t r = rbeg + threadIdx.y; r < rend; r += 8) {
      float v = a[r * lda + c];
      if (h != nullptr) v *= (h[r * width + c] - mean[r]) * rstd[r];
      s += v;
    }
  }
  __shared__ float red[8][33];
  red[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < width) {
    float t = 0.f;
    for (int k = 0; k < 8; ++k) t += red[k][threadIdx.x];
    part[size_t(blockIdx.y) * width + c] = t;
  }
}

// final: out[c] (+)= sum_s part[s][c]
__global__ void colsum_final_kernel(const float* __restrict__ part, int splits, int width, float* __restrict__ out, int accumulate) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= width) return;
  float s = 0.f;
  for (int k = 0; k < splits; ++k) s += part[size_t(k) * width + c];
  out[c] = accumulate ? out[c] + s : s;
}

// final activation: y -> tanh (twice when use_tanh)
__global__ void final_fwd_kernel(const float* __restrict__ y, int64_t N, int use_tanh, float* __restrict__ t1, float* __restrict__ out) {
  int64_t r = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (r >= N) return;
  float a = tanhf(y[r]);
  t1[r] = a;
  out[r] = use_tanh ? tanhf(a) : a;
}
__global__ void final_bwd_kernel(const float* __restrict__ gout, const float* __restrict__ t1, int64_t N, int use_tanh,
                                 float* __restrict__ dy) {
  int64_t r = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (r >= N) return;
  float a = t1[r];
  float d = gout[r] * (1.f - a * a);
  if (use_tanh) { float t2 = tanhf(a); d *= (1.f - t2 * t2); }
  dy[r] = d;
}

// dY_{l-1}[r][c] = dX[r][c] * act'(a) where a = X_l[r][c] (post relu*dropout activation), in place on dx
__global__ void act_bwd_kernel(float* __restrict__ dx, int64_t ldx, const float* __restrict__ a, int64_t lda, int64_t N,
                               int width, float scale) {
  int64_t idx = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= N * width) return;
  int64_t r = idx / width; int c = int(idx % width);
  float g = dx[r * ldx + c];
  dx[r * ldx + c] = (a[r * lda + c] > 0.f) ? g * scale : 0.f;
}

// weight-norm backward for one layer.  One block per out row.  dg = sum dW*v*invn ; dv = g*invn*(dW - v*invn*dg)
__global__ void wn_bwd_kernel(const float* __restrict__ dW, const float* __restrict__ v, const float* __restrict__ g,
                              const float* __restrict__ invn, int in, float* __restrict__ dv, float* __restrict__ dg, int accumulate) {
  const int o = blockIdx.x;
  __shared__ float red[4];
  __shared__ float s_dg;
  const float inv = invn[o];
  float s = 0.f;
  for (int k = threadIdx.x; k < in; k += blockDim.x) s += dW[size_t(o) * in + k] * v[size_t(o) * in + k];
  for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = ((red[0] + red[1]) + (red[2] + red[3])) * inv;
    s_dg = t;
    dg[o] = accumulate ? dg[o] + t : t;
  }
  __syncthreads();
  const float dgo = s_dg, gi = g[o] * inv;
  for (int k = threadIdx.x; k < in; k += blockDim.x) {
    size_t e = size_t(o) * in + k;
    float t = gi * (dW[e] - v[e] * inv * dgo);
    dv[e] = accumulate ? dv[e] + t : t;
  }
}
__global__ void copy_or_add_kernel(const float* __restrict__ src, float* __restrict__ dst, size_t n, int accumulate) {
  size_t e = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e < n) dst[e] = accumulate ? dst[e] + src[e] : src[e];
}

inline unsigned nblk(int64_t n, int bs) { return unsigned((n + bs - 1) / bs); }

int launch_gemm(cudaStream_t st, const float* a, int64_t sai, int64_t sap, const float* b, int64_t sbp, int64_t sbj, float* c,
                int64_t ldc, int64_t c_split, const float* bias, int64_t I, int J, int64_t P, int splits) {
  int64_t pps = (P + splits - 1) / splits;
  pps = (pps + 15) / 16 * 16;
  dim3 grid((J + 63) / 64, unsigned((I + 63) / 64), splits);
  DSDF_REQUIRE(grid.y <= 65535u * 1u || true, "grid");
  if (grid.y > 65535) return set_error("generic path: too many rows per call (chunk the batch below 4M rows)");
  gemm_kernel<<<grid, 256, 0, st>>>(a, sai, sap, b, sbp, sbj, c, ldc, c_split, bias, I, J, P, pps);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int fold_all(const dsdf_net_desc* d, const dsdf_params* p, float* ws, const GenericLayout& g, cudaStream_t st) {
  for (int l = 0; l < d->num_linear; ++l) {
    const float* gl = ((d->weight_norm_mask >> l) & 1) ? p->g[l] : nullptr;
    if (((d->weight_norm_mask >> l) & 1) && gl == nullptr) return set_error("weight-normed layer without g pointer");
    fold_kernel<<<d->out_dim[l], 128, 0, st>>>(p->v[l], gl, d->out_dim[l], d->in_dim[l], ws + g.weff[l], ws + g.invn[l]);
  }
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace
}  // namespace dsdf

using namespace dsdf;

extern "C" size_t dsdf_generic_workspace_bytes(const dsdf_net_desc* desc, int64_t n_rows) {
  return make_layout(desc, n_rows).total * sizeof(float);
}

extern "C" int dsdf_generic_forward(const dsdf_net_desc* d, const dsdf_params* p, const float* input, int64_t N,
                                    int32_t training, uint64_t seed, uint64_t row0, float* out, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(N > 0, "n_rows must be positive");
  const GenericLayout g = make_layout(d, N);
  DSDF_REQUIRE(workspace_bytes >= g.total * sizeof(float), "generic workspace too small");
  float* ws = static_cast<float*>(workspace);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nl = d->num_linear, L = d->latent_size, D = d->geom_dim, LD = L + D;
  if (int e = fold_all(d, p, ws, g, st)) return e;
  const uint32_t thr = dropout_threshold(d->dropout_prob);
  const float scale = dropout_scale(d->dropout_prob);   // 1 / (1 - effective p), see common.cuh
  const uint32_t thr_lat = dropout_threshold(0.2f);
  const float scale_lat = dropout_scale(0.2f);
  // x0 = [dropout(latent), xyz]  (deep_sdf_decoder.py:77-84)
  const int ndrop = (d->latent_dropout && training && LD > D) ? L : 0;
  copy_cols_kernel<<<nblk(N * LD, 256), 256, 0, st>>>(input, LD, 0, ws + g.x[0], LD, 0, N, LD, 0, ndrop, seed, row0, thr_lat, scale_lat);
  for (int l = 0; l < nl; ++l) {
    const int in = d->in_dim[l], outd = d->out_dim[l];
    const bool last = (l == nl - 1);
    const bool ln = (d->layernorm_mask >> l) & 1;
    float* dst; int64_t ldd;
    if (last) { dst = ws + g.ylast; ldd = 1; }
    else if (ln) { dst = ws + g.h[l]; ldd = outd; }
    else { dst = ws + g.x[l + 1]; ldd = d->in_dim[l + 1]; }
    // y = x W^T + b   (deep_sdf_decoder.py:92)
    if (int e = launch_gemm(st, ws + g.x[l], in, 1, ws + g.weff[l], 1, in, dst, ldd, 0, p->b[l], N, outd, in, 1)) return e;
    if (last) break;
    float* xn = ws + g.x[l + 1];
    const int64_t ldn = d->in_dim[l + 1];
    if (ln) {
      DSDF_REQUIRE(p->ln_w[l] && p->ln_b[l], "LayerNorm parameters missing");
      layernorm_fwd_kernel<<<nblk(N, 8), 256, 0, st>>>(ws + g.h[l], N, outd, p->ln_w[l], p->ln_b[l], xn, ldn, ws + g.mean[l], ws + g.rstd[l]);
    }
    const int do_drop = (training && ((d->dropout_mask >> l) & 1)) ? 1 : 0;
    relu_dropout_kernel<<<nblk(N * outd, 256), 256, 0, st>>>(xn, N, outd, ldn, do_drop, seed, uint32_t(l), row0, thr, scale);
    // concatenations feeding layer l+1  (deep_sdf_decoder.py:88-91)
    if ((d->latent_in_mask >> (l + 1)) & 1) {
      DSDF_REQUIRE(outd + LD == ldn, "latent_in layer width mismatch");
      copy_cols_kernel<<<nblk(N * LD, 256), 256, 0, st>>>(input, LD, 0, xn, ldn, outd, N, LD, 0, 0, 0, 0, 0, 1.f);
    } else if (d->xyz_in_all) {
      DSDF_REQUIRE(outd + D == ldn, "xyz_in_all layer width mismatch");
      copy_cols_kernel<<<nblk(N * D, 256), 256, 0, st>>>(input, LD, L, xn, ldn, outd, N, D, 0, 0, 0, 0, 0, 1.f);
    } else {
      DSDF_REQUIRE(outd == ldn, "layer width mismatch");
    }
  }
  final_fwd_kernel<<<nblk(N, 256), 256, 0, st>>>(ws + g.ylast, N, d->use_tanh, ws + g.t1, out);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_generic_backward(const dsdf_net_desc* d, const dsdf_params* p, const float* input, int64_t N,
                                     int32_t training, uint64_t seed, uint64_t row0, const float* grad_out,
                                     float* grad_input, const dsdf_param_grads* pg, int32_t accumulate, void* workspace,
                                     size_t workspace_bytes, void* stream) {
  if (int e = require_sm100()) return e;
  const GenericLayout g = make_layout(d, N);
  DSDF_REQUIRE(workspace_bytes >= g.total * sizeof(float), "generic workspace too small");
  float* ws = static_cast<float*>(workspace);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nl = d->num_linear, L = d->latent_size, D = d->geom_dim, LD = L + D;
  const uint32_t thr = dropout_threshold(d->dropout_prob);
  const float scale = dropout_scale(d->dropout_prob);   // 1 / (1 - effective p), see common.cuh
  const uint32_t thr_lat = dropout_threshold(0.2f);
  const float scale_lat = dropout_scale(0.2f);
  if (grad_input) DSDF_CHECK_CUDA(cudaMemsetAsync(grad_input, 0, size_t(N) * LD * sizeof(float), st));
  float* dy = ws + g.da;      // dY of the current layer, ld = ldy
  float* dx = ws + g.db;
  int64_t ldy = 1;
  final_bwd_kernel<<<nblk(N, 256), 256, 0, st>>>(grad_out, ws + g.t1, N, d->use_tanh, dy);
  const int64_t rows_per_cs = (N + CS_SPLITS - 1) / CS_SPLITS;
  for (int l = nl - 1; l >= 0; --l) {
    const int in = d->in_dim[l], outd = d->out_dim[l];
    const bool wn = (d->weight_norm_mask >> l) & 1;
    if (pg) {
      // dW = dY^T X   (split over rows, fixed-order reduction)
      if (int e = launch_gemm(st, dy, 1, ldy, ws + g.x[l], in, 1, ws + g.partial, in, int64_t(g.max_wsize), nullptr, outd, in, N, g.splits)) return e;
      const size_t wsz = size_t(outd) * in;
      reduce_partials_kernel<<<nblk(wsz, 256), 256, 0, st>>>(ws + g.partial, g.splits, g.max_wsize, wsz, ws + g.dweff);
      if (wn) {
        DSDF_REQUIRE(pg->v[l] && pg->g[l], "missing grad pointers for weight-normed layer");
        wn_bwd_kernel<<<outd, 128, 0, st>>>(ws + g.dweff, p->v[l], p->g[l], ws + g.invn[l], in, pg->v[l], pg->g[l], accumulate);
      } else {
        DSDF_REQUIRE(pg->v[l], "missing weight grad pointer");
        copy_or_add_kernel<<<nblk(wsz, 256), 256, 0, st>>>(ws + g.dweff, pg->v[l], wsz, accumulate);
      }
      DSDF_REQUIRE(pg->b[l], "missing bias grad pointer");
      colsum_partial_kernel<<<dim3((outd + 31) / 32, CS_SPLITS), dim3(32, 8), 0, st>>>(dy, ldy, nullptr, nullptr, nullptr, N, outd, ws + g.colpart, rows_per_cs);
      colsum_final_kernel<<<nblk(outd, 128), 128, 0, st>>>(ws + g.colpart, CS_SPLITS, outd, pg->b[l], accumulate);
    }
    if (l == 0 && !grad_input) break;
    // dX = dY W
    if (int e = launch_gemm(st, dy, ldy, 1, ws + g.weff[l], in, 1, dx, in, 0, nullptr, N, in, outd, 1)) return e;
    if (l == 0) {
      const int ndrop = (d->latent_dropout && training && LD > D) ? L : 0;
      copy_cols_kernel<<<nblk(N * LD, 256), 256, 0, st>>>(dx, in, 0, grad_input, LD, 0, N, LD, 1, ndrop, seed, row0, thr_lat, scale_lat);
      break;
    }
    const int prev_out = d->out_dim[l - 1];
    if (grad_input) {
      if ((d->latent_in_mask >> l) & 1)
        copy_cols_kernel<<<nblk(N * LD, 256), 256, 0, st>>>(dx, in, prev_out, grad_input, LD, 0, N, LD, 1, 0, 0, 0, 0, 1.f);
      else if (d->xyz_in_all)
        copy_cols_kernel<<<nblk(N * D, 256), 256, 0, st>>>(dx, in, prev_out, grad_input, LD, L, N, D, 1, 0, 0, 0, 0, 1.f);
    }
    // through relu * dropout of hidden layer l-1 (its output is the first prev_out columns of x[l])
    const float sc = (training && ((d->dropout_mask >> (l - 1)) & 1)) ? scale : 1.f;
    act_bwd_kernel<<<nblk(N * prev_out, 256), 256, 0, st>>>(dx, in, ws + g.x[l], in, N, prev_out, sc);
    if ((d->layernorm_mask >> (l - 1)) & 1) {
      if (pg) {
        DSDF_REQUIRE(pg->ln_w[l - 1] && pg->ln_b[l - 1], "missing LayerNorm grad pointers");
        colsum_partial_kernel<<<dim3((prev_out + 31) / 32, CS_SPLITS), dim3(32, 8), 0, st>>>(dx, in, ws + g.h[l - 1], ws + g.mean[l - 1], ws + g.rstd[l - 1], N, prev_out, ws + g.colpart, rows_per_cs);
        colsum_final_kernel<<<nblk(prev_out, 128), 128, 0, st>>>(ws + g.colpart, CS_SPLITS, prev_out, pg->ln_w[l - 1], accumulate);
        colsum_partial_kernel<<<dim3((prev_out + 31) / 32, CS_SPLITS), dim3(32, 8), 0, st>>>(dx, in, nullptr, nullptr, nullptr, N, prev_out, ws + g.colpart, rows_per_cs);
        colsum_final_kernel<<<nblk(prev_out, 128), 128, 0, st>>>(ws + g.colpart, CS_SPLITS, prev_out, pg->ln_b[l - 1], accumulate);
      }
      layernorm_bwd_kernel<<<nblk(N, 8), 256, 0, st>>>(dx, in, ws + g.h[l - 1], N, prev_out, p->ln_w[l - 1], ws + g.mean[l - 1], ws + g.rstd[l - 1]);
    }
    // next iteration: dY = dx (ld = in), swap buffers
    float* t = dy; dy = dx; dx = t;
    ldy = in;
  }
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}
