// Small fp32 kernels that close the fused training step: deterministic reductions of the per-tile / per-split partials,
// assembly of the effective-weight gradient, weight-norm backward, latent-code gradient (+ code regulariser), loss, Adam.
// Replaces train_deep_sdf.py:519-533 (loss assembly, regulariser) and :545 (optimizer_all.step()).
#pragma once
#include "common.cuh"
#include "fused.cuh"
#include "fused_host.h"

namespace dsdf {

struct FinishArgs {
  dsdf_net_desc desc;
  FusedLayout lo;
  dsdf_params params;
  dsdf_param_grads grads;
  const uint8_t* pack;
  const float* tile_part;       // [n_tiles][n_slots][512]
  const int32_t* tile_scene;    // [n_tiles]
  int32_t n_tiles, n_slots, n_scenes;
  float* scene_sums;            // [S][n_slots][512]
  float* scene_reg;             // [S] rows_s * ||z_s|| (written by latent_grad_kernel)
  const float* wg_part;         // wgrad partials
  WgradPlan plan;               // wgrad segments: partial [4 x 128 outputs][n] per segment
  const float* last_part;       // [n_last_blocks][513]
  int32_t n_last_blocks;
  const float* lat_table;       // [S_total][L]
  const int64_t* scene_ids;     // [S]
  float* lat_grad;              // [S_total][L] dense
  float reg_scale;              // lambda * min(1, epoch/100) * inv_n   (0: no regulariser)
  float inv_n;
  const float* tile_loss;       // [n_tiles]
  float* loss_out;              // [2]: total loss, data term
  int32_t xyz_slot0;            // first xyz-sum slot of lin0 (3 slots), skip layer follows (+3)
  int32_t accumulate;           // != 0: add to the parameter / latent gradients and to the loss (batch_split chunks,
                                // train_deep_sdf.py:495-535) instead of overwriting them
  float in_scale[DSDF_MAX_LAYERS + 1];   // the activation stash is UNSCALED (relu * keep): the dropout scale of hidden layer
                                         // l-1 multiplies dW of layer l's activation columns here (1 when dropout is off)
};

// scene_sums[s][slot][c] = sum over the tiles of scene s (ascending tile order) of tile_part[tile][slot][c].
// grid (S, n_slots), block 512.  The tile list of the scene is compacted (in order) 512 tiles at a time with ballots.
__global__ void __launch_bounds__(FH) scene_sums_kernel(const float* __restrict__ tile_part, const int32_t* __restrict__ tile_scene,
                                                        int n_tiles, int n_slots, float* __restrict__ scene_sums) {
  const int s = blockIdx.x, slot = blockIdx.y, c = threadIdx.x;
  const int warp = c >> 5, lane = c & 31;
  __shared__ int s_list[FH];
  __shared__ int s_wcount[FH / 32];
  float acc = 0.f;
  for (int base = 0; base < n_tiles; base += FH) {
    const int t = base + c;
    const bool mine = (t < n_tiles) && (__ldg(tile_scene + t) == s);
    const unsigned bal = __ballot_sync(0xffffffffu, mine);
    if (lane == 0) s_wcount[warp] = __popc(bal);
    __syncthreads();
    int off = 0, total = 0;
#pragma unroll
    for (int w = 0; w < FH / 32; ++w) { const int n = s_wcount[w]; if (w < warp) off += n; total += n; }
    if (mine) s_list[off + __popc(bal & ((1u << lane) - 1u))] = t;
    __syncthreads();
    for (int i = 0; i < total; ++i) acc += __ldg(tile_part + (size_t(s_list[i]) * n_slots + slot) * FH + c);
    __syncthreads();
  }
  scene_sums[(size_t(s) * n_slots + slot) * FH + c] = acc;
}

// last_part[0][k] = sum_b last_part[b][k] (fixed order: warp w adds blocks w, w+32, ...; then warps 0..31 in order).
// grid ceil(513/32), block 1024: lane = column within the 32-column group -> 128-byte reads.  In place: a block only ever
// touches its own columns, and writes them after all its reads.
__global__ void __launch_bounds__(1024) last_reduce_kernel(float* __restrict__ last_part, int n_blocks) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;      // 32 warps: a quarter of the serial depth of 8
  const int k = blockIdx.x * 32 + lane;
  __shared__ float red[32][32];
  float acc = 0.f;
  if (k <= FH) {
#pragma unroll 8
    for (int b = warp; b < n_blocks; b += 32) acc += last_part[size_t(b) * (FH + 1) + k];
  }
  red[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && k <= FH) {
    float t = red[0][lane];
#pragma unroll
    for (int w = 1; w < 32; ++w) t += red[w][lane];
    last_part[k] = t;
  }
}

// One block per (output row o, layer l): assemble dW_eff[o, :], then weight-norm backward (or plain copy) and bias grad.
__global__ void __launch_bounds__(128) assemble_grads_kernel(const __grid_constant__ FinishArgs a) {
  const int l = blockIdx.y, o = blockIdx.x;
  const dsdf_net_desc& d = a.desc;
  if (o >= d.out_dim[l]) return;
  const int in = d.in_dim[l], nh = a.lo.num_hidden, L = d.latent_size, S = a.n_scenes;
  __shared__ float s_dw[FH + 8];
  __shared__ float red[4];
  __shared__ float s_dg;
  const float* v = a.params.v[l] + size_t(o) * in;
  float dot = 0.f;
  for (int k = threadIdx.x; k < in; k += blockDim.x) {
    float g = 0.f;
    if (l == nh) {
      for (int b = 0; b < a.n_last_blocks; ++b) g += a.last_part[size_t(b) * (FH + 1) + k];
      g *= a.in_scale[l];
    } else if (k < a.lo.xw[l]) {
      for (int sg = a.plan.layer_seg0[l]; sg < a.plan.layer_seg0[l + 1]; ++sg)
        g += a.wg_part[a.plan.seg[sg].part_offset + size_t(o) * a.plan.seg[sg].n + k];
      g *= a.in_scale[l];
    } else if (a.lo.lat_col0[l] >= 0 && k < a.lo.lat_col0[l] + L) {
      const int kk = k - a.lo.lat_col0[l];
      for (int s = 0; s < S; ++s)
        g += a.scene_sums[(size_t(s) * a.n_slots + l) * FH + o] * a.lat_table[size_t(a.scene_ids[s]) * L + kk];
    } else {
      const int dd = k - a.lo.xyz_col0[l];
      const int slot = a.xyz_slot0 + (l == 0 ? 0 : 3) + dd;
      for (int s = 0; s < S; ++s) g += a.scene_sums[(size_t(s) * a.n_slots + slot) * FH + o];
    }
    s_dw[k] = g;
    dot += g * v[k];
  }
  for (int off = 16; off > 0; off >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = dot;
  __syncthreads();
  const bool wn = (d.weight_norm_mask >> l) & 1;
  const float invn = reinterpret_cast<const float*>(a.pack + a.lo.off_invn[l])[o];
  if (threadIdx.x == 0) {
    const float dg = ((red[0] + red[1]) + (red[2] + red[3])) * invn;
    s_dg = dg;
    if (wn) a.grads.g[l][o] = (a.accumulate ? a.grads.g[l][o] : 0.f) + dg;
    float db = 0.f;
    if (l == nh) { for (int b = 0; b < a.n_last_blocks; ++b) db += a.last_part[size_t(b) * (FH + 1) + FH]; }
    else { for (int s = 0; s < S; ++s) db += a.scene_sums[(size_t(s) * a.n_slots + l) * FH + o]; }
    a.grads.b[l][o] = (a.accumulate ? a.grads.b[l][o] : 0.f) + db;
  }
  __syncthreads();
  float* dv = a.grads.v[l] + size_t(o) * in;
  if (wn) {
    const float gi = a.params.g[l][o] * invn, dg = s_dg;
    for (int k = threadIdx.x; k < in; k += blockDim.x) dv[k] = (a.accumulate ? dv[k] : 0.f) + gi * (s_dw[k] - v[k] * invn * dg);
  } else {
    for (int k = threadIdx.x; k < in; k += blockDim.x) dv[k] = (a.accumulate ? dv[k] : 0.f) + s_dw[k];
  }
}

// lat_grad[scene_ids[s]][k] = colsum_s(dY_0) . W0[:, k] + colsum_s(dY_skip) . Wskip[:, xw+k] + reg * rows_s * z/||z||
// grid (S, ceil(L/8)), block 256: warp w handles latent column k = 8*blockIdx.y + w; lanes stride the 512 outputs
// (coalesced wlatT reads, fixed shuffle-tree order -> deterministic).
__global__ void __launch_bounds__(256) latent_grad_kernel(const __grid_constant__ FinishArgs a) {
  const int s = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int L = a.desc.latent_size;
  const float* wlatT = reinterpret_cast<const float*>(a.pack + a.lo.off_wlatT);
  const float* ss0 = a.scene_sums + (size_t(s) * a.n_slots + 0) * FH;
  const float* ss1 = a.lo.lat_layer > 0 ? a.scene_sums + (size_t(s) * a.n_slots + a.lo.lat_layer) * FH : nullptr;
  const float* z = a.lat_table + size_t(a.scene_ids[s]) * L;
  __shared__ float s_norm;
  __shared__ int s_cnt[8];
  {
    int cnt = 0;
    for (int t = threadIdx.x; t < a.n_tiles; t += 256) cnt += (__ldg(a.tile_scene + t) == s);
    for (int off = 16; off > 0; off >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, off);
    if (lane == 0) s_cnt[warp] = cnt;
  }
  if (warp == 0) {
    float q = 0.f;
    for (int k = lane; k < L; k += 32) q += z[k] * z[k];
    for (int off = 16; off > 0; off >>= 1) q += __shfl_xor_sync(0xffffffffu, q, off);
    if (lane == 0) s_norm = sqrtf(q);
  }
  __syncthreads();
  int rows = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) rows += s_cnt[w];
  rows *= TILE_M;
  if (blockIdx.y == 0 && threadIdx.x == 0) a.scene_reg[s] = s_norm * float(rows);
  const int k = blockIdx.y * 8 + warp;
  if (k >= L) return;
  float acc = 0.f;
  for (int o = lane; o < FH; o += 32) {
    acc += ss0[o] * wlatT[size_t(k) * FH + o];
    if (ss1) acc += ss1[o] * wlatT[(size_t(L) + k) * FH + o];
  }
  for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if (lane == 0) {
    if (a.reg_scale != 0.f && s_norm > 0.f) acc += a.reg_scale * float(rows) * z[k] / s_norm;
    float* dst = a.lat_grad + size_t(a.scene_ids[s]) * L + k;
    *dst = (a.accumulate ? *dst : 0.f) + acc;
  }
}

// loss_out[0] = inv_n * sum(tile_loss) + reg ; loss_out[1] = data term.  Single block, fixed summation order.
__global__ void __launch_bounds__(256) loss_kernel(const __grid_constant__ FinishArgs a) {
  __shared__ float red[256];
  float acc = 0.f;
  for (int t = threadIdx.x; t < a.n_tiles; t += 256) acc += a.tile_loss[t];
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) {
    if (threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float data = red[0] * a.inv_n;
    float reg = 0.f;
    if (a.reg_scale != 0.f) {
      for (int s = 0; s < a.n_scenes; ++s) reg += a.scene_reg[s];
      reg *= a.reg_scale;
    }
    a.loss_out[0] = (a.accumulate ? a.loss_out[0] : 0.f) + data + reg;
    a.loss_out[1] = (a.accumulate ? a.loss_out[1] : 0.f) + data;
  }
}

// torch.optim.Adam (no weight decay, no amsgrad), one flat tensor.  step >= 1.
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, int64_t n,
                            float lr, float beta1, float beta2, float eps, float bc1, float bc2_sqrt, float grad_scale,
                            const float* __restrict__ grad_scale_dev) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (grad_scale_dev != nullptr) grad_scale *= __ldg(grad_scale_dev);
  const float gi = g[i] * grad_scale;
  const float mi = beta1 * m[i] + (1.f - beta1) * gi;
  const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;
  m[i] = mi;
  v[i] = vi;
  const float denom = sqrtf(vi) / bc2_sqrt + eps;
  p[i] -= (lr / bc1) * (mi / denom);
}

// the same update with the bias corrections taken from the device-resident step state (CUDA-graph replayable)
__global__ void adam_state_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, int64_t n,
                                  float lr, float beta1, float beta2, float eps, const dsdf_step_state* __restrict__ state, float grad_scale,
                                  const float* __restrict__ grad_scale_dev) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float bc1 = state->bc1, bc2_sqrt = state->bc2_sqrt;
  if (grad_scale_dev != nullptr) grad_scale *= __ldg(grad_scale_dev);
  const float gi = g[i] * grad_scale;
  const float mi = beta1 * m[i] + (1.f - beta1) * gi;
  const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;
  m[i] = mi;
  v[i] = vi;
  const float denom = sqrtf(vi) / bc2_sqrt + eps;
  p[i] -= (lr / bc1) * (mi / denom);
}

__global__ void step_advance_kernel(dsdf_step_state* state, double beta1, double beta2, int set, uint32_t value) {
  const uint32_t s = set ? value : state->step + 1u;
  state->step = s;
  state->bc1 = float(1.0 - pow(beta1, double(s)));
  state->bc2_sqrt = sqrtf(float(1.0 - pow(beta2, double(s))));
}

// ||g||^2 in two deterministic stages: 1024 blocks of fixed element ranges, then one block.
__global__ void __launch_bounds__(256) sumsq_partial_kernel(const float* __restrict__ g, int64_t n, float* __restrict__ part) {
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
__global__ void __launch_bounds__(1024) clip_coef_kernel(const float* __restrict__ part, float max_norm, float* __restrict__ out) {
  __shared__ float red[1024];
  red[threadIdx.x] = part[threadIdx.x];
  __syncthreads();
  for (int w = 512; w > 0; w >>= 1) {
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

}  // namespace dsdf
