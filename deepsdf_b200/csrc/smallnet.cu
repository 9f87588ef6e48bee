// Fused eval-mode forward + input gradient for the SMALL decoders the reference's optimisation / analysis consumers
// actually run (optimization_runs/*/config.json: experiments/double_lattice_3D_small_network 4x32,
// round_cross_big_network 6x128, ...; corner_spheres_only_small_network 4x64): hidden widths <= 128, weight-norm or plain
// layers, any latent_in set, CodeLength + geom_dimension <= 16, no LayerNorm / xyz_in_all.
//
// These networks are far too small for the 512-wide tcgen05 chain and - at 4 k .. 83 k MAC per row - neither HBM- nor
// tensor-bound in any useful sense: what the reference pays for is ~30 kernel launches and [N, width] HBM round trips per
// layer, times L x n_cpt re-evaluations in the jvp loop (deep_sdf/mesh.py:405-422, 0.8-5.8 s per call in its logs).  Here
// ONE kernel keeps a 128-row tile's activations in shared memory across all layers (feature-major [feature][row] fp32),
// streams each layer's folded fp32 weights through shared memory once per tile, runs the layer as a register-tiled FFMA
// GEMM (8 rows x WP/16 features per thread), records the ReLU gates as one byte per (feature, 8 rows), and - when the
// caller wants G = dSDF/d[latent, xyz] - walks the chain back in the same kernel with the same tile still resident.
// fp32 throughout and fixed summation order: SDF and G agree with the fp32 oracle to ~1e-6 (tests/test_smallnet_gpu.py).
// Replaces Decoder.forward (deep_sdf/networks/deep_sdf_decoder.py:76-111) + torch.autograd for these networks in
// decode_sdf (deep_sdf/utils.py:54-65), create_mesh (deep_sdf/mesh.py:37-73) and evaluate_network (mesh.py:457-528).
#include "common.cuh"
#include "host_util.h"
#include "../../include/dsdf.h"

namespace dsdf {

namespace {

constexpr int SN_ROWS = 128;                 // rows per tile
constexpr int SN_PITCH = SN_ROWS + 4;        // activation row pitch in floats (bank spread for the feature-strided stores)
constexpr int SN_THREADS = 256;              // 16 row groups (8 rows) x 16 feature lanes
constexpr int SN_MAX_IN = 16;                // CodeLength + geom_dimension
constexpr int SN_MAX_HIDDEN = 8;

struct SmallLayout {
  int wp;                                    // padded width: 32, 64, 96 or 128
  int nh;                                    // hidden layers
  size_t off_wt[DSDF_MAX_LAYERS];            // f32 [wp][wp]  W^T: row = input feature, col = output feature (forward operand)
  size_t off_w[DSDF_MAX_LAYERS];             // f32 [wp][wp]  W  : row = output feature, col = input feature (backward operand)
  size_t off_b[DSDF_MAX_LAYERS];             // f32 [wp]
  size_t off_wlast, off_blast;               // f32 [wp], [1]
  size_t total;                              // floats
};

int small_unsupported(const dsdf_net_desc* d, const char** why) {
  auto fail = [&](const char* m) { if (why) *why = m; return 1; };
  const int nl = d->num_linear, nh = nl - 1;
  if (nl < 2 || nh > SN_MAX_HIDDEN) return fail("needs 1..8 hidden layers");
  if (d->layernorm_mask) return fail("LayerNorm not supported by the small-network kernel");
  if (d->xyz_in_all) return fail("xyz_in_all not supported by the small-network kernel");
  if (d->latent_dropout) return fail("latent_dropout not supported by the small-network kernel");
  if (d->out_dim[nl - 1] != 1) return fail("last layer must have one output");
  const int LD = d->latent_size + d->geom_dim;
  if (LD < 1 || LD > SN_MAX_IN) return fail("CodeLength + geom_dimension must be <= 16");
  if (d->in_dim[0] != LD) return fail("first layer must take [latent, xyz]");
  if (d->latent_in_mask & 1u) return fail("latent_in=[0] not supported");
  for (int l = 0; l < nl; ++l) {
    if (d->in_dim[l] > 128 || (l < nh && d->out_dim[l] > 128)) return fail("hidden width must be <= 128");
    if (l > 0) {
      const bool lat = (d->latent_in_mask >> l) & 1;
      if (d->in_dim[l] != d->out_dim[l - 1] + (lat ? LD : 0)) return fail("layer widths do not chain");
    }
  }
  if (why) *why = "";
  return 0;
}

int small_layout(const dsdf_net_desc* d, SmallLayout* lo) {
  const char* why = nullptr;
  if (small_unsupported(d, &why)) return set_error("small-network kernel unsupported: %s", why);
  memset(lo, 0, sizeof(*lo));
  const int nl = d->num_linear;
  int mx = 1;
  for (int l = 0; l < nl; ++l) {
    if (d->in_dim[l] > mx) mx = d->in_dim[l];
    if (l < nl - 1 && d->out_dim[l] > mx) mx = d->out_dim[l];
  }
  lo->wp = (mx + 31) / 32 * 32;
  lo->nh = nl - 1;
  size_t off = 0;
  const size_t sq = size_t(lo->wp) * lo->wp;
  for (int l = 0; l < lo->nh; ++l) {
    lo->off_wt[l] = off; off += sq;
    lo->off_w[l] = off;  off += sq;
    lo->off_b[l] = off;  off += lo->wp;
  }
  lo->off_wlast = off; off += lo->wp;
  lo->off_blast = off; off += 4;
  lo->total = off;
  return 0;
}

struct SmallPackArgs {
  dsdf_net_desc desc;
  dsdf_params params;
  SmallLayout lo;
  float* pack;
};

// grid (max out rows, num_linear), block 128: fold g * v / ||v|| of one output row (torch._weight_norm, dim 0) and scatter
// it into the padded operand blocks
__global__ void small_pack_kernel(const __grid_constant__ SmallPackArgs a) {
  const int l = blockIdx.y, o = blockIdx.x;
  const dsdf_net_desc& d = a.desc;
  if (o >= d.out_dim[l]) return;
  const int in = d.in_dim[l], wp = a.lo.wp, nh = a.lo.nh;
  const float* vrow = a.params.v[l] + size_t(o) * in;
  __shared__ float red[4];
  __shared__ float s_scale;
  float ss = 0.f;
  for (int k = threadIdx.x; k < in; k += blockDim.x) { const float t = vrow[k]; ss += t * t; }
  for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x == 0) {
    const float tot = (red[0] + red[1]) + (red[2] + red[3]);
    s_scale = (a.params.g[l] != nullptr) ? a.params.g[l][o] / sqrtf(tot) : 1.0f;
    if (l < nh) a.pack[a.lo.off_b[l] + o] = a.params.b[l][o];
    else a.pack[a.lo.off_blast] = a.params.b[l][0];
  }
  __syncthreads();
  const float sc = s_scale;
  for (int k = threadIdx.x; k < in; k += blockDim.x) {
    const float w = vrow[k] * sc;
    if (l == nh) { a.pack[a.lo.off_wlast + k] = w; continue; }
    a.pack[a.lo.off_wt[l] + size_t(k) * wp + o] = w;
    a.pack[a.lo.off_w[l] + size_t(o) * wp + k] = w;
  }
}

struct SmallArgs {
  dsdf_net_desc desc;
  SmallLayout lo;
  const float* pack;
  const float* queries;        // row r at queries + r * ldq: [latent(L), xyz(D)], or xyz only when `latent` is given
  int64_t ldq;
  const float* latent;         // one code [L] broadcast over all rows, or null
  int64_t n_rows;
  float* sdf_out;              // [n_rows]
  float* G;                    // [n_rows][ldg] = dSDF / d[latent, xyz], or null
  int64_t ldg;
  int32_t n_tiles;
  // frozen-decoder latent optimisation (upstream reconstruct.py): one code per `rows_per_latent` consecutive rows taken from
  // lat_table, and the clamped-L1 seed dL/dsdf = sign(clamp(sdf) - clamp(gt)) * [|sdf| <= clamp] * inv_n instead of 1
  const float* lat_table;      // [n_rows / rows_per_latent][L], or null
  int64_t rows_per_latent;
  const float* gt;             // value of row r at gt[r * gt_stride], or null
  int64_t gt_stride;
  float clamp, inv_n;
  float* row_loss;             // [n_rows] |clamp(sdf) - clamp(gt)|, or null
};

// acc[8 rows][CPT features] += sum_k A[k][8 rows of this thread] * B[k][tx + 16 j]
template <int WP>
__device__ __forceinline__ void gemm_tile(const float* __restrict__ A, const float* __restrict__ B, int K, int ty, int tx,
                                          float (&acc)[8][WP / 16]) {
  constexpr int CPT = WP / 16;
  const float* ap = A + ty * 8;
  const float* bp = B + tx;
#pragma unroll 4
  for (int k = 0; k < K; ++k) {
    const float4 a0 = *reinterpret_cast<const float4*>(ap + k * SN_PITCH);
    const float4 a1 = *reinterpret_cast<const float4*>(ap + k * SN_PITCH + 4);
    const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    float bv[CPT];
#pragma unroll
    for (int j = 0; j < CPT; ++j) bv[j] = bp[k * WP + 16 * j];
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int j = 0; j < CPT; ++j) acc[r][j] = fmaf(av[r], bv[j], acc[r][j]);
  }
}

template <int WP>
__global__ void __launch_bounds__(SN_THREADS) small_forward_kernel(const __grid_constant__ SmallArgs a) {
  constexpr int CPT = WP / 16;
  extern __shared__ __align__(16) float sm[];
  float* act0 = sm;                                   // [WP][SN_PITCH]
  float* act1 = act0 + WP * SN_PITCH;
  float* sW = act1 + WP * SN_PITCH;                   // [WP][WP]
  float* sIn = sW + WP * WP;                          // [SN_MAX_IN][SN_ROWS]: this tile's decoder input, feature-major
  float* sG = sIn + SN_MAX_IN * SN_ROWS;              // [SN_MAX_IN][SN_ROWS]: input-gradient accumulator
  float* sY = sG + SN_MAX_IN * SN_ROWS;               // [SN_ROWS]: d sdf / d (pre-tanh output)
  uint8_t* gates = reinterpret_cast<uint8_t*>(sY + SN_ROWS);   // [nh][WP][16]: bit i of byte (f, ty) = relu gate of row 8 ty + i
  const dsdf_net_desc& d = a.desc;
  const int nh = a.lo.nh, L = d.latent_size, D = d.geom_dim, LD = L + D;
  const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
  const bool with_grad = a.G != nullptr;
  const float* pack = a.pack;

  for (int i = tid; i < 2 * WP * SN_PITCH; i += SN_THREADS) act0[i] = 0.f;      // padded features must stay finite (x 0 weights)
  __syncthreads();

  for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
    const int64_t row0 = int64_t(tile) * SN_ROWS;
    // ---- this tile's input, feature-major (deep_sdf/utils.py:57-63: [latent, xyz] per row)
    for (int i = tid; i < LD * SN_ROWS; i += SN_THREADS) {
      const int c = i / SN_ROWS, r = i % SN_ROWS;      // r fastest within a feature: conflict-free stores
      const int64_t row = row0 + r;
      float v = 0.f;
      if (row < a.n_rows) {
        if (a.lat_table != nullptr) v = (c < L) ? __ldg(a.lat_table + (row / a.rows_per_latent) * L + c) : __ldg(a.queries + row * a.ldq + (c - L));
        else if (a.latent != nullptr) v = (c < L) ? __ldg(a.latent + c) : __ldg(a.queries + row * a.ldq + (c - L));
        else v = __ldg(a.queries + row * a.ldq + c);
      }
      sIn[c * SN_ROWS + r] = v;
      act0[c * SN_PITCH + r] = v;
      if (with_grad) sG[c * SN_ROWS + r] = 0.f;
    }
    float* cur = act0;
    float* nxt = act1;
    // ================================================================ forward chain
    for (int l = 0; l < nh; ++l) {
      const int in = d.in_dim[l], out = d.out_dim[l];
      __syncthreads();                               // cur complete; sW free
      if ((d.latent_in_mask >> l) & 1) {             // x = cat([x, input], 1)   (deep_sdf_decoder.py:88-89)
        const int x0 = in - LD;
        for (int i = tid; i < LD * SN_ROWS; i += SN_THREADS) cur[(x0 + i / SN_ROWS) * SN_PITCH + (i % SN_ROWS)] = sIn[i];
      }
      {
        const float4* src = reinterpret_cast<const float4*>(pack + a.lo.off_wt[l]);
        float4* dst = reinterpret_cast<float4*>(sW);
        for (int i = tid; i < WP * WP / 4; i += SN_THREADS) dst[i] = __ldg(src + i);
      }
      __syncthreads();
      float acc[8][CPT];
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        const float b = __ldg(pack + a.lo.off_b[l] + tx + 16 * j);
#pragma unroll
        for (int r = 0; r < 8; ++r) acc[r][j] = b;
      }
      gemm_tile<WP>(cur, sW, (in + 3) & ~3, ty, tx, acc);
      // relu (deep_sdf_decoder.py:104; eval mode: no dropout), gates, write the next activation tile
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        const int f = tx + 16 * j;
        uint32_t gb = 0;
        float v[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          const bool on = acc[r][j] > 0.f && f < out;
          v[r] = on ? acc[r][j] : 0.f;
          gb |= uint32_t(on) << r;
        }
        gates[(l * WP + f) * 16 + ty] = uint8_t(gb);
        float* o = nxt + f * SN_PITCH + ty * 8;
        *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
      }
      float* t = cur; cur = nxt; nxt = t;
    }
    __syncthreads();
    // ================================================================ last layer (-> 1), tanh (twice with use_tanh)
    if (tid < SN_ROWS) {
      const int in = d.in_dim[nh];
      float y = __ldg(pack + a.lo.off_blast);
      for (int f = 0; f < in; ++f) y = fmaf(cur[f * SN_PITCH + tid], __ldg(pack + a.lo.off_wlast + f), y);
      float t1 = tanhf(y), dt = 1.f;
      if (d.use_tanh) { dt = 1.f - t1 * t1; t1 = tanhf(t1); }      // deep_sdf_decoder.py:94-95 then the final self.th (:108-109)
      dt *= 1.f - t1 * t1;
      if (row0 + tid < a.n_rows) a.sdf_out[row0 + tid] = t1;
      if (a.gt != nullptr) {
        // clamped-L1 seed of the latent optimisation (same expression as the fused chain's last epilogue)
        float diff = 0.f, inside = 1.f;
        if (row0 + tid < a.n_rows) {
          const float g = __ldg(a.gt + (row0 + tid) * a.gt_stride);
          float pc = t1, gc = g;
          if (a.clamp > 0.f) {
            pc = fminf(fmaxf(t1, -a.clamp), a.clamp);
            gc = fminf(fmaxf(g, -a.clamp), a.clamp);
            inside = (t1 >= -a.clamp && t1 <= a.clamp) ? 1.f : 0.f;
          }
          diff = pc - gc;
          if (a.row_loss != nullptr) a.row_loss[row0 + tid] = fabsf(diff);
        }
        const float sgn = (diff > 0.f) ? 1.f : ((diff < 0.f) ? -1.f : 0.f);
        dt *= sgn * inside * a.inv_n;
      }
      sY[tid] = dt;
    }
    __syncthreads();               // the last-layer reads of `cur` are done before anybody stages the next tile / seeds dY
    if (!with_grad) continue;
    // ================================================================ backward chain: dY_l = (dY_{l+1} W_{l+1}) (.) gate_l
    // seed: dX[f][r] = dy[r] * w_last[f], gated by the last hidden layer
    {
      const int lh = nh - 1;
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        const int f = tx + 16 * j;
        const float w = (f < d.in_dim[nh]) ? __ldg(pack + a.lo.off_wlast + f) : 0.f;
        const uint32_t gb = gates[(lh * WP + f) * 16 + ty];
        float v[8];
#pragma unroll
        for (int r = 0; r < 8; ++r) v[r] = ((gb >> r) & 1u) ? sY[ty * 8 + r] * w : 0.f;
        float* o = nxt + f * SN_PITCH + ty * 8;
        *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
      }
      float* t = cur; cur = nxt; nxt = t;            // cur = dY of the last hidden layer
    }
    for (int l = nh - 1; l >= 1; --l) {
      const int in = d.in_dim[l], out = d.out_dim[l];
      const bool lat = (d.latent_in_mask >> l) & 1;
      const int x0 = lat ? in - LD : in;               // width of the hidden part of this layer's input = out_dim[l-1]
      __syncthreads();
      {
        const float4* src = reinterpret_cast<const float4*>(pack + a.lo.off_w[l]);
        float4* dst = reinterpret_cast<float4*>(sW);
        for (int i = tid; i < WP * WP / 4; i += SN_THREADS) dst[i] = __ldg(src + i);
      }
      __syncthreads();
      float acc[8][CPT];
#pragma unroll
      for (int j = 0; j < CPT; ++j)
#pragma unroll
        for (int r = 0; r < 8; ++r) acc[r][j] = 0.f;
      gemm_tile<WP>(cur, sW, (out + 3) & ~3, ty, tx, acc);       // dX[i][r] = sum_o dY[o][r] W[o][i]
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        const int f = tx + 16 * j;
        float v[8];
        if (f < x0) {
          const uint32_t gb = gates[((l - 1) * WP + f) * 16 + ty];
#pragma unroll
          for (int r = 0; r < 8; ++r) v[r] = ((gb >> r) & 1u) ? acc[r][j] : 0.f;
        } else {
          // the re-injected [latent, xyz] columns of a latent_in layer feed G directly
          if (lat && f < in) {
#pragma unroll
            for (int r = 0; r < 8; ++r) sG[(f - x0) * SN_ROWS + ty * 8 + r] += acc[r][j];
          }
#pragma unroll
          for (int r = 0; r < 8; ++r) v[r] = 0.f;
        }
        float* o = nxt + f * SN_PITCH + ty * 8;
        *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(o + 4) = make_float4(v[4], v[5], v[6], v[7]);
      }
      float* t = cur; cur = nxt; nxt = t;
    }
    __syncthreads();
    // layer 0: G[r][c] = (latent_in contributions) + sum_o dY_0[o][r] W_0[o][c]
    for (int i = tid; i < LD * SN_ROWS; i += SN_THREADS) {
      const int c = i / SN_ROWS, r = i % SN_ROWS;
      float g = sG[i];
      const int out0 = d.out_dim[0];
      const float* w0 = pack + a.lo.off_wt[0] + size_t(c) * WP;      // W^T row c = column c of W_0
      for (int o = 0; o < out0; ++o) g = fmaf(cur[o * SN_PITCH + r], __ldg(w0 + o), g);
      sG[i] = g;
    }
    __syncthreads();
    for (int i = tid; i < SN_ROWS * LD; i += SN_THREADS) {
      const int r = i / LD, c = i % LD;                  // c fastest: consecutive addresses within a row of G
      if (row0 + r < a.n_rows) a.G[(row0 + r) * a.ldg + c] = sG[c * SN_ROWS + r];
    }
    __syncthreads();               // sG / sIn are re-staged by the next tile
  }
}

// Per shape s (rows [s R, s R + R)): lat_grad[s][c] = sum_r G[r][c] + 2 l2reg z[c] / L,  shape_loss[s] = inv_n sum_r row_loss[r]
// + l2reg mean(z^2).  One block per shape, fixed summation order (thread-strided partial sums, then a shared-memory tree).
__global__ void __launch_bounds__(256) small_latent_finish_kernel(const float* __restrict__ G, int64_t ldg, const float* __restrict__ row_loss,
                                                                  const float* __restrict__ lat_table, int L, int64_t R, float inv_n,
                                                                  float l2reg, float* __restrict__ lat_grad, float* __restrict__ shape_loss) {
  const int s = blockIdx.x;
  __shared__ float red[256];
  const float* z = lat_table + size_t(s) * L;
  for (int c = 0; c <= L; ++c) {
    float acc = 0.f;
    for (int64_t r = threadIdx.x; r < R; r += 256) acc += (c < L) ? G[(int64_t(s) * R + r) * ldg + c] : row_loss[int64_t(s) * R + r];
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
      if (threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w];
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      if (c < L) {
        lat_grad[size_t(s) * L + c] = red[0] + 2.f * l2reg * z[c] / float(L);
      } else {
        float q = 0.f;
        for (int k = 0; k < L; ++k) q += z[k] * z[k];
        shape_loss[s] = red[0] * inv_n + l2reg * q / float(L);
      }
    }
    __syncthreads();
  }
}

template <int WP>
size_t small_smem_bytes(int nh) {
  return size_t(2 * WP * SN_PITCH + WP * WP + 2 * SN_MAX_IN * SN_ROWS + SN_ROWS) * 4 + size_t(nh) * WP * 16;
}

template <int WP>
int launch_small(const SmallArgs& a, cudaStream_t st) {
  const size_t smem = small_smem_bytes<WP>(a.lo.nh);
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(small_forward_kernel<WP>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  int per_sm = 0;
  DSDF_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, small_forward_kernel<WP>, SN_THREADS, smem));
  if (per_sm < 1) per_sm = 1;
  int grid = sm_count() * per_sm;
  if (grid > a.n_tiles) grid = a.n_tiles;
  DSDF_LAUNCHED;
  small_forward_kernel<WP><<<grid, SN_THREADS, smem, st>>>(a);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace
}  // namespace dsdf

using namespace dsdf;

extern "C" int dsdf_small_supported(const dsdf_net_desc* desc) { return small_unsupported(desc, nullptr) == 0; }

extern "C" size_t dsdf_small_pack_bytes(const dsdf_net_desc* desc) {
  SmallLayout lo;
  if (small_layout(desc, &lo)) return 0;
  return lo.total * sizeof(float);
}

extern "C" int dsdf_small_pack_weights(const dsdf_net_desc* desc, const dsdf_params* params, void* pack, void* stream) {
  if (int e = require_sm100()) return e;
  SmallPackArgs a;
  a.desc = *desc;
  a.params = *params;
  if (int e = small_layout(desc, &a.lo)) return e;
  DSDF_REQUIRE((reinterpret_cast<uintptr_t>(pack) & 15) == 0, "small pack buffer must be 16-byte aligned");
  a.pack = static_cast<float*>(pack);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DSDF_CHECK_CUDA(cudaMemsetAsync(pack, 0, a.lo.total * sizeof(float), st));
  int max_out = 1;
  for (int l = 0; l < desc->num_linear; ++l) max_out = desc->out_dim[l] > max_out ? desc->out_dim[l] : max_out;
  DSDF_LAUNCHED;
  small_pack_kernel<<<dim3(max_out, desc->num_linear), 128, 0, st>>>(a);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

static int small_dispatch(const SmallArgs& a, cudaStream_t st) {
  switch (a.lo.wp) {
    case 32: return launch_small<32>(a, st);
    case 64: return launch_small<64>(a, st);
    case 96: return launch_small<96>(a, st);
    case 128: return launch_small<128>(a, st);
    default: return set_error("small-network kernel: unexpected padded width");
  }
}

extern "C" int dsdf_small_latent_step(const dsdf_net_desc* desc, const void* pack, const float* samples, int64_t lds,
                                      const float* lat_table, int32_t n_shapes, int64_t rows_per_shape, float clamp, float l2reg,
                                      float* sdf_scratch, float* G_scratch, float* row_loss_scratch, float* lat_grad,
                                      float* shape_loss, void* stream) {
  if (int e = require_sm100()) return e;
  SmallArgs a;
  memset(&a, 0, sizeof(a));
  a.desc = *desc;
  if (int e = small_layout(desc, &a.lo)) return e;
  DSDF_REQUIRE(desc->geom_dim == 3 && lds >= 4, "dsdf_small_latent_step: samples are [rows][>= 4] (x, y, z, sdf)");
  DSDF_REQUIRE(n_shapes > 0 && rows_per_shape > 0 && samples && lat_table && sdf_scratch && G_scratch && row_loss_scratch && lat_grad &&
               shape_loss, "dsdf_small_latent_step: bad arguments");
  const int L = desc->latent_size;
  a.pack = static_cast<const float*>(pack);
  a.queries = samples; a.ldq = lds; a.n_rows = int64_t(n_shapes) * rows_per_shape; a.sdf_out = sdf_scratch;
  a.G = G_scratch; a.ldg = L + 3;
  a.lat_table = lat_table; a.rows_per_latent = rows_per_shape;
  a.gt = samples + 3; a.gt_stride = lds; a.clamp = clamp; a.inv_n = 1.0f / float(rows_per_shape); a.row_loss = row_loss_scratch;
  a.n_tiles = int((a.n_rows + SN_ROWS - 1) / SN_ROWS);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (int e = small_dispatch(a, st)) return e;
  DSDF_LAUNCHED;
  small_latent_finish_kernel<<<n_shapes, 256, 0, st>>>(G_scratch, L + 3, row_loss_scratch, lat_table, L, rows_per_shape, a.inv_n, l2reg,
                                                       lat_grad, shape_loss);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_small_forward(const dsdf_net_desc* desc, const void* pack, const float* queries, int64_t ldq,
                                  const float* latent, int64_t n_rows, float* sdf_out, float* G, int64_t ldg, void* stream) {
  if (int e = require_sm100()) return e;
  SmallArgs a;
  memset(&a, 0, sizeof(a));
  a.desc = *desc;
  if (int e = small_layout(desc, &a.lo)) return e;
  DSDF_REQUIRE(n_rows > 0 && queries != nullptr && sdf_out != nullptr, "dsdf_small_forward: bad arguments");
  const int LD = desc->latent_size + desc->geom_dim;
  DSDF_REQUIRE(ldq >= (latent ? desc->geom_dim : LD), "dsdf_small_forward: ldq too small for the query rows");
  DSDF_REQUIRE(G == nullptr || ldg >= LD, "dsdf_small_forward: ldg must be >= CodeLength + geom_dimension");
  a.pack = static_cast<const float*>(pack);
  a.queries = queries; a.ldq = ldq; a.latent = latent; a.n_rows = n_rows; a.sdf_out = sdf_out; a.G = G; a.ldg = ldg;
  a.n_tiles = int((n_rows + SN_ROWS - 1) / SN_ROWS);
  return small_dispatch(a, static_cast<cudaStream_t>(stream));
}
