// Host side of the fused tcgen05 path: support check, packed-weight layout, weight-norm fold + bf16 packing,
// per-scene bias preparation (latent gather + Embedding max_norm renorm), and the C-ABI launch wrappers.
#include <stdlib.h>
#include "common.cuh"
#include "fused.cuh"
#include "fused_host.h"
#include "host_util.h"

namespace dsdf {

// ---------------------------------------------------------------------------------------------- layout
static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

int fused_unsupported_reason(const dsdf_net_desc* d, const char** why) {
  const int nl = d->num_linear, nh = nl - 1;
  auto fail = [&](const char* m) { if (why) *why = m; return 1; };
  if (nl < 3 || nl > DSDF_MAX_LAYERS) return fail("needs 2..15 hidden layers");
  if (d->geom_dim != 3) return fail("geom_dimension must be 3");
  if (d->xyz_in_all) return fail("xyz_in_all not supported by the fused path");
  if (d->layernorm_mask) return fail("LayerNorm not supported by the fused path");
  if (d->latent_dropout) return fail("latent_dropout not supported by the fused path");
  if (d->latent_size < 1 || d->latent_size > 256) return fail("latent_size must be 1..256");
  if (d->out_dim[nl - 1] != 1) return fail("last layer must have one output");
  if (d->latent_in_mask & 1u) return fail("latent_in=[0] not supported");
  if (d->latent_in_mask >> nh) return fail("latent_in on the last layer not supported");
  int n_lat = 0;
  for (int l = 1; l < nh; ++l) n_lat += (d->latent_in_mask >> l) & 1;
  if (n_lat > 1) return fail("at most one latent_in layer");
  const int LD = d->latent_size + d->geom_dim;
  if (d->in_dim[0] != LD) return fail("first layer must take [latent, xyz]");
  for (int l = 0; l < nh; ++l) {
    const bool next_lat = (d->latent_in_mask >> (l + 1)) & 1;
    if (d->out_dim[l] != (next_lat ? FH - LD : FH)) return fail("hidden width must be 512");
    if (l > 0 && d->in_dim[l] != FH) return fail("hidden width must be 512");
  }
  if (d->in_dim[nl - 1] != FH) return fail("hidden width must be 512");
  if (why) *why = "";
  return 0;
}

int fused_layout(const dsdf_net_desc* d, FusedLayout* lo) {
  const char* why = nullptr;
  if (fused_unsupported_reason(d, &why)) return set_error("fused path unsupported: %s", why);
  memset(lo, 0, sizeof(*lo));
  const int nl = d->num_linear, nh = nl - 1;
  const int L = d->latent_size, LD = L + d->geom_dim;
  lo->num_hidden = nh;
  lo->lat_layer = -1;
  int rows_f = 0, rows_b = 0;
  for (int l = 0; l < nh; ++l) {
    const bool lat = (d->latent_in_mask >> l) & 1;
    if (lat) lo->lat_layer = l;
    lo->xw[l] = (l == 0) ? 0 : (lat ? d->in_dim[l] - LD : d->in_dim[l]);
    lo->lat_col0[l] = (l == 0) ? 0 : (lat ? lo->xw[l] : -1);
    lo->xyz_col0[l] = (l == 0) ? L : (lat ? lo->xw[l] + L : -1);
    lo->fwd_row0[l] = rows_f;
    lo->fwd_halves[l] = (d->out_dim[l] + HALF_N - 1) / HALF_N;
    lo->fwd_kchunks[l] = (lo->xw[l] + 63) / 64;
    if (l > 0) rows_f += lo->fwd_halves[l] * HALF_N;
    lo->bwd_row0[l] = rows_b;
    lo->bwd_halves[l] = (lo->xw[l] + HALF_N - 1) / HALF_N;
    lo->bwd_kchunks[l] = (d->out_dim[l] + 63) / 64;
    if (l > 0) rows_b += lo->bwd_halves[l] * HALF_N;
  }
  // lin0's dense latent block sits behind the others (scene mode never reads it: lin0 has no K chunks there)
  lo->fwd_row0[0] = rows_f;
  rows_f += FH;
  lo->fwd_rows = rows_f;
  lo->bwd_rows = rows_b;
  lo->dense_lat_chunks = 0;
  lo->dense_kx = 0;
  if (lo->lat_layer > 1) {
    const int xw = lo->xw[lo->lat_layer], xc = (xw + 63) / 64, lc = (L + 63) / 64;
    // the layer before the latent_in layer must leave the upper half of the activation tile free (one accumulator half)
    if (L > AUX_MAX_ROW_LAT1 && xc <= 4 && lc <= 4 && lo->fwd_halves[lo->lat_layer - 1] == 1) {
      lo->dense_lat_chunks = lc;
      lo->dense_kx = 256;
    }
  }
  size_t off = 0;
  lo->off_wp = off;      off = align_up(off + size_t(rows_f) * FH * 2, 1024);
  lo->off_wpt = off;     off = align_up(off + size_t(rows_b) * FH * 2, 1024);
  const float sc = dropout_scale(d->dropout_prob);
  bool any_drop = false;
  lo->in_scale[0] = 1.0f;
  for (int l = 1; l <= nh; ++l) {
    const bool dr = ((d->dropout_mask >> (l - 1)) & 1) && d->dropout_prob > 0.f;
    lo->in_scale[l] = dr ? sc : 1.0f;
    any_drop |= dr;
  }
  lo->off_wp_tr = lo->off_wp;
  lo->off_wpt_tr = lo->off_wpt;
  if (any_drop) {
    lo->off_wp_tr = off;   off = align_up(off + size_t(rows_f) * FH * 2, 1024);
    lo->off_wpt_tr = off;  off = align_up(off + size_t(rows_b) * FH * 2, 1024);
  }
  lo->off_aux = off;     off = align_up(off + size_t(nh) * FH * AUX_W * 2, 1024);
  lo->off_bias = off;    off = align_up(off + size_t(nh) * FH * 4, 256);
  lo->off_wxyz = off;    off = align_up(off + size_t(2) * 3 * FH * 4, 256);        // slot 0: lin0, slot 1: skip layer
  lo->off_wlatT = off;   off = align_up(off + size_t(2) * L * FH * 4, 256);        // [slot][k][512]
  lo->off_wlast = off;   off = align_up(off + size_t(FH) * 4, 256);
  lo->off_wlast_tr = lo->off_wlast;
  if (any_drop) { lo->off_wlast_tr = off; off = align_up(off + size_t(FH) * 4, 256); }
  lo->jac_rows = (LD + 15) / 16 * 16;
  lo->off_wjac = align_up(off, 1024);   off = align_up(lo->off_wjac + size_t(lo->jac_rows) * 2 * FH * 2, 1024);
  for (int l = 0; l < nl; ++l) {
    lo->off_weff[l] = off;                 // (boundary marker: the fp32 folded copy was never read and is no longer written)
    lo->off_invn[l] = off;  off = align_up(off + size_t(d->out_dim[l]) * 4, 256);
  }
  lo->total_bytes = off;
  return 0;
}

// ---------------------------------------------------------------------------------------------- fold + pack
struct PackArgs {
  dsdf_net_desc desc;
  dsdf_params params;
  FusedLayout lo;
  uint8_t* pack;
  int32_t want_eval, want_train;      // which variants of the big bf16 blocks / w_last to write
};

// grid: (max out rows, num_linear); block 128.  One block folds one output row of one layer.
__global__ void pack_weights_kernel(const __grid_constant__ PackArgs a) {
  const int l = blockIdx.y, o = blockIdx.x;
  const dsdf_net_desc& d = a.desc;
  if (o >= d.out_dim[l]) return;
  const int in = d.in_dim[l];
  const int nh = a.lo.num_hidden;
  const float* vrow = a.params.v[l] + size_t(o) * in;
  __shared__ float red[4];
  __shared__ float s_scale;
  float ss = 0.f;
  for (int k = threadIdx.x; k < in; k += blockDim.x) { float t = vrow[k]; ss += t * t; }
  for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tot = (red[0] + red[1]) + (red[2] + red[3]);
    float invn = 1.0f, scale = 1.0f;
    if (a.params.g[l] != nullptr) { invn = 1.0f / sqrtf(tot); scale = a.params.g[l][o] * invn; }
    s_scale = scale;
    reinterpret_cast<float*>(a.pack + a.lo.off_invn[l])[o] = invn;
    if (l < nh) {
      const float b = a.params.b[l][o];
      reinterpret_cast<float*>(a.pack + a.lo.off_bias)[l * FH + o] = b;
      // aux B row of this output feature: the bias, once for each CTA of a pair (fused.cuh)
      __nv_bfloat16* ar = reinterpret_cast<__nv_bfloat16*>(a.pack + a.lo.off_aux) + (size_t(l) * FH + o) * AUX_W;
      __nv_bfloat16 hi, lo2;
      split_bf16(b, hi, lo2);
      ar[AUX_COL_ONE0] = hi; ar[AUX_COL_ONE0 + 1] = lo2;
      ar[AUX_COL_ONE1] = hi; ar[AUX_COL_ONE1 + 1] = lo2;
    }
  }
  __syncthreads();
  const float scale = s_scale;
  __nv_bfloat16* wp = reinterpret_cast<__nv_bfloat16*>(a.pack + a.lo.off_wp);
  // training variants (dropout scale of the previous hidden layer folded into the activation columns).  The W^T blocks of
  // both variants are written by pack_transpose_kernel from the W blocks (a row-wise kernel would scatter 2-byte stores).
  const bool two = a.lo.off_wp_tr != a.lo.off_wp;
  const bool ev = a.want_eval != 0 || !two, tr = a.want_train != 0 && two;
  __nv_bfloat16* wp_tr = reinterpret_cast<__nv_bfloat16*>(a.pack + a.lo.off_wp_tr);
  float* wlast_tr = reinterpret_cast<float*>(a.pack + a.lo.off_wlast_tr);
  const float ins = a.lo.in_scale[l];
  __nv_bfloat16* aux_row = reinterpret_cast<__nv_bfloat16*>(a.pack + a.lo.off_aux) + (size_t(l < nh ? l : 0) * FH + o) * AUX_W;
  float* wxyz = reinterpret_cast<float*>(a.pack + a.lo.off_wxyz);
  float* wlatT = reinterpret_cast<float*>(a.pack + a.lo.off_wlatT);
  float* wlast = reinterpret_cast<float*>(a.pack + a.lo.off_wlast);
  __nv_bfloat16* wjac = reinterpret_cast<__nv_bfloat16*>(a.pack + a.lo.off_wjac);
  const int L = d.latent_size;
  for (int k = threadIdx.x; k < in; k += blockDim.x) {
    const float w = vrow[k] * scale;
    if (l == nh) { if (ev) wlast[k] = w; if (tr) wlast_tr[k] = w * ins; continue; }
    const int xw = a.lo.xw[l];
    if (k < xw) {
      if (l > 0) {
        if (ev) wp[size_t(a.lo.fwd_row0[l] + o) * FH + k] = __float2bfloat16_rn(w);
        if (tr) wp_tr[size_t(a.lo.fwd_row0[l] + o) * FH + k] = __float2bfloat16_rn(w * ins);
      }
    } else if (a.lo.lat_col0[l] >= 0 && k < a.lo.lat_col0[l] + L) {
      const int slot = (l == 0) ? 0 : 1;
      const int dlat = k - a.lo.lat_col0[l];
      wlatT[(size_t(slot) * L + dlat) * FH + o] = w;
      wjac[size_t(dlat) * (2 * FH) + slot * FH + o] = __float2bfloat16_rn(w);
      if (a.lo.dense_lat_chunks > 0 && ev) {
        // dense per-row codes: the latent columns as tensor-core K columns (eval variant: the mode is eval / gates only)
        const int kcol = (l == 0 ? 0 : a.lo.dense_kx) + dlat;
        wp[size_t(a.lo.fwd_row0[l] + o) * FH + kcol] = __float2bfloat16_rn(w);
      }
      if (L <= AUX_MAX_ROW_LAT) {           // row-latent mode: (w_hi, w_hi, w_lo) meets the A side's (z_hi, z_lo, z_hi)
        __nv_bfloat16 hi, lo2;
        split_bf16(w, hi, lo2);
        __nv_bfloat16* c = aux_row + AUX_COL_LAT + 3 * dlat;
        c[0] = hi; c[1] = hi; c[2] = lo2;
      } else if (L <= AUX_MAX_ROW_LAT1) {   // 4 < L <= 16: one bf16 term per latent column
        aux_row[AUX_COL_LAT + dlat] = __float2bfloat16_rn(w);
      }
    } else if (a.lo.xyz_col0[l] >= 0 && k >= a.lo.xyz_col0[l]) {
      const int slot = (l == 0) ? 0 : 1;
      const int dx = k - a.lo.xyz_col0[l];
      wxyz[(size_t(slot) * 3 + dx) * FH + o] = w;
      wjac[size_t(L + dx) * (2 * FH) + slot * FH + o] = __float2bfloat16_rn(w);
      __nv_bfloat16 hi, lo2;
      split_bf16(w, hi, lo2);
      __nv_bfloat16* c = aux_row + AUX_COL_XYZ + 3 * dx;
      c[0] = hi; c[1] = hi; c[2] = lo2;
    }
  }
}

// W^T blocks (the dgrad chain's B operand: rows = input features, K = output features) from the W blocks just written:
// 32 x 32 tiles through shared memory, so that both the reads and the writes are 64-byte row segments.
// grid (16, 16, hidden layers 1..nh-1), block (32, 8); src / dst select the variant.
__global__ void pack_transpose_kernel(const __nv_bfloat16* __restrict__ src, __nv_bfloat16* __restrict__ dst, FusedLayout lo,
                                      dsdf_net_desc d) {
  const int l = blockIdx.z + 1;
  const int out = d.out_dim[l], xw = lo.xw[l];
  const int o0 = blockIdx.y * 32, k0 = blockIdx.x * 32;
  if (o0 >= out || k0 >= xw) return;
  __shared__ __nv_bfloat16 t[32][34];
  for (int r = threadIdx.y; r < 32; r += 8) {
    const int o = o0 + r, k = k0 + threadIdx.x;
    t[r][threadIdx.x] = (o < out && k < xw) ? src[size_t(lo.fwd_row0[l] + o) * FH + k] : __float2bfloat16_rn(0.f);
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += 8) {
    const int k = k0 + r, o = o0 + threadIdx.x;
    if (k < xw && o < out) dst[size_t(lo.bwd_row0[l] + k) * FH + o] = t[threadIdx.x][r];
  }
}

// ---------------------------------------------------------------------------------------------- scene prepare
// Embedding(max_norm) lookup side effect: rows of the batch with ||z|| > max_norm are rescaled in place
// (train_deep_sdf.py:385,509).  One warp per scene.
__global__ void latent_renorm_kernel(float* table, const int64_t* scene_ids, int S, int L, float max_norm) {
  const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (s >= S) return;
  const int lane = threadIdx.x & 31;
  // a scene listed twice must be renormalised once: only its first occurrence acts
  const int64_t id = scene_ids[s];
  for (int t = 0; t < s; ++t) if (scene_ids[t] == id) return;
  float* row = table + size_t(id) * L;
  float ss = 0.f;
  for (int k = lane; k < L; k += 32) ss += row[k] * row[k];
  for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
  const float nrm = sqrtf(ss);
  if (nrm > max_norm) {
    const float sc = max_norm / (nrm + 1e-7f);
    for (int k = lane; k < L; k += 32) row[k] *= sc;
  }
}

// Per-scene aux B rows (fused.cuh): scene_aux[s][slot][o][16] bf16 with
//   b = bias_l[o] + sum_k Weff_l[o, lat_col0 + k] * z[scene_ids[s]][k]   (fp32, then split hi + lo)
// at columns 0,1 and 11,12 and the xyz columns of the layer at 2..10.
// grid (S, 2, 512 / SB_OUT), block (SB_OUT, SB_KS): the K = L sum of an output is cut into SB_KS slices that run side by
// side (a single thread walking all L latent columns is one long chain of dependent L2 loads: 19-29 us for 8 scenes at
// L = 256), folded in a fixed order through shared memory.
constexpr int SB_OUT = 128, SB_KS = 8;
__global__ void __launch_bounds__(SB_OUT * SB_KS)
scene_bias_kernel(const float* __restrict__ table, const int64_t* __restrict__ scene_ids, int L,
                  const float* __restrict__ wlatT, const float* __restrict__ bias, const float* __restrict__ wxyz,
                  int lat_layer, __nv_bfloat16* __restrict__ scene_aux) {
  const int s = blockIdx.x, slot = blockIdx.y, o = blockIdx.z * SB_OUT + threadIdx.x, ks = threadIdx.y;
  extern __shared__ float z[];                     // [L] the scene's code, then [SB_KS][SB_OUT] partial sums
  float* part = z + ((L + 3) & ~3);
  const float* row = table + size_t(scene_ids[s]) * L;
  const int tid = threadIdx.y * SB_OUT + threadIdx.x;
  for (int k = tid; k < L; k += SB_OUT * SB_KS) z[k] = row[k];
  __syncthreads();
  const bool live = (slot == 0 || lat_layer > 0);
  float acc = 0.f;
  if (live) {
    const float* wl = wlatT + size_t(slot) * L * FH + o;
    const int per = (L + SB_KS - 1) / SB_KS;
    const int k0 = ks * per, k1 = (k0 + per < L) ? k0 + per : L;
#pragma unroll 8
    for (int k = k0; k < k1; ++k) acc += wl[size_t(k) * FH] * z[k];
  }
  part[ks * SB_OUT + threadIdx.x] = acc;
  __syncthreads();
  if (ks != 0) return;
  uint32_t w[8] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
  if (live) {
    const int l = (slot == 0) ? 0 : lat_layer;
    float tot = bias[l * FH + o];
#pragma unroll
    for (int i = 0; i < SB_KS; ++i) tot += part[i * SB_OUT + threadIdx.x];
    __nv_bfloat16 e[AUX_SCENE_W];
#pragma unroll
    for (int i = 0; i < AUX_SCENE_W; ++i) e[i] = __float2bfloat16_rn(0.f);
    __nv_bfloat16 hi, lo2;
    split_bf16(tot, hi, lo2);
    e[AUX_COL_ONE0] = hi; e[AUX_COL_ONE0 + 1] = lo2; e[AUX_COL_ONE1] = hi; e[AUX_COL_ONE1 + 1] = lo2;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      split_bf16(wxyz[(size_t(slot) * 3 + d) * FH + o], hi, lo2);
      e[AUX_COL_XYZ + 3 * d] = hi; e[AUX_COL_XYZ + 3 * d + 1] = hi; e[AUX_COL_XYZ + 3 * d + 2] = lo2;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = bf16_pair_bits(e[2 * i], e[2 * i + 1]);
  }
  uint4* dst = reinterpret_cast<uint4*>(scene_aux + ((size_t(s) * 2 + slot) * FH + o) * AUX_SCENE_W);
  dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
  dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
}

static void fill_programs(const dsdf_net_desc* d, const FusedLayout& lo, ChainLayer* fwd, ChainLayer* bwd) {
  for (int l = 0; l < lo.num_hidden; ++l) {
    ChainLayer& F = fwd[l];
    F.has_mma = 1;                 // every forward layer has at least its aux step (lin0 in scene mode: nothing else)
    F.w_row0 = lo.fwd_row0[l];
    F.n_halves = lo.fwd_halves[l];
    F.k_chunks = lo.fwd_kchunks[l];
    F.bias_slot = (l == 0) ? 0 : (l == lo.lat_layer ? 1 : -1);
    F.has_xyz = (l == 0 || l == lo.lat_layer);
    F.dropout = (d->dropout_mask >> l) & 1;
    F.out_dim = d->out_dim[l];
    ChainLayer& B = bwd[l];
    B.has_mma = (l > 0);
    B.w_row0 = lo.bwd_row0[l];
    B.n_halves = lo.bwd_halves[l];
    B.k_chunks = lo.bwd_kchunks[l];
    B.bias_slot = -1;
    B.has_xyz = (l == 0 || l == lo.lat_layer);
    B.dropout = F.dropout;
    B.out_dim = d->out_dim[l];
  }
}

int launch_chain_forward(const CUtensorMap& tmW, const CUtensorMap& tmActs, const CUtensorMap& tmAux, const CUtensorMap& tmScene,
                         const FwdArgs& args, int n_sms, cudaStream_t stream);

}  // namespace dsdf

using namespace dsdf;

extern "C" int dsdf_fused_supported(const dsdf_net_desc* desc) { return fused_unsupported_reason(desc, nullptr) == 0; }

extern "C" int dsdf_fused_rows_supported(const dsdf_net_desc* desc) {
  FusedLayout lo;
  if (fused_unsupported_reason(desc, nullptr) != 0 || fused_layout(desc, &lo) != 0) return 0;
  return desc->latent_size <= AUX_MAX_ROW_LAT1 || lo.dense_lat_chunks > 0;
}

extern "C" size_t dsdf_pack_bytes(const dsdf_net_desc* desc) {
  FusedLayout lo;
  if (fused_layout(desc, &lo)) return 0;
  return lo.total_bytes;
}

extern "C" int dsdf_pack_weights_ex(const dsdf_net_desc* desc, const dsdf_params* params, void* pack, uint32_t flags, void* stream) {
  if (int e = require_sm100()) return e;
  PackArgs a;
  a.desc = *desc;
  a.params = *params;
  if (int e = fused_layout(desc, &a.lo)) return e;
  a.pack = static_cast<uint8_t*>(pack);
  a.want_eval = (flags & DSDF_PACK_EVAL) != 0;
  a.want_train = (flags & DSDF_PACK_TRAIN) != 0;
  DSDF_REQUIRE(a.want_eval || a.want_train, "dsdf_pack_weights_ex: flags must name DSDF_PACK_EVAL and / or DSDF_PACK_TRAIN");
  DSDF_REQUIRE((reinterpret_cast<uintptr_t>(pack) & 1023) == 0, "pack buffer must be 1024-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // the padded regions (rows / columns beyond the layer shapes) are never written: they only need zeroing once per buffer
  if (flags & DSDF_PACK_ZERO) DSDF_CHECK_CUDA(cudaMemsetAsync(pack, 0, a.lo.off_weff[0], st));
  int max_out = 0;
  for (int l = 0; l < desc->num_linear; ++l) max_out = desc->out_dim[l] > max_out ? desc->out_dim[l] : max_out;
  DSDF_LAUNCHED;
  pack_weights_kernel<<<dim3(max_out, desc->num_linear), 128, 0, st>>>(a);
  const bool two = a.lo.off_wp_tr != a.lo.off_wp;
  const uint8_t* p = a.pack;
  if (a.lo.num_hidden > 1) {
    const dim3 grid(FH / 32, FH / 32, a.lo.num_hidden - 1), block(32, 8);
    if (a.want_eval || !two) {
      DSDF_LAUNCHED;
      pack_transpose_kernel<<<grid, block, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(p + a.lo.off_wp),
                                                    reinterpret_cast<__nv_bfloat16*>(a.pack + a.lo.off_wpt), a.lo, a.desc);
    }
    if (a.want_train && two) {
      DSDF_LAUNCHED;
      pack_transpose_kernel<<<grid, block, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(p + a.lo.off_wp_tr),
                                                    reinterpret_cast<__nv_bfloat16*>(a.pack + a.lo.off_wpt_tr), a.lo, a.desc);
    }
  }
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_pack_weights(const dsdf_net_desc* desc, const dsdf_params* params, void* pack, void* stream) {
  return dsdf_pack_weights_ex(desc, params, pack, DSDF_PACK_EVAL | DSDF_PACK_TRAIN | DSDF_PACK_ZERO, stream);
}

extern "C" size_t dsdf_scene_aux_bytes(int32_t n_scenes) { return size_t(n_scenes > 0 ? n_scenes : 0) * 2 * FH * AUX_SCENE_W * 2; }

extern "C" int dsdf_scene_prepare(const dsdf_net_desc* desc, const void* pack, float* lat_table, const int64_t* scene_ids,
                                  int32_t n_scenes, float max_norm, void* scene_aux, void* stream) {
  if (int e = require_sm100()) return e;
  FusedLayout lo;
  if (int e = fused_layout(desc, &lo)) return e;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int L = desc->latent_size;
  if (max_norm > 0.f) {
    DSDF_LAUNCHED;
    latent_renorm_kernel<<<(n_scenes + 3) / 4, 128, 0, st>>>(lat_table, scene_ids, n_scenes, L, max_norm);
    DSDF_CHECK_CUDA(cudaGetLastError());
  }
  const uint8_t* p = static_cast<const uint8_t*>(pack);
  if (scene_aux != nullptr) {
    DSDF_REQUIRE((reinterpret_cast<uintptr_t>(scene_aux) & 15) == 0, "scene_aux must be 16-byte aligned");
    DSDF_LAUNCHED;
    scene_bias_kernel<<<dim3(n_scenes, 2, FH / SB_OUT), dim3(SB_OUT, SB_KS), (((L + 3) & ~3) + SB_KS * SB_OUT) * sizeof(float), st>>>(
        lat_table, scene_ids, L, reinterpret_cast<const float*>(p + lo.off_wlatT), reinterpret_cast<const float*>(p + lo.off_bias),
        reinterpret_cast<const float*>(p + lo.off_wxyz), lo.lat_layer, static_cast<__nv_bfloat16*>(scene_aux));
  }
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

static int fused_forward_impl(const dsdf_net_desc* desc, const void* pack, const void* cbias, const int32_t* tile_scene,
                              const float* xyz, int32_t xyz_stride, int64_t n_rows, float* sdf_out, const dsdf_train_io* tr,
                              const float* row_lat, int32_t row_lat_stride, void* stream) {
  if (int e = require_sm100()) return e;
  FusedLayout lo;
  if (int e = fused_layout(desc, &lo)) return e;
  DSDF_REQUIRE(n_rows > 0, "n_rows must be positive");
  const uint8_t* p = static_cast<const uint8_t*>(pack);
  FwdArgs a;
  memset(&a, 0, sizeof(a));
  ChainLayer bwd_unused[DSDF_MAX_LAYERS];
  fill_programs(desc, lo, a.prog, bwd_unused);
  a.num_hidden = lo.num_hidden;
  a.n_rows = n_rows;
  a.n_tiles = int((n_rows + TILE_M - 1) / TILE_M);
  a.n_rows_pad = int64_t(a.n_tiles) * TILE_M;
  // training with dropout: the packs whose activation columns carry the previous layer's dropout scale
  const bool tr_variant = tr->training && !tr->dropout_off;
  for (int l = 0; l < lo.num_hidden; ++l)
    a.aux_steps[l] = (row_lat != nullptr && (l == 0 || l == lo.lat_layer)) ? 2 : 1;
  a.w_last = reinterpret_cast<const float*>(p + (tr_variant ? lo.off_wlast_tr : lo.off_wlast));
  a.b_last = tr && tr->b_last ? tr->b_last : nullptr;
  DSDF_REQUIRE(a.b_last != nullptr, "b_last (bias of the last layer) must be given");
  a.tile_scene = tile_scene;
  a.xyz = xyz;
  DSDF_REQUIRE(xyz_stride >= 3, "xyz_stride must be >= 3");
  a.xyz_stride = xyz_stride;
  a.gt_stride = tr->gt_stride > 0 ? tr->gt_stride : 1;
  a.sdf_out = sdf_out;
  a.use_tanh = desc->use_tanh;
  a.training = tr->training;
  a.drop_thr = dropout_threshold(desc->dropout_prob);   // survivors' scale 1 / (1 - effective p): folded into the tr packs
  a.seed = tr->seed;
  philox_round_keys(tr->seed, a.philox_rk);
  a.row0 = tr->row0;
  a.step_dev = tr->step_state ? &tr->step_state->step : nullptr;
  a.gt = tr->gt;
  a.clamp = tr->clamp;
  a.inv_n = tr->inv_n;
  a.dy_out = tr->dy_out;
  a.tile_loss = tr->tile_loss;
  a.bitmask = tr->bitmask;
  a.store_acts = (tr->acts != nullptr);
  a.dropout_on = tr->dropout_off ? 0 : 1;
  a.dy_unit = tr->dy_unit;
  a.tile_rows = tr->tile_rows;
  if (row_lat != nullptr) {
    DSDF_REQUIRE(desc->latent_size <= AUX_MAX_ROW_LAT1 || lo.dense_lat_chunks > 0,
                 "row-latent mode supports latent_size <= 16, or long codes whose latent_in layer has <= 256 hidden inputs (e.g. L = 256)");
    DSDF_REQUIRE(row_lat_stride >= desc->latent_size, "row_lat_stride must be >= latent_size");
    a.row_lat = row_lat;
    a.row_lat_stride = row_lat_stride;
    a.row_lat_dim = desc->latent_size;
    if (desc->latent_size > AUX_MAX_ROW_LAT1) {
      // dense codes: the latent columns are K chunks of lin0 / the latent_in layer (eval-variant pack)
      DSDF_REQUIRE(!tr_variant, "dense per-row codes run in eval / gates mode only (no dropout)");
      a.dense_chunks = lo.dense_lat_chunks;
      a.dense_layer = lo.lat_layer;
      a.prog[0].k_chunks = lo.dense_lat_chunks;
      a.prog[lo.lat_layer].k_chunks = lo.dense_kx / 64 + lo.dense_lat_chunks;
      for (int l = 0; l < lo.num_hidden; ++l) a.aux_steps[l] = 1;
    }
  } else {
    DSDF_REQUIRE(cbias != nullptr && tile_scene != nullptr, "scene mode needs cbias and tile_scene");
  }
  if (a.training) {
    DSDF_REQUIRE(tr->n_rows_pad == a.n_rows_pad, "n_rows_pad must be n_rows rounded up to 128");
    if (a.gt) DSDF_REQUIRE(a.dy_out && a.tile_loss, "training with gt needs dy_out and tile_loss");
    if (a.dy_unit) DSDF_REQUIRE(a.dy_out && !a.gt, "dy_unit needs dy_out and no gt");
  }
  CUtensorMap tmW, tmActs, tmAux, tmScene;
  if (int e = make_tmap_bf16_2d(&tmW, p + (tr_variant ? lo.off_wp_tr : lo.off_wp), FH, lo.fwd_rows, FH * 2, 64, 128)) return e;
  if (int e = make_tmap_bf16_2d_noswz(&tmAux, p + lo.off_aux, AUX_W, uint64_t(lo.num_hidden) * FH, AUX_W * 2, 8, 128)) return e;
  if (cbias != nullptr) {
    // per-scene aux rows [n_scenes][2][512][16]; the scene count is the caller's business (tile_scene indexes it)
    DSDF_REQUIRE((reinterpret_cast<uintptr_t>(cbias) & 15) == 0, "scene_aux must be 16-byte aligned");
    if (int e = make_tmap_bf16_2d_noswz(&tmScene, cbias, AUX_SCENE_W, uint64_t(1) << 30, AUX_SCENE_W * 2, 8, 128)) return e;
  } else {
    tmScene = tmAux;
  }
  const void* acts_ptr = tr->acts ? tr->acts : static_cast<const void*>(p + lo.off_wp);   // dummy map when unused
  const uint64_t acts_rows = tr->acts ? uint64_t(lo.num_hidden) * a.n_rows_pad : uint64_t(lo.fwd_rows);
  if (int e = make_tmap_bf16_2d(&tmActs, acts_ptr, FH, acts_rows, FH * 2, 64, TILE_M)) return e;
  return launch_chain_forward(tmW, tmActs, tmAux, tmScene, a, sm_count(), static_cast<cudaStream_t>(stream));
}

extern "C" int dsdf_fused_forward(const dsdf_net_desc* desc, const void* pack, const void* cbias,
                                  const int32_t* tile_scene, const float* xyz, int32_t xyz_stride, int64_t n_rows,
                                  float* sdf_out, const dsdf_train_io* tr, void* stream) {
  return fused_forward_impl(desc, pack, cbias, tile_scene, xyz, xyz_stride, n_rows, sdf_out, tr, nullptr, 0, stream);
}

extern "C" int dsdf_fused_forward_rows(const dsdf_net_desc* desc, const void* pack, const float* queries, int32_t ldq,
                                       int64_t n_rows, float* sdf_out, const dsdf_train_io* tr, void* stream) {
  DSDF_REQUIRE(desc != nullptr && queries != nullptr && ldq >= desc->latent_size + 3, "queries must be [n_rows][ldq >= L + 3]");
  return fused_forward_impl(desc, pack, nullptr, nullptr, queries + desc->latent_size, ldq, n_rows, sdf_out, tr, queries, ldq, stream);
}
