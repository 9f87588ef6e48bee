// C-ABI wrappers of the fused training step: buffer sizing, dgrad chain, wgrad, finish (reductions + weight-norm
// backward + latent gradient + loss) and Adam.
#include <stdlib.h>
#include "train_kernels.cuh"
#include "host_util.h"

namespace dsdf {

int launch_chain_dgrad(const CUtensorMap& tmW, const CUtensorMap& tmDys, const BwdArgs& args, int n_sms, cudaStream_t stream);
int launch_wgrad(const CUtensorMap& tmDys, const CUtensorMap& tmActs, const WgradArgs& args, cudaStream_t stream);
int wgrad_max_active_clusters();
int launch_wgrad_last(const void* x_last, const float* dy, int64_t n_rows_pad, float* part, int n_blocks, cudaStream_t stream);

namespace {

constexpr int kLastBlocks = 1184;   // fixed (not device dependent): 8 x 148

// Cut the (layer, 32-row step) space into WG_CLUSTERS contiguous pieces of equal cost.  A cluster gets one piece = at most
// a few segments.
int build_plan(const dsdf_net_desc* d, const FusedLayout& lo, int64_t n_rows, WgradPlan* p) {
  memset(p, 0, sizeof(*p));
  (void)d;
  const int64_t n_steps = n_rows / 32;
  int64_t total = 0;
  // Cost model: wgrad runs at the HBM roofline and a step's time is set by its load latency, not by its bytes or MMAs, so
  // a 32-row step of a 256-wide layer costs the SAME as one of a 512-wide layer.  (Weighting it by its tensor work, 1:2,
  // gave the clusters that own the narrow layer twice the steps: measured 4.0 ms instead of 2.4 ms for the kernel.)
  auto cost = [&](int) { return 1; };
  for (int l = 1; l < lo.num_hidden; ++l) total += n_steps * cost(l);
  const int ncl = WG_CLUSTERS;
  p->n_clusters = ncl;
  const int64_t share = (total + ncl - 1) / ncl;
  int l = 1, nseg = 0;
  int64_t step = 0, off = 0;
  for (int l2 = 0; l2 <= DSDF_MAX_LAYERS; ++l2) p->layer_seg0[l2] = -1;
  for (int c = 0; c < ncl; ++c) {
    p->cluster_seg0[c] = nseg;
    int64_t remaining = share;
    while (remaining > 0 && l < lo.num_hidden) {
      const int u = lo.bwd_halves[l], cu = cost(l);
      int64_t take = (remaining + cu - 1) / cu;
      if (take > n_steps - step) take = n_steps - step;
      if (take > 0) {
        if (nseg >= WG_MAX_SEGS) return set_error("too many wgrad segments");
        WgradSeg& s = p->seg[nseg];
        s.layer = l; s.n = u * HALF_N; s.step_beg = int32_t(step); s.step_end = int32_t(step + take);
        s.part_offset = off;
        off += int64_t(4) * 128 * s.n;
        if (p->layer_seg0[l] < 0) p->layer_seg0[l] = nseg;
        ++nseg;
        step += take;
        remaining -= take * cu;
      }
      if (step >= n_steps) { ++l; step = 0; }
    }
  }
  p->cluster_seg0[ncl] = nseg;
  p->n_segs = nseg;
  // close the per-layer ranges: layer_seg0[l+1] = first segment of the next layer that has one (or n_segs)
  int next = nseg;
  for (int l2 = DSDF_MAX_LAYERS; l2 >= 0; --l2) {
    if (p->layer_seg0[l2] < 0) p->layer_seg0[l2] = next; else next = p->layer_seg0[l2];
  }
  p->total_floats = off;
  return 0;
}

inline int n_slots(const FusedLayout& lo) { return lo.num_hidden + 6; }

}  // namespace
}  // namespace dsdf

using namespace dsdf;

// sizes[]: 0 acts, 1 dys, 2 bitmask, 3 dy, 4 tile_loss, 5 tile_part, 6 wg_part, 7 last_part, 8 scene_sums (+scene_reg), 9 cbias
extern "C" int dsdf_train_buf_sizes(const dsdf_net_desc* desc, int64_t n_rows, int32_t n_scenes, size_t* sizes) {
  FusedLayout lo;
  if (int e = fused_layout(desc, &lo)) return e;
  const int64_t n_tiles = (n_rows + TILE_M - 1) / TILE_M, n_pad = n_tiles * TILE_M;
  WgradPlan plan;
  if (int e = build_plan(desc, lo, n_pad, &plan)) return e;
  const size_t nh = lo.num_hidden;
  sizes[0] = nh * n_pad * FH * 2;
  sizes[1] = nh * n_pad * FH * 2;
  sizes[2] = nh * n_pad * MASK_WORDS * 4;
  sizes[3] = size_t(n_pad) * 4;
  sizes[4] = size_t(n_tiles) * 4;
  sizes[5] = size_t(n_tiles) * n_slots(lo) * FH * 4;
  // split-K partials: one [4 x 128][n <= 512] block per segment, at most one segment per cluster plus one per layer boundary.
  // The bound (not the plan's own total) is reported so that the size is monotonic in n_rows: a workspace allocated for a
  // larger batch also serves every smaller one (the chunks of --batch_split).
  const size_t wg_bound = size_t(WG_CLUSTERS + lo.num_hidden) * 4 * 128 * FH;
  sizes[6] = (size_t(plan.total_floats) > wg_bound ? size_t(plan.total_floats) : wg_bound) * 4;
  sizes[7] = size_t(kLastBlocks) * (FH + 1) * 4;
  sizes[8] = (size_t(n_scenes) * n_slots(lo) * FH + size_t(n_scenes)) * 4;
  sizes[9] = size_t(n_scenes) * 2 * FH * AUX_SCENE_W * 2;
  return 0;
}

extern "C" int dsdf_fused_dgrad(const dsdf_net_desc* desc, const void* pack, const dsdf_train_bufs* bufs, const float* xyz,
                                int32_t xyz_stride, int64_t n_rows, int32_t training, uint32_t store_mask,
                                uint32_t sum_mask, void* stream) {
  if (int e = require_sm100()) return e;
  FusedLayout lo;
  if (int e = fused_layout(desc, &lo)) return e;
  DSDF_REQUIRE(n_rows > 0 && n_rows % TILE_M == 0, "fused training needs n_rows to be a positive multiple of 128");
  const uint8_t* p = static_cast<const uint8_t*>(pack);
  BwdArgs a;
  memset(&a, 0, sizeof(a));
  ChainLayer fwd_unused[DSDF_MAX_LAYERS];
  {
    // same program builder as the forward (fused_host.cu keeps it static); rebuild the backward half here
    for (int l = 0; l < lo.num_hidden; ++l) {
      ChainLayer& B = a.prog[l];
      B.has_mma = (l > 0); B.w_row0 = lo.bwd_row0[l]; B.n_halves = lo.bwd_halves[l]; B.k_chunks = lo.bwd_kchunks[l];
      B.bias_slot = -1; B.has_xyz = (l == 0 || l == lo.lat_layer); B.dropout = (desc->dropout_mask >> l) & 1;
      B.out_dim = desc->out_dim[l];
    }
    (void)fwd_unused;
  }
  a.num_hidden = lo.num_hidden;
  a.n_rows = n_rows;
  a.n_tiles = int(n_rows / TILE_M);
  a.n_rows_pad = n_rows;
  // training: the pack variant with the dropout scales folded into W^T and w_last (the forward ran on its twin)
  a.w_last = reinterpret_cast<const float*>(p + (training ? lo.off_wlast_tr : lo.off_wlast));
  a.dy = bufs->dy;
  a.bitmask = bufs->bitmask;
  for (int l = 0; l < lo.num_hidden; ++l) {
    const bool sums = (sum_mask >> l) & 1u;
    a.colsum_slot[l] = sums ? l : -1;
    a.xyzsum_slot[l] = !sums ? -1 : ((l == 0) ? lo.num_hidden : (l == lo.lat_layer ? lo.num_hidden + 3 : -1));
  }
  a.xyz = xyz;
  DSDF_REQUIRE(xyz_stride >= 3, "xyz_stride must be >= 3");
  a.xyz_stride = xyz_stride;
  a.tile_part = bufs->tile_part;
  a.part_stride = n_slots(lo) * FH;
  a.store_dys = int32_t(store_mask & ((1u << lo.num_hidden) - 1u));
  DSDF_REQUIRE(a.store_dys == 0 || bufs->dys != nullptr, "store_mask needs bufs->dys");
  DSDF_REQUIRE(sum_mask == 0 || bufs->tile_part != nullptr, "sum_mask needs bufs->tile_part");
  CUtensorMap tmW, tmDys;
  if (int e = make_tmap_bf16_2d(&tmW, p + (training ? lo.off_wpt_tr : lo.off_wpt), FH, lo.bwd_rows, FH * 2, 64, 128)) return e;
  const void* dys_ptr = bufs->dys ? bufs->dys : static_cast<const void*>(p + lo.off_wpt);   // dummy map when nothing is stored
  const uint64_t dys_rows = bufs->dys ? uint64_t(lo.num_hidden) * n_rows : uint64_t(lo.bwd_rows);
  if (int e = make_tmap_bf16_2d(&tmDys, dys_ptr, FH, dys_rows, FH * 2, 64, TILE_M)) return e;
  return launch_chain_dgrad(tmW, tmDys, a, sm_count(), static_cast<cudaStream_t>(stream));
}

extern "C" int dsdf_fused_wgrad(const dsdf_net_desc* desc, const dsdf_train_bufs* bufs, int64_t n_rows, void* stream) {
  if (int e = require_sm100()) return e;
  FusedLayout lo;
  if (int e = fused_layout(desc, &lo)) return e;
  DSDF_REQUIRE(n_rows > 0 && n_rows % TILE_M == 0, "fused training needs n_rows to be a positive multiple of 128");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  WgradArgs a;
  memset(&a, 0, sizeof(a));
  if (int e = build_plan(desc, lo, n_rows, &a.plan)) return e;
  a.n_rows_pad = n_rows;
  a.partials = bufs->wg_part;
  CUtensorMap tmDys, tmActs;
  const uint64_t rows = uint64_t(lo.num_hidden) * n_rows;
  if (int e = make_tmap_bf16_2d(&tmDys, bufs->dys, FH, rows, FH * 2, 64, 32)) return e;
  if (int e = make_tmap_bf16_2d(&tmActs, bufs->acts, FH, rows, FH * 2, 64, 32)) return e;
  if (int e = launch_wgrad(tmDys, tmActs, a, st)) return e;
  const uint8_t* x_last = static_cast<const uint8_t*>(bufs->acts) + size_t(lo.num_hidden - 1) * n_rows * FH * 2;
  return launch_wgrad_last(x_last, bufs->dy, n_rows, bufs->last_part, kLastBlocks, st);
}

extern "C" int dsdf_fused_finish(const dsdf_net_desc* desc, const void* pack, const dsdf_params* params,
                                 const dsdf_train_bufs* bufs, const int32_t* tile_scene, int64_t n_rows,
                                 const float* lat_table, const int64_t* scene_ids, int32_t n_scenes, int64_t n_scenes_total,
                                 float reg_lambda_eff, float inv_n, int32_t training, int32_t accumulate, int32_t phases,
                                 const dsdf_param_grads* pgrads, float* lat_grad, float* loss_out, void* stream) {
  if (int e = require_sm100()) return e;
  FinishArgs a;
  memset(&a, 0, sizeof(a));
  if (int e = fused_layout(desc, &a.lo)) return e;
  if (int e = build_plan(desc, a.lo, n_rows, &a.plan)) return e;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  a.desc = *desc;
  a.params = *params;
  a.grads = *pgrads;
  a.pack = static_cast<const uint8_t*>(pack);
  a.tile_part = bufs->tile_part;
  a.tile_scene = tile_scene;
  a.n_tiles = int(n_rows / TILE_M);
  a.n_slots = n_slots(a.lo);
  a.n_scenes = n_scenes;
  a.scene_sums = bufs->scene_sums;
  a.scene_reg = bufs->scene_sums + size_t(n_scenes) * a.n_slots * FH;
  a.wg_part = bufs->wg_part;
  a.last_part = bufs->last_part;
  a.n_last_blocks = 1;               // last_reduce_kernel folds the kLastBlocks partials into row 0 first
  a.lat_table = lat_table;
  a.scene_ids = scene_ids;
  a.lat_grad = lat_grad;
  a.reg_scale = reg_lambda_eff * inv_n;
  a.inv_n = inv_n;
  a.tile_loss = bufs->tile_loss;
  a.loss_out = loss_out;
  a.xyz_slot0 = a.lo.num_hidden;
  a.accumulate = accumulate;
  for (int l = 0; l <= a.lo.num_hidden; ++l) a.in_scale[l] = training ? a.lo.in_scale[l] : 1.0f;
  for (int l = 0; l < desc->num_linear; ++l) {
    DSDF_REQUIRE(pgrads->v[l] && pgrads->b[l], "missing gradient pointers");
    if ((desc->weight_norm_mask >> l) & 1) DSDF_REQUIRE(pgrads->g[l] && params->g[l], "missing g pointers");
  }
  const bool do_scene = (phases & DSDF_FINISH_SCENE) != 0, do_params = (phases & DSDF_FINISH_PARAMS) != 0;
  DSDF_REQUIRE(do_scene || do_params, "dsdf_fused_finish: phases must name DSDF_FINISH_SCENE and / or DSDF_FINISH_PARAMS");
  if (do_params) {
    DSDF_LAUNCHED;
    last_reduce_kernel<<<(FH + 1 + 31) / 32, 1024, 0, st>>>(bufs->last_part, kLastBlocks);
  }
  if (do_scene) {
    DSDF_LAUNCHED;
    scene_sums_kernel<<<dim3(n_scenes, a.n_slots), FH, 0, st>>>(a.tile_part, tile_scene, a.n_tiles, a.n_slots, a.scene_sums);
  }
  if (lat_grad && do_scene) {
    if (!accumulate) DSDF_CHECK_CUDA(cudaMemsetAsync(lat_grad, 0, size_t(n_scenes_total) * desc->latent_size * sizeof(float), st));
    DSDF_LAUNCHED;
    latent_grad_kernel<<<dim3(n_scenes, (desc->latent_size + 7) / 8), 256, 0, st>>>(a);
  }
  int max_out = 0;
  for (int l = 0; l < desc->num_linear; ++l) max_out = desc->out_dim[l] > max_out ? desc->out_dim[l] : max_out;
  if (do_params) {
    DSDF_LAUNCHED;
    assemble_grads_kernel<<<dim3(max_out, desc->num_linear), 128, 0, st>>>(a);
  }
  if (loss_out && do_params) {
    DSDF_REQUIRE(lat_grad || reg_lambda_eff == 0.f, "loss with regulariser needs the latent-grad pass");
    DSDF_LAUNCHED;
    loss_kernel<<<1, 256, 0, st>>>(a);
  }
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_debug_wgrad_clusters(void) { return wgrad_max_active_clusters(); }

extern "C" int dsdf_adam(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps,
                         int32_t step, float grad_scale, const float* grad_scale_dev, void* stream) {
  if (int e = require_sm100()) return e;
  if (n <= 0) return 0;
  DSDF_REQUIRE(step >= 1, "Adam step counter starts at 1");
  const float bc1 = float(1.0 - pow(double(beta1), double(step)));
  const float bc2 = float(1.0 - pow(double(beta2), double(step)));
  DSDF_LAUNCHED;
  adam_kernel<<<unsigned((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(p, g, m, v, n, lr, beta1, beta2, eps, bc1,
                                                                                       sqrtf(bc2), grad_scale, grad_scale_dev);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_adam_state(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps,
                               const dsdf_step_state* state, float grad_scale, const float* grad_scale_dev, void* stream) {
  if (int e = require_sm100()) return e;
  if (n <= 0) return 0;
  DSDF_REQUIRE(state != nullptr, "dsdf_adam_state needs the device step state");
  DSDF_LAUNCHED;
  adam_state_kernel<<<unsigned((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(p, g, m, v, n, lr, beta1, beta2, eps, state,
                                                                                             grad_scale, grad_scale_dev);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_step_advance(dsdf_step_state* state, float beta1, float beta2, void* stream) {
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(state != nullptr, "dsdf_step_advance: null state");
  DSDF_LAUNCHED;
  step_advance_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(state, double(beta1), double(beta2), 0, 0u);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_step_set(dsdf_step_state* state, uint32_t step, void* stream) {
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(state != nullptr, "dsdf_step_set: null state");
  DSDF_LAUNCHED;
  step_advance_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(state, 0.9, 0.999, 1, step);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_clip_coef(const float* g, int64_t n, float max_norm, float* scratch, float* out, void* stream) {
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(n > 0 && g && scratch && out, "dsdf_clip_coef: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DSDF_LAUNCHED;
  sumsq_partial_kernel<<<1024, 256, 0, st>>>(g, n, scratch);
  DSDF_LAUNCHED;
  clip_coef_kernel<<<1, 1024, 0, st>>>(scratch, max_norm, out);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}
