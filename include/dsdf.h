/* deepsdf_b200 C ABI  (libdsdf_b200.so)
 *
 * Drop-in boundary for the DeepSDF auto-decoder hot path of mkofler96/DeepSDF.  Every entry point takes plain
 * pointers and sizes only (CUDA *device* pointers owned by the caller unless stated otherwise), an explicit
 * cudaStream_t passed as void*, and returns 0 on success or a non-zero code; dsdf_last_error() then holds a
 * thread-local message.  The library never allocates device memory behind the caller's back: every scratch
 * buffer is a caller-provided workspace whose size a *_bytes() query returns.  sm_100a only - any other device
 * is an error, there is no CPU or other-GPU fallback.
 *
 * What each group replaces in the reference (paths relative to the reference root):
 *   net desc / params        Decoder.__init__                  deep_sdf/networks/deep_sdf_decoder.py:10-73
 *   dsdf_generic_forward     Decoder.forward                   deep_sdf/networks/deep_sdf_decoder.py:76-111
 *   dsdf_generic_backward    autograd of the above             train_deep_sdf.py:533, deep_sdf/mesh.py:405-422
 *   dsdf_pack_weights        torch._weight_norm per call       deep_sdf/networks/deep_sdf_decoder.py:50-55
 *   dsdf_scene_prepare       lat_vecs(indices) + max_norm      train_deep_sdf.py:385,509
 *   dsdf_fused_forward       decoder(input) (+clamp, L1)       train_deep_sdf.py:511-519 ; deep_sdf/utils.py:54-65
 *   dsdf_fused_dgrad/wgrad   chunk_loss.backward()             train_deep_sdf.py:533
 *   dsdf_fused_finish        weight_norm backward, latent grad, code regulariser   train_deep_sdf.py:523-533
 *   dsdf_adam                optimizer_all.step()              train_deep_sdf.py:400-411,545
 *   dsdf_grid_points         create_mesh sample grid           deep_sdf/mesh.py:37-53
 *   dsdf_sample_batch        SDFSamples.__getitem__ + DataLoader collate      deep_sdf/data.py:75-110,142-194
 *   dsdf_fused_input_grad    autograd.functional.jvp loop      deep_sdf/mesh.py:405-422
 *   dsdf_fused_latent_finish reconstruct.py latent optimisation (upstream; README.md:134-142)
 */
#ifndef DSDF_B200_H
#define DSDF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DSDF_MAX_LAYERS 16
#define DSDF_ABI_VERSION 3

/* Network description == the reference's NetworkSpecs after Decoder.__init__ resolved the layer sizes. */
typedef struct dsdf_net_desc {
  int32_t latent_size;                 /* L  (CodeLength) */
  int32_t geom_dim;                    /* D  (geom_dimension) */
  int32_t num_linear;                  /* number of nn.Linear layers = len(dims)+1 */
  int32_t in_dim[DSDF_MAX_LAYERS];     /* lin{l}.in_features  */
  int32_t out_dim[DSDF_MAX_LAYERS];    /* lin{l}.out_features */
  uint32_t weight_norm_mask;           /* bit l: lin{l} is weight-normed (g,v) */
  uint32_t layernorm_mask;             /* bit l: bn{l} LayerNorm applied after lin{l} (hidden layers only) */
  uint32_t dropout_mask;               /* bit l: F.dropout after hidden layer l */
  uint32_t latent_in_mask;             /* bit l: x = cat([x, input]) before lin{l} */
  int32_t xyz_in_all;                  /* x = cat([x, xyz]) before every lin{l}, l != 0, not in latent_in */
  int32_t use_tanh;                    /* extra tanh on the last layer (the final self.th tanh is always applied) */
  int32_t latent_dropout;              /* F.dropout(latent, 0.2) on the first-layer copy of the latent */
  float dropout_prob;
} dsdf_net_desc;

/* Parameter pointers in the reference's own layout (fp32, row-major [out,in]); g == NULL for plain layers. */
typedef struct dsdf_params {
  const float* v[DSDF_MAX_LAYERS];     /* lin{l}.weight  or  lin{l}.parametrizations.weight.original1 */
  const float* g[DSDF_MAX_LAYERS];     /* lin{l}.parametrizations.weight.original0 [out]  or NULL */
  const float* b[DSDF_MAX_LAYERS];     /* lin{l}.bias */
  const float* ln_w[DSDF_MAX_LAYERS];  /* bn{l}.weight or NULL */
  const float* ln_b[DSDF_MAX_LAYERS];  /* bn{l}.bias   or NULL */
} dsdf_params;

typedef struct dsdf_param_grads {      /* same shapes as dsdf_params; written (not accumulated) unless stated */
  float* v[DSDF_MAX_LAYERS];
  float* g[DSDF_MAX_LAYERS];
  float* b[DSDF_MAX_LAYERS];
  float* ln_w[DSDF_MAX_LAYERS];
  float* ln_b[DSDF_MAX_LAYERS];
} dsdf_param_grads;

/* Device-resident per-step scalars (64 bytes, caller-allocated device memory): what changes from one optimiser step to the next.
 * Kernels read them from memory instead of from their parameters, which makes the whole training step replayable as one CUDA
 * graph.  dsdf_step_advance: step += 1 and Adam's bias corrections bc1 = 1 - beta1^step, bc2_sqrt = sqrt(1 - beta2^step)
 * (computed in double, like torch.optim.Adam's host arithmetic).  dsdf_step_set: set the counter (resume from a checkpoint). */
typedef struct dsdf_step_state {
  uint32_t step;
  float bc1, bc2_sqrt;
  uint32_t reserved[13];
} dsdf_step_state;
int dsdf_step_advance(dsdf_step_state* state, float beta1, float beta2, void* stream);
int dsdf_step_set(dsdf_step_state* state, uint32_t step, void* stream);

int dsdf_abi_version(void);
const char* dsdf_last_error(void);
/* 0 if the current CUDA device is sm_100; error otherwise (used by the Python side to fail loudly). */
int dsdf_check_device(void);
/* kernels launched by this library since it was loaded (fused, small-network, sampler and consumer paths) */
int64_t dsdf_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * Dropout: Philox4x32-7, counter (row, layer, 2*(col/32)+k), key seed; 8 random bits per decision, bit-sliced over 32
 * columns; keep iff the 8-bit draw >= round(p*256) (csrc/common.cuh).  Survivors are scaled by dsdf_dropout_scale(p) =
 * 1 / (1 - effective p); on the fused path that factor is folded into the NEXT layer's packed weights (training variant of
 * the pack), so the stashed activations are relu(h) * keep, unscaled.
 * Writes the {0,1} keep mask of one hidden layer as uint8 [n_rows, width] (tests / oracle hand-over). */
int dsdf_dropout_mask(uint64_t seed, uint32_t layer, uint64_t row0, int64_t n_rows, int32_t width, float p,
                      uint8_t* mask_out, void* stream);
/* effective drop probability round(p*256)/256 actually applied (host function) */
float dsdf_dropout_effective_p(float p);
/* the factor survivors are multiplied by: 1 / (1 - dsdf_dropout_effective_p(p)), so that E[keep * scale] == 1 exactly and
 * train- and eval-mode activations agree in expectation like F.dropout's (deep_sdf_decoder.py:105-106).  Host function. */
float dsdf_dropout_scale(float p);

/* ------------------------------------------------------------------------------------------------
 * Generic path: any NetworkSpecs (small nets, LayerNorm, xyz_in_all, geom_dim != 3).  fp32 CUDA-core kernels.
 * forward:  input [N, L+D] fp32 -> out [N] fp32.  `save` != 0 keeps what backward needs inside `workspace`.
 * backward: grad_out [N]; grad_input [N, L+D] or NULL; pgrads or NULL.  accumulate != 0 adds into pgrads. */
size_t dsdf_generic_workspace_bytes(const dsdf_net_desc* desc, int64_t n_rows);
int dsdf_generic_forward(const dsdf_net_desc* desc, const dsdf_params* params, const float* input, int64_t n_rows,
                         int32_t training, uint64_t seed, uint64_t row0, float* out, void* workspace,
                         size_t workspace_bytes, void* stream);
int dsdf_generic_backward(const dsdf_net_desc* desc, const dsdf_params* params, const float* input, int64_t n_rows,
                          int32_t training, uint64_t seed, uint64_t row0, const float* grad_out, float* grad_input,
                          const dsdf_param_grads* pgrads, int32_t accumulate, void* workspace, size_t workspace_bytes,
                          void* stream);

/* ------------------------------------------------------------------------------------------------
 * Small-network path (csrc/smallnet.cu): hidden widths <= 128, weight-norm or plain layers, any latent_in set, CodeLength +
 * geom_dimension <= 16, no LayerNorm / xyz_in_all / latent_dropout - the 4x32 / 4x64 / 6x128 decoders of the reference's
 * optimisation runs (optimization_runs/config.json:8).  ONE kernel, eval mode: the 128-row tile stays in shared memory
 * across all layers (fp32 FFMA), and with G != NULL the same kernel walks the chain back and writes
 * G[r] = dSDF/d[latent, xyz] of row r - what the L x n_cpt jvp loop of deep_sdf/mesh.py:405-422 re-derives per tangent.
 *   pack     caller buffer of dsdf_small_pack_bytes(), 16-byte aligned, rebuilt by dsdf_small_pack_weights when the
 *            parameters change (fold g*v/||v||, padded W and W^T per layer)
 *   queries  row r at queries + r*ldq: [latent(L), xyz(D)]; with `latent` != NULL the rows hold xyz only and the one code
 *            latent[L] is broadcast (decode_sdf with a latent_vector, deep_sdf/utils.py:57-61; create_mesh)
 *   G        [n_rows][ldg] fp32 or NULL, ldg >= L + D */
int dsdf_small_supported(const dsdf_net_desc* desc);
size_t dsdf_small_pack_bytes(const dsdf_net_desc* desc);
int dsdf_small_pack_weights(const dsdf_net_desc* desc, const dsdf_params* params, void* pack, void* stream);
int dsdf_small_forward(const dsdf_net_desc* desc, const void* pack, const float* queries, int64_t ldq, const float* latent,
                       int64_t n_rows, float* sdf_out, float* G, int64_t ldg, void* stream);
/* One iteration's loss and latent gradient of the frozen-decoder latent optimisation (upstream reconstruct.py) on a small
 * network: samples [n_shapes * rows_per_shape][lds >= 4] (x, y, z, sdf) shape-major, one code per shape from lat_table
 * [n_shapes][L];   lat_grad[s] = d/dz ( mean |clamp(pred) - clamp(gt)| + l2reg * mean(z^2) ),  shape_loss[s] = that
 * expression.  Scratch: sdf [rows], G [rows][L+3], row_loss [rows].  Deterministic (fixed summation order). */
int dsdf_small_latent_step(const dsdf_net_desc* desc, const void* pack, const float* samples, int64_t lds,
                           const float* lat_table, int32_t n_shapes, int64_t rows_per_shape, float clamp, float l2reg,
                           float* sdf_scratch, float* G_scratch, float* row_loss_scratch, float* lat_grad, float* shape_loss,
                           void* stream);

/* ------------------------------------------------------------------------------------------------
 * Fused tcgen05 path (hidden width 512, weight-norm or plain layers, one latent_in layer, geom_dim 3).
 *
 * Scene mode: every 128-row tile of points belongs to ONE scene, so the latent columns of lin0 and of the
 * latent_in layer collapse into a per-scene bias vector (dsdf_scene_prepare) and are never materialised per row.
 *   pack       caller buffer of dsdf_pack_bytes(), 1024-byte aligned; rebuilt by dsdf_pack_weights whenever the
 *              parameters change (fold g*v/||v||, bf16 pack of W and W^T, fp32 side tables).
 *   scene_aux  [n_scenes][2][512][16] bf16, written by dsdf_scene_prepare
 *   tile_scene [ceil(n_rows/128)] int32: scene (row of scene_aux) each tile uses; the two tiles of a CTA pair may differ
 */
typedef struct dsdf_train_io {
  const float* b_last;     /* device pointer to the last layer's bias [1] (always required) */
  int32_t training;        /* 1: dropout on, masks/activations/loss outputs below are honoured */
  uint64_t seed;           /* Philox key */
  uint64_t row0;           /* global row id of local row 0 (dropout stream is keyed by global row) */
  const float* gt;         /* ground-truth SDF or NULL; value of row r at gt[r * gt_stride] */
  int32_t gt_stride;       /* 1 for a dense [n_rows] vector, 4 when gt points at column 3 of the reference's [N,4] samples;
                              0 is read as 1 */
  float clamp;             /* ClampingDistance; <= 0 disables clamping */
  float inv_n;             /* 1 / (global batch rows): loss normaliser, train_deep_sdf.py:519 */
  float* dy_out;           /* [n_rows_pad] dL/d(pre-tanh output) */
  float* tile_loss;        /* [n_tiles] sum |pred - gt| per tile (not yet divided by N) */
  uint32_t* bitmask;       /* [num_hidden][n_rows_pad][16] relu*dropout keep bits, or NULL */
  void* acts;              /* bf16 [num_hidden][n_rows_pad][512] post-activation stash for wgrad, or NULL */
  int64_t n_rows_pad;      /* n_rows rounded up to a multiple of 128 */
  int32_t dropout_off;     /* training=1 with dropout disabled: an eval-mode forward that still records the relu gates
                              (what the input-gradient path and the frozen-decoder latent optimisation need) */
  int32_t dy_unit;         /* gt == NULL: dy_out = d sdf / d(pre-tanh output), the seed of the input gradient */
  const int32_t* tile_rows;/* [n_tiles] valid rows per 128-row tile (ragged scenes padded to tiles), or NULL = all */
  const struct dsdf_step_state* step_state;   /* device-resident optimiser step counter, or NULL: the dropout stream of row r is then
                              keyed by row0 + (step << 40) + r, so that a step captured in a CUDA graph draws fresh masks at
                              every replay without a single changed kernel parameter */
} dsdf_train_io;

int dsdf_fused_supported(const dsdf_net_desc* desc);
/* 1 if dsdf_fused_forward_rows (per-row latent codes) supports this network: CodeLength <= 16 (codes through the aux step) or a
 * long code whose latent_in layer has <= 256 hidden inputs, e.g. the original DeepSDF L = 256 (codes as bf16 K chunks) */
int dsdf_fused_rows_supported(const dsdf_net_desc* desc);
size_t dsdf_pack_bytes(const dsdf_net_desc* desc);
int dsdf_pack_weights(const dsdf_net_desc* desc, const dsdf_params* params, void* pack, void* stream);
/* The same with a choice of what to (re)write.  The big bf16 blocks and w_last exist in two variants: EVAL (plain folded
 * weights: eval-mode forwards, input gradients, latent optimisation) and TRAIN (the inverted-dropout scale of hidden layer
 * l-1 multiplied into the activation columns of layer l: the training step).  A training loop re-packs DSDF_PACK_TRAIN only.
 * DSDF_PACK_ZERO also zeroes the padded regions; needed once per buffer (dsdf_pack_weights = all three). */
#define DSDF_PACK_EVAL 1u
#define DSDF_PACK_TRAIN 2u
#define DSDF_PACK_ZERO 4u
int dsdf_pack_weights_ex(const dsdf_net_desc* desc, const dsdf_params* params, void* pack, uint32_t flags, void* stream);
/* nn.Embedding(max_norm) renorm of the listed rows (in place, max_norm <= 0: off) + the per-scene aux rows
 * scene_aux [n_scenes][2][512][16] bf16 (dsdf_scene_aux_bytes, 16-byte aligned; NULL: renorm only): per output feature the
 * scene's bias of lin0 / the latent_in layer (layer bias + latent columns . z, fp32, split hi + lo) and the layer's xyz
 * columns, in the K-column order of the forward chain's aux tensor-core step (csrc/fused.cuh). */
size_t dsdf_scene_aux_bytes(int32_t n_scenes);
int dsdf_scene_prepare(const dsdf_net_desc* desc, const void* pack, float* lat_table, const int64_t* scene_ids,
                       int32_t n_scenes, float max_norm, void* scene_aux, void* stream);
/* xyz: row r at xyz + r * xyz_stride (3 floats).  xyz_stride = 3 for [N,3] queries, 4 to read the reference's [N,4]
 * (xyz, sdf) sample rows in place (deep_sdf/data.py:102-110, train_deep_sdf.py:483-490). */
int dsdf_fused_forward(const dsdf_net_desc* desc, const void* pack, const void* scene_aux, const int32_t* tile_scene,
                       const float* xyz, int32_t xyz_stride, int64_t n_rows, float* sdf_out, const dsdf_train_io* io,
                       void* stream);
/* High-precision eval forward on the tensor cores (csrc/chain_exact.cu): the "1e-5" mode of Decoder.forward
 * (deep_sdf/networks/deep_sdf_decoder.py:76-111) for grid evaluation (deep_sdf/mesh.py:37-73) and decode_sdf with one code
 * (deep_sdf/utils.py:54-65).  Activations, weights and aux operands are carried as (hi, lo) fp16 pairs (~22 mantissa bits), all
 * partial products accumulate in fp32: <= 1e-5 against the fp32 arithmetic at natural output scale, at ~4x the tensor work of
 * dsdf_fused_forward per point.  Scene mode, eval only.
 *   exact_pack   dsdf_exact_pack_bytes() bytes, written by dsdf_exact_pack from the parameters (weight norm folded, hidden
 *                columns and layer biases split hi / lo)
 *   scene_aux    dsdf_scene_aux_bytes(n_scenes) bytes, written by dsdf_exact_scene_aux (the fp16 counterpart of the rows
 *                dsdf_scene_prepare writes; no renorm - run dsdf_scene_prepare first if max_norm applies)
 *   pack         the regular pack (fp32 side tables, last layer) */
size_t dsdf_exact_pack_bytes(const dsdf_net_desc* desc);
int dsdf_exact_pack(const dsdf_net_desc* desc, const dsdf_params* params, void* exact_pack, void* stream);
int dsdf_exact_scene_aux(const dsdf_net_desc* desc, const void* pack, const float* lat_table, const int64_t* scene_ids,
                         int32_t n_scenes, void* scene_aux, void* stream);
int dsdf_fused_forward_exact(const dsdf_net_desc* desc, const void* pack, const void* exact_pack, const void* scene_aux,
                             const int32_t* tile_scene, const float* xyz, int32_t xyz_stride, int64_t n_rows,
                             const float* b_last, float* sdf_out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Fused training step (scene mode).  All buffers are caller-owned device memory; dsdf_train_buf_sizes() fills
 * sizes[10] (bytes) in the order of the struct members below (+ sizes[9] = scene_aux).  n_rows must be a multiple of 128
 * and every 128-row tile must belong to one scene (SamplesPerScene % 128 == 0: 16384 and 16000 both are).
 *   forward (training=1, gt given)  -> sdf, dy, tile_loss, bitmask, acts
 *   dsdf_fused_dgrad                -> dys (bf16 dY per hidden layer), tile_part (per-tile column sums)
 *   dsdf_fused_wgrad                -> wg_part (split-K partials of dW), last_part (last layer)
 *   dsdf_fused_finish               -> parameter grads in the reference layout (dg, dv, db), dense latent grad
 *                                      [n_scenes_total, L] (zeroed, rows of scene_ids written; scene_ids must be
 *                                      unique within a batch), loss_out[0] = loss incl. regulariser, [1] = data term.
 *                                      reg_lambda_eff = CodeRegularizationLambda * min(1, epoch/100) or 0.
 *                                      training != 0: the step ran with dropout (the stash is unscaled; dW gets the scales).
 *                                      accumulate != 0: add to the gradients and to loss_out instead of overwriting them - the
 *                                      chunks of --batch_split (train_deep_sdf.py:495-535); a scene may then span chunks.
 * Every reduction has a fixed order: two runs on the same inputs give bit-identical gradients. */
typedef struct dsdf_train_bufs {
  void* acts;            /* bf16 [num_hidden][n_rows][512] */
  void* dys;             /* bf16 [num_hidden][n_rows][512] */
  uint32_t* bitmask;     /* [num_hidden][n_rows][16] */
  float* dy;             /* [n_rows] */
  float* tile_loss;      /* [n_rows/128] */
  float* tile_part;      /* [n_rows/128][num_hidden+6][512] */
  float* wg_part;
  float* last_part;
  float* scene_sums;
} dsdf_train_bufs;

int dsdf_train_buf_sizes(const dsdf_net_desc* desc, int64_t n_rows, int32_t n_scenes, size_t* sizes);
/* store_mask: bit t set = keep the bf16 dY tile of hidden layer t in bufs->dys (training: every t > 0 =
 * DSDF_STORE_DYS_TRAIN; input gradient: layer 0 and the latent_in layer; latent-only optimisation: 0).
 * sum_mask: bit t set = emit the per-tile column sums of dY_t (and the xyz-weighted ones of lin0 / latent_in layer)
 * into bufs->tile_part (training: all layers; latent-only optimisation: layer 0 and the latent_in layer). */
/* dsdf_fused_finish phases.  SCENE needs only dsdf_fused_dgrad's per-tile sums (per-scene sums, dense latent gradient): a
 * caller may run it on a second stream beside dsdf_fused_wgrad.  PARAMS needs SCENE and dsdf_fused_wgrad (decoder gradients,
 * loss).  Both bits = the whole finish in order. */
#define DSDF_FINISH_SCENE 1
#define DSDF_FINISH_PARAMS 2
#define DSDF_FINISH_ALL 3
#define DSDF_STORE_DYS_TRAIN 0xFFFFFFFEu
#define DSDF_SUMS_ALL 0xFFFFFFFFu
int dsdf_fused_dgrad(const dsdf_net_desc* desc, const void* pack, const dsdf_train_bufs* bufs, const float* xyz,
                     int32_t xyz_stride, int64_t n_rows, int32_t training, uint32_t store_mask, uint32_t sum_mask,
                     void* stream);
int dsdf_fused_wgrad(const dsdf_net_desc* desc, const dsdf_train_bufs* bufs, int64_t n_rows, void* stream);
int dsdf_fused_finish(const dsdf_net_desc* desc, const void* pack, const dsdf_params* params,
                      const dsdf_train_bufs* bufs, const int32_t* tile_scene, int64_t n_rows, const float* lat_table,
                      const int64_t* scene_ids, int32_t n_scenes, int64_t n_scenes_total, float reg_lambda_eff,
                      float inv_n, int32_t training, int32_t accumulate, int32_t phases, const dsdf_param_grads* pgrads,
                      float* lat_grad, float* loss_out, void* stream);
/* Row-latent forward: every row brings its own latent code (queries[r] = [latent(L), xyz(3)], row stride ldq floats) - the
 * microstructure consumers of deep_sdf/mesh.py:457-497, whose codes come from a spline over the domain (L = 1 or 2), and
 * decoder(input[N, L+3]) calls of the L = 16 experiments; supported: L <= 16.  Same kernel and same dsdf_train_io as
 * dsdf_fused_forward; the latent columns of lin0 and of the latent_in layer enter through the aux tensor-core step like the
 * xyz columns (L <= 4: hi/lo-split products, ~16 mantissa bits; 4 < L <= 16: one bf16 term per component).  Followed by dsdf_fused_dgrad (xyz = queries + L,
 * xyz_stride = ldq) and dsdf_fused_input_grad it yields G = dSDF/d[latent, xyz] for per-point codes. */
int dsdf_fused_forward_rows(const dsdf_net_desc* desc, const void* pack, const float* queries, int32_t ldq, int64_t n_rows,
                            float* sdf_out, const dsdf_train_io* tr, void* stream);
/* ------------------------------------------------------------------------------------------------
 * Input gradient (the derivative consumers of deep_sdf/mesh.py:405-454 and analysis/: dSDF/dlatent, dSDF/dxyz).
 * After  dsdf_fused_forward(training=1, dropout_off=1, dy_unit=1)  and  dsdf_fused_dgrad(store_mask = layer 0 |
 * latent_in layer)  this contracts the stashed dY tiles of those two layers with the latent / xyz columns of their
 * weights on the tensor cores:   G[r, 0:L] = dSDF/dlatent,  G[r, L:L+3] = dSDF/dxyz   - the column order of the
 * reference's decoder input [latent, xyz] (deep_sdf/utils.py:57-63), so G == input.grad of Decoder.forward.
 * G is fp32 [n_rows][ldg], ldg >= L+3. */
int dsdf_fused_input_grad(const dsdf_net_desc* desc, const void* pack, const dsdf_train_bufs* bufs, int64_t n_rows,
                          float* G, int64_t ldg, void* stream);

/* create_mesh's sample grid (deep_sdf/mesh.py:37-53): linear index i = (ix*N + iy)*N + iz  ->  xyz = -1 + idx*2/(N-1).
 * Writes rows [i0, i0+n) of the N^3 grid as fp32 [n][3] (a z-slab of the slowest index is a contiguous index range). */
int dsdf_grid_points(int32_t N, int64_t i0, int64_t n, float* xyz_out, void* stream);

/* Frozen-decoder latent optimisation (upstream DeepSDF reconstruct.py; README.md:134-142 of the reference): after
 * forward(training=1, dropout_off=1, gt, clamp, inv_n = 1/rows_per_shape) + dgrad(sum_mask = layer 0 | latent_in layer)
 * reduce the per-tile column sums per shape and form   lat_grad[s] = d/dz ( mean|clamp(pred)-clamp(gt)| +
 * l2reg * mean(z^2) )   and   shape_loss[s] (same expression).  lat_grad is [n_scenes][L] (row s = scene_ids[s]'s
 * gradient, dense in batch order), deterministic. */
int dsdf_fused_latent_finish(const dsdf_net_desc* desc, const void* pack, const dsdf_train_bufs* bufs,
                             const int32_t* tile_scene, int64_t n_rows, const float* lat_table, const int64_t* scene_ids,
                             int32_t n_scenes, float l2reg, float inv_n, float* lat_grad, float* shape_loss,
                             void* stream);

/* ------------------------------------------------------------------------------------------------
 * Device-resident dataset + sampler (deep_sdf/data.py:75-110 unpack_sdf_samples, :142-194 SDFSamples.__getitem__).
 *   store [rows][4] fp32 (x, y, z, sdf): per scene its positive rows then its negative rows, NaN rows dropped (remove_nans)
 *   meta  [n_scenes_total][4] int64: {pos_off, n_pos, neg_off, n_neg} in rows
 * Writes the scene-major batch [n_scenes * samples_per_scene][4]: per scene `half` positive then `half` negative rows (or
 * all of the scarcer sign and the rest of the other, data.py:92-99), drawn without replacement by a keyed bijection of
 * [0, n) (same role as torch.randperm(n)[:take]); rows beyond what the scene can supply are zero. */
int dsdf_sample_batch(const float* store, const int64_t* meta, const int64_t* scene_ids, int32_t n_scenes,
                      int32_t samples_per_scene, uint64_t seed, float* batch_out, void* stream);
/* the bijection itself, on the host: source row of batch position i within the scene's positive / negative rows */
uint32_t dsdf_sample_perm(uint32_t i, uint32_t n, uint64_t seed, int64_t scene, int32_t negative);

/* ------------------------------------------------------------------------------------------------
 * Pre/post kernels of the microstructure evaluation (deep_sdf/mesh.py:457-514, csrc/microstructure.cu).
 * dsdf_tile_queries: queries[r] = [lat[r, 0:L], T(x), T(y), T(z)] with the tiling transform of prepare_samples
 *   (mesh.py:470-477: p = 2/t, T(x) = (2/p) * |((x - t%2) mod 2p) - p| - 1) of samples_orig [n][ld_s]; rows [n, n_pad) are
 *   zero-filled (pad to 128-row tiles).  tiling3: HOST array of 3 floats.
 * dsdf_cap_sdf: the border caps of evaluate_network (mesh.py:499-514) applied in place to sdf [n]: for every entry
 *   border = (x[dim] - multiplier*(1-measure)) * -multiplier; cap == -1: sdf = max(sdf, -border), cap == +1: sdf =
 *   min(sdf, border); then sdf = max(sdf, distance outside the unit cube).  Where a cap decides the value the latent
 *   derivative vanishes: if G != NULL its columns [0, L) of that row are zeroed.  dims/multipliers/measures/caps: HOST. */
int dsdf_tile_queries(const float* samples_orig, int32_t ld_s, const float* lat, int32_t ld_l, int32_t latent_size, int64_t n,
                      int64_t n_pad, const float* tiling3, float* queries, void* stream);
int dsdf_cap_sdf(float* sdf, float* G, int64_t ldg, int32_t latent_size, const float* samples_orig, int32_t ld_s, int64_t n,
                 int32_t n_caps, const int32_t* dims, const double* multipliers, const double* measures, const int32_t* caps,
                 void* stream);

/* ------------------------------------------------------------------------------------------------
 * Iso-surface of the device-resident SDF volume (the tail of create_mesh: deep_sdf/mesh.py:88-155 hands a host copy of
 * the [N,N,N] tensor to skimage.measure.marching_cubes).  Marching TETRAHEDRA over the six Kuhn tetrahedra of every cube
 * (csrc/isosurface.cu): a watertight triangle soup, not skimage's Lewiner triangulation.
 * volume[(ix*ny + iy)*nz + iz] fp32 (create_mesh's ordering), inside <=> value < level.
 *   1. dsdf_iso_count  -> block_counts[dsdf_iso_num_blocks(nx,ny,nz)] (int32 triangles per block of 256 cubes)
 *   2. the caller forms block_offsets = exclusive prefix sum (int64) and allocates verts [3 * n_triangles][3] fp32
 *   3. dsdf_iso_emit   -> verts: triangle t = rows 3t, 3t+1, 3t+2, normal towards increasing SDF;
 *                         vertex = origin3 + grid_coordinate * voxel3 (origin3 / voxel3 are HOST arrays of 3 floats). */
int64_t dsdf_iso_num_blocks(int32_t nx, int32_t ny, int32_t nz);
int dsdf_iso_count(const float* volume, int32_t nx, int32_t ny, int32_t nz, float level, int32_t* block_counts, void* stream);
int dsdf_iso_emit(const float* volume, int32_t nx, int32_t ny, int32_t nz, float level, const float* origin3,
                  const float* voxel3, const int64_t* block_offsets, float* verts, void* stream);

/* torch.optim.Adam update of one flat fp32 tensor (dense, no weight decay); step counts from 1; g is multiplied by
 * grad_scale and, if grad_scale_dev != NULL, by the device scalar *grad_scale_dev first (gradient clipping without a
 * host round trip). */
int dsdf_adam(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps,
              int32_t step, float grad_scale, const float* grad_scale_dev, void* stream);
/* the same update with the bias corrections read from the device-resident step state: replayable inside a CUDA graph */
int dsdf_adam_state(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2, float eps,
                    const dsdf_step_state* state, float grad_scale, const float* grad_scale_dev, void* stream);
/* All-reduce (sum, in place) of a flat fp32 buffer that every rank holds in symmetric memory mapped through one NVSwitch
 * multicast address (csrc/nvls_allreduce.cu): the gradient exchange of the multi-GPU training path (replaces nn.DataParallel's
 * reduction, train_deep_sdf.py:353) as this library's own kernel - multimem.ld_reduce / multimem.st on rank-owned slices between
 * two cross-GPU barriers on the symmetric signal pads.  multicast_ptr: multicast address of the buffer; signal_pad_ptrs_dev: device
 * array [world] of the ranks' signal-pad addresses (zero-initialised, >= dsdf_nvls_signal_words(world) 32-bit words each, used by
 * nothing else); n: floats, a multiple of 4.  Every rank of the group must make the call, stream-ordered behind its producers. */
int32_t dsdf_nvls_signal_words(int32_t world);
int dsdf_nvls_allreduce_f32(float* multicast_ptr, const void* signal_pad_ptrs_dev, int32_t rank, int32_t world, int64_t n,
                            void* stream);
/* torch.nn.utils.clip_grad_norm_ (train_deep_sdf.py:541-543) over one flat fp32 gradient: out[0] = ||g||_2,
 * out[1] = min(1, max_norm / (||g||_2 + 1e-6)).  Deterministic two-stage reduction; `scratch` holds 1024 floats. */
int dsdf_clip_coef(const float* g, int64_t n, float max_norm, float* scratch, float* out, void* stream);
/* The same coefficient without a caller workspace and without any allocation (csrc/noalloc.cu).  NO counterpart in the
 * reference: added for callers that cannot call cudaMalloc or hand over scratch - code
 * under CUDA-graph capture, tight serving loops, harnesses that forbid allocation.  The 1024 partial sums live in a file-scope
 * `static __device__ __align__(256)` array of 4 KB that every GPU loading the library holds until exit; its device address is
 * resolved once per device (cudaGetSymbolAddress with that device current) on the first call there.  n must be
 * <= DSDF_CLIP_STATIC_MAX_N, larger sizes are rejected with an error.  Calls no cudaMalloc / cudaFree / allocating routine.
 * Every call on a device shares the array: it keeps the last call's partials (each call overwrites all of them, nothing needs
 * clearing), and two calls that OVERLAP on the GPU (different streams, parallel graph branches) race - serialise them.
 * Bit-identical to dsdf_clip_coef (same block ranges, same summation order). */
#define DSDF_CLIP_STATIC_MAX_N (1LL << 28)
int dsdf_clip_coef_static(const float* g, int64_t n, float max_norm, float* out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * tcgen05 bring-up self test (see csrc/selftest.cu). */
int dsdf_selftest_umma(int mode, const void* A, const void* B, float* C, void* A_store_out, int K, int N,
                       unsigned a_lbo, unsigned a_sbo, unsigned b_lbo, unsigned b_sbo, void* stream);

/* diagnostics: how many 4-CTA wgrad clusters can be co-resident on the current device (-1: query failed) */
int dsdf_debug_wgrad_clusters(void);
/* CTA-pair (cta_group::2) bring-up: C[256,256] = A[256,K] * B[256,K]^T on one 2-CTA cluster; mode bit2 = A written by
 * threads and handed over with a remote mbarrier arrive (see csrc/selftest.cu). */
int dsdf_selftest_umma_pair(int mode, const void* A, const void* B, float* C, int K, void* stream);
/* aux-step bring-up: the pair GEMM plus aux_k (16 | 32) extra K columns in the SWIZZLE_NONE K-major canonical layout
 * (A_aux, B_aux: [256][32] bf16); lbo / sbo = descriptor offsets under test (the kernels use 2048 / 128). */
int dsdf_selftest_umma_aux(const void* A, const void* B, const void* A_aux, const void* B_aux, float* C, int K, int aux_k,
                           unsigned lbo, unsigned sbo, void* stream);
/* Tensor-pipe rate probe (bring-up / DESIGN.md evidence): `iters` x 8 K-chunks of cta_group::2 MMAs, M = 256, N = n (256 or
 * 128), issued back to back by one CTA pair from shared-memory operands; out_dev[0] = SM clocks from first issue to the last
 * commit, out_dev[1] = STS.128 issued meanwhile per thread when lsu_load != 0 (an epilogue's share of the shared-memory port). */
int dsdf_selftest_mma_rate(int n, int iters, int lsu_load, unsigned long long* out_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DSDF_B200_H */
