// Shared definitions of the fused tcgen05 path (hidden width 512): packed-weight layout, layer programs and
// the kernel argument blocks.  Host (fused_host.cu) and device (fused_fwd.cu, fused_bwd.cu, wgrad.cu) agree on these.
#pragma once
#include <cuda.h>
#include <stdint.h>
#include "../../include/dsdf.h"

namespace dsdf {

constexpr int FH = 512;              // hidden width of the fused path
constexpr int TILE_M = 128;          // rows per CTA tile == TMEM lanes
constexpr int A_CHUNK_BYTES = TILE_M * 128;       // [128 rows][64 bf16], SW128
constexpr int A_BYTES = 8 * A_CHUNK_BYTES;        // 128 KB: one full 512-wide activation tile
constexpr int HALF_N = 256;                       // accumulator half (one UMMA N)
constexpr int STAGE_BYTES = HALF_N * 128;         // weight stage: [256 out rows][64 K] bf16, SW128 (32 KB)
constexpr int NUM_STAGES = 3;
constexpr int EPI_WARPS = 8;
constexpr int CHAIN_THREADS = (EPI_WARPS + 2) * 32;   // + TMA warp + MMA warp
constexpr int CHAIN_SMEM = A_BYTES + NUM_STAGES * STAGE_BYTES + 1024;
constexpr int MASK_WORDS = FH / 32;               // 16 x u32 keep/relu bits per row per hidden layer

// ---- "aux" K columns of the forward chain: bias, xyz columns and (row-latent mode) short latent codes enter every layer
// as one or two extra K = 16 tensor-core steps instead of per-element CUDA-core work in the epilogue.  The A side is a
// [128 rows x 32] bf16 block per tile (written once per tile by the epilogue threads), the B side one [512 x 32] bf16 row
// block per layer (pack) or one [512 x 16] block per (scene, slot) (scene mode: bias + latent part of lin0 / skip layer).
// Values are split hi + lo (bf16 each) so that the products carry ~16 mantissa bits:
//   col  0, 1   A = 1, 1 (rank-0 CTA of the pair) / 0, 0 (rank 1)     B = bias_hi, bias_lo
//   col  2..10  A = (x_hi, x_lo, x_hi), (y..), (z..)                  B = (wx_hi, wx_hi, wx_lo), (wy..), (wz..)
//   col 11, 12  A = 0, 0 (rank 0) / 1, 1 (rank 1)                     B = bias_hi, bias_lo   (the pair shares ONE B operand:
//               the two CTAs' tiles may belong to different scenes, so k-core 0 comes from rank 0's scene and k-core 1 from
//               rank 1's, and each CTA switches on its own copy of the bias with its 'one' columns)
//   col 16+3d.. A = (z_d hi, lo, hi), d < 4 (row-latent mode, L <= 4) B = (w hi, hi, lo) of latent column d
//   col 16+d    A = bf16(z_d), d < 16 (row-latent mode, 4 < L <= 16)   B = bf16(w) of latent column d - one term each: the
//               precision of every other bf16 operand of the chain (codes of the reference's L = 16 experiments)
// Both sides use the SWIZZLE_NONE K-major canonical layout: 8-row x 16-byte core matrices, rows of a k-core contiguous
// (SBO = 128 B), k-cores 2 KB apart (LBO = 2048 B).
constexpr int AUX_W = 32;
constexpr int AUX_SCENE_W = 16;
constexpr int AUX_CORE_BYTES = TILE_M * 16;       // one k-core of 128 rows: 2 KB
constexpr int AUX_COL_ONE0 = 0, AUX_COL_XYZ = 2, AUX_COL_ONE1 = 11, AUX_COL_LAT = 16;
constexpr int AUX_MAX_ROW_LAT = 4;            // hi/lo-split (3-term) per-row codes
constexpr int AUX_MAX_ROW_LAT1 = 16;          // single-term per-row codes

// One hidden layer of a chain (forward: lin{l}; backward: dgrad of lin{l}).
struct ChainLayer {
  int32_t has_mma;      // backward: 0 = layer 0 (no dgrad product needed).  forward: always 1 (every layer has its aux step)
  int32_t w_row0;       // first row of this layer's block in the packed weight tensor
  int32_t n_halves;     // 1 or 2 accumulator halves (N = 256 each)
  int32_t k_chunks;     // 64-wide K chunks read from the A tile (forward layer 0: none, the aux step is all there is)
  int32_t bias_slot;    // forward: -1 = layer-constant aux block; 0/1 = per-scene aux slot (latent folded into the bias)
  int32_t has_xyz;      // forward: the aux block carries xyz columns; backward: emit xyz-weighted column sums
  int32_t dropout;      // dropout after this hidden layer (training)
  int32_t out_dim;      // true width (<= 512)
};

struct FwdArgs {
  int32_t num_hidden;                       // hidden layers (the final 512->1 layer is fused into the last epilogue)
  int32_t n_tiles;
  int64_t n_rows;
  int64_t n_rows_pad;                       // rows per layer region in acts / bitmask (multiple of 128)
  ChainLayer prog[DSDF_MAX_LAYERS];
  int32_t aux_steps[DSDF_MAX_LAYERS];       // K = 16 aux steps of layer l: 1, or 2 where row-latent columns are live
  const float* w_last;                      // [512] fp32 effective last-layer weight (dropout scale of the last hidden layer folded in when training)
  const float* b_last;                      // [1]
  const int32_t* tile_scene;                // [n_tiles] -> scene row of the per-scene aux table (scene mode)
  const float* xyz;                         // row r at xyz + r*xyz_stride (3 floats)
  int32_t xyz_stride;                       // floats between rows: 3 ([N,3]) or 4 ([N,4] xyz+sdf samples)
  int32_t gt_stride;                        // floats between gt values: 1 or 4
  float* sdf_out;                           // [N]
  int32_t use_tanh;
  int32_t training;
  uint32_t drop_thr;
  uint64_t seed;
  uint64_t row0;                            // global row id of local row 0 (dropout stream is keyed by global rows)
  const uint32_t* step_dev;                 // device step counter or null: rows are keyed row0 + (step << 40) + r
  const float* gt;                          // [N] or null
  float clamp;                              // ClampingDistance (<=0: no clamp)
  float inv_n;                              // 1 / N_total (global batch)
  float* dy_out;                            // [N_pad] dL/d(pre-tanh output)
  float* tile_loss;                         // [n_tiles] sum |pred-gt| of the tile's rows
  uint32_t* bitmask;                        // [num_hidden][N_pad][16]
  int32_t store_acts;
  int32_t dropout_on;                       // 0: gates mode without dropout (eval-mode forward that still records relu gates)
  int32_t dy_unit;                          // gt == null: dy_out = d sdf / d(pre-tanh output)  (input-gradient path)
  const int32_t* tile_rows;                 // [n_tiles] valid rows of each tile (ragged scenes padded to 128), or null
  uint32_t philox_rk[16];                   // Philox round keys of `seed` (common.cuh philox_round_keys)
  // row-latent mode (every row brings its own latent code, L <= 16: the microstructure consumers of deep_sdf/mesh.py:457-497
  // with L = 1, 2, and decoder(input[N, L+3]) of the L = 16 experiments)
  const float* row_lat;                     // row r's code at row_lat + r*row_lat_stride, or null (scene mode)
  int32_t row_lat_stride;
  int32_t row_lat_dim;                      // L
  // dense per-row codes (L > 16, e.g. 256): the code is a real tensor-core operand - bf16 K chunks of the activation tile:
  // chunks [0, dense_chunks) for lin0 (the tile is empty there), chunks [4, 4 + dense_chunks) for the latent_in layer
  // (written by the epilogue of the layer before it, whose single accumulator half leaves them free)
  int32_t dense_chunks;                     // 0: codes (if any) ride the aux step
  int32_t dense_layer;                      // the latent_in layer
};

struct BwdArgs {
  int32_t num_hidden;
  int32_t n_tiles;
  int64_t n_rows;
  int64_t n_rows_pad;
  ChainLayer prog[DSDF_MAX_LAYERS];         // prog[l] = dgrad of lin{l}: produces dY_{l-1}; l = num_hidden-1 .. 1
  const float* w_last;                      // [512]
  const float* dy;                          // [N_pad]
  const uint32_t* bitmask;                  // [num_hidden][N_pad][16]
  const float* xyz;                         // row r at xyz + r*xyz_stride
  int32_t xyz_stride;
  float* tile_part;                         // [n_tiles][part_stride] per-tile column sums
  int32_t part_stride;
  int32_t colsum_slot[DSDF_MAX_LAYERS];     // slot (x512 floats) of colsum(dY_l) in tile_part, or -1
  int32_t xyzsum_slot[DSDF_MAX_LAYERS];     // first of 3 slots of sum_r xyz[r,d]*dY_l[r,:], or -1
  int32_t store_dys;                        // bit t: TMA-store the bf16 dY tile of hidden layer t
};

// wgrad work plan.  The row axis of every layer is cut into contiguous segments, one per CLUSTER of four CTAs (the four
// 128-feature output blocks of the layer share every activation box through TMA multicast).  The plan depends only on
// the network and the row count - never on the device - so the split-K summation order, and with it every gradient
// bit, is reproducible.
constexpr int WG_CLUSTERS = 33;      // 33 clusters x 4 CTAs = 132 SMs: a 4-CTA cluster must sit inside one GPC, and the
                                     // B200's GPCs leave SMs over (cudaOccupancyMaxActiveClusters = 33); a constant, so that
                                     // the split-K order never depends on the device
constexpr int WG_MAX_CLUSTERS = 40;
constexpr int WG_MAX_SEGS = 128;
struct WgradSeg {
  int32_t layer;        // hidden layer l: dW_l = dY_l^T . X_{l-1}
  int32_t n;            // input features covered (256 or 512)
  int32_t step_beg;     // 32-row steps [step_beg, step_end)
  int32_t step_end;
  int64_t part_offset;  // float offset of this segment's partial [4 x 128 output features][n]
};
struct WgradPlan {
  int32_t n_segs;
  int32_t n_clusters;                          // clusters launched (WG_CLUSTERS; other values only for experiments)
  int32_t cluster_seg0[WG_MAX_CLUSTERS + 1];   // segments of cluster c: [cluster_seg0[c], cluster_seg0[c+1])
  int32_t layer_seg0[DSDF_MAX_LAYERS + 1];     // segments of layer l:   [layer_seg0[l], layer_seg0[l+1])
  WgradSeg seg[WG_MAX_SEGS];
  int64_t total_floats;
};
struct WgradArgs {
  WgradPlan plan;
  int64_t n_rows_pad;
  float* partials;
};

}  // namespace dsdf
