// Packed-weight layout of the fused path (offsets into the caller-provided pack buffer).
#pragma once
#include <stddef.h>
#include <stdint.h>
#include "../../include/dsdf.h"

namespace dsdf {

struct FusedLayout {
  int32_t num_hidden;
  int32_t lat_layer;                     // index of the latent_in layer, or -1
  int32_t xw[DSDF_MAX_LAYERS];           // width of the hidden-activation part of layer l's input (0 for layer 0)
  int32_t lat_col0[DSDF_MAX_LAYERS];     // first latent column of layer l's input, or -1
  int32_t xyz_col0[DSDF_MAX_LAYERS];     // first xyz column of layer l's input, or -1
  int32_t fwd_row0[DSDF_MAX_LAYERS], fwd_halves[DSDF_MAX_LAYERS], fwd_kchunks[DSDF_MAX_LAYERS];
  int32_t bwd_row0[DSDF_MAX_LAYERS], bwd_halves[DSDF_MAX_LAYERS], bwd_kchunks[DSDF_MAX_LAYERS];
  int32_t fwd_rows, bwd_rows;
  // dense per-row codes (row-latent mode for long codes, e.g. L = 256): lin0's latent columns as a real [512 x Lpad] B block
  // (rows fwd_row0[0] ..: behind all other layers' blocks of the same array) and the latent_in layer's latent columns
  // appended to its block at K offset dense_kx; 0 chunks = the network is outside this mode (see fused_layout)
  int32_t dense_lat_chunks;              // ceil(L / 64), or 0
  int32_t dense_kx;                      // first K column of the latent part in the latent_in layer's block (64 * ceil(xw / 64))
  size_t off_wp;                         // bf16 [fwd_rows][512]   forward B operand  (rows = out features, K = in)
  size_t off_wpt;                        // bf16 [bwd_rows][512]   dgrad B operand    (rows = in features,  K = out)
  // training variants: the inverted-dropout scale 1/(1 - p_eff) of hidden layer l-1 multiplied into the hidden-activation
  // columns of layer l's weights (and into the last layer), so that activations / dY are stored unscaled and neither chain
  // epilogue multiplies.  Equal to the plain offsets when the network has no dropout.
  size_t off_wp_tr, off_wpt_tr, off_wlast_tr;
  size_t off_aux;                        // bf16 [num_hidden][512][AUX_W]  layer-constant aux B blocks (fused.cuh)
  float in_scale[DSDF_MAX_LAYERS + 1];   // training: factor folded into the activation columns of layer l (l = num_hidden: last layer)
  size_t off_bias;                       // f32  [num_hidden][512]
  size_t off_wxyz;                       // f32  [2][3][512]       xyz columns of lin0 / skip layer
  size_t off_wlatT;                      // f32  [2][L][512]       latent columns of lin0 / skip layer, transposed
  size_t off_wlast;                      // f32  [512]             effective last-layer weight
  size_t off_wjac;                       // bf16 [jac_rows][1024]  input-gradient B operand: row c < L latent column c, rows
                                         //                        L..L+2 xyz; K < 512 from lin0, K >= 512 from the latent_in layer
  int32_t jac_rows;                      // L+3 rounded up to a multiple of 16 (<= 272)
  size_t off_weff[DSDF_MAX_LAYERS];      // f32  [out][in]         effective (folded) weights, reference layout
  size_t off_invn[DSDF_MAX_LAYERS];      // f32  [out]             1/||v|| per row (1 for plain layers)
  size_t total_bytes;
};

int fused_unsupported_reason(const dsdf_net_desc* d, const char** why);
int fused_layout(const dsdf_net_desc* d, FusedLayout* lo);

}  // namespace dsdf
