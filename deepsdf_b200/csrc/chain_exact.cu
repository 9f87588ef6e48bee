// High-precision eval forward of the 8x512-class decoder on the tensor cores ("split-fp16" mode): the north star's
// "1e-5 (tf32 mode)" bound, met with 16-bit MMAs (kind::f16, fp32 accumulate) instead of the fp32 CUDA-core path.
//
// Every activation x, every weight w and every aux operand is carried as a PAIR of fp16 numbers, x = x_hi + x_lo
// (x_lo = fp16(x - x_hi): together ~22 mantissa bits; a first build with bf16 pairs, ~16 bits, missed the 1e-5 bound by 4x at
// natural output scale, and bf16-pair aux operands alone still left 1.4e-5.  fp16's range is ample for weight-normed
// activations, which are clamped to 6e4).  The activation tile of a CTA keeps the chain kernel's geometry, [128 rows x 512]
// 16-bit in shared memory, but its 128 MMA rows are 64 points TWICE: rows 0..63 hold x_hi of the CTA's 64 points, rows
// 64..127 hold x_lo of the same points.  Each layer streams its weights twice, W_hi then W_lo, into the same accumulators,
// so that TMEM lane r ends up with x_hi . (W_hi + W_lo) and lane r + 64 with x_lo . (W_hi + W_lo); the epilogue adds the two
// lanes (exchanged through shared memory between the two warps that own them), which is (x_hi + x_lo) . (W_hi + W_lo) with
// fp32 accumulation - all four partial products, nothing dropped.  Layer bias, per-scene latent part and xyz columns enter
// through the aux step of chain_fwd.cu's design, here with fp16 (hi, lo) operands of their own (exact_scene_aux_kernel and the
// constant aux rows behind the weights in the exact pack); the x_lo rows of the aux A block are zero, so they are added once.
//
// Cost: two weight passes for half as many points per tile = 4x the tensor work of the bf16 chain per point - still
// ~20x the throughput of the fp32 CUDA-core path (csrc/generic.cu), which stays the bit-level reference.  Eval mode,
// scene mode (one latent code per 128-point tile; the grid evaluation of deep_sdf/mesh.py:37-73 and decode_sdf with one
// code, deep_sdf/utils.py:54-65).  Structure (CTA pairs, cta_group::2 MMAs, TMA weight ring, per-quarter hand-over) as in
// chain_fwd.cu; replaces Decoder.forward (deep_sdf/networks/deep_sdf_decoder.py:76-111) where 1e-5 is asked for.
#include <cuda_fp16.h>

#include "common.cuh"
#include "fused.cuh"
#include "fused_host.h"
#include "host_util.h"

namespace dsdf {

namespace {

constexpr int X_EPI_WARPS = 16;
constexpr int X_EPI_THREADS = X_EPI_WARPS * 32;
constexpr int X_THREADS = (X_EPI_WARPS + 4) * 32;      // + TMA warp + MMA warp + two idle warps (whole service warpgroup)
constexpr int X_EPI_REGS = 112, X_SVC_REGS = 24;
constexpr int X_STAGES = 4;
constexpr int X_STAGE_BYTES = 128 * 128;               // this CTA's half ([128 out rows][64 K] bf16) of a weight block
constexpr int X_ROWS = 64;                             // points per CTA tile (MMA rows: 64 hi + 64 lo)
constexpr int X_OFF_B = A_BYTES;
constexpr int X_OFF_AUX = X_OFF_B + X_STAGES * X_STAGE_BYTES;    // aux A block: 2 k-cores x [128 rows][8 bf16]
constexpr int X_OFF_WLAST = X_OFF_AUX + 2 * AUX_CORE_BYTES;      // f32 [512]
constexpr int X_OFF_YPART = X_OFF_WLAST + FH * 4;                // f32 [4][128]
constexpr int X_OFF_XCHG = X_OFF_YPART + 4 * TILE_M * 4;         // f32 [8 warp pairs][16][32]: lane-pair exchange
constexpr int X_SMEM = X_OFF_XCHG + 8 * 16 * 32 * 4 + 1024;

struct ExactArgs {
  int32_t num_hidden;
  int32_t n_tiles;                          // 128-point tiles (one per CTA pair; each CTA takes 64 of the points)
  int64_t n_rows;
  int32_t lo_row0;                          // first row of the W_lo blocks in the exact pack (= fwd_rows)
  ChainLayer prog[DSDF_MAX_LAYERS];
  const float* w_last;                      // [512] fp32
  const float* b_last;                      // [1]
  const int32_t* tile_scene;                // [n_tiles]
  const float* xyz;
  int32_t xyz_stride;
  int32_t use_tanh;
  float* sdf_out;                           // [n_rows]
};

// fold g * v / ||v|| (weight norm, deep_sdf_decoder.py:50-55) and split every hidden-activation column of the hidden
// layers into (hi, lo) fp16: exact pack = 16-bit [2][fwd_rows][512], rows arranged like the forward pack (fused_host.cu)
struct ExactPackArgs {
  dsdf_net_desc desc;
  dsdf_params params;
  FusedLayout lo;
  __nv_bfloat16* out;      // 16-bit container: [2][fwd_rows][512] fp16 weights, then [num_hidden][512][16] fp16 layer-constant aux rows
};

__device__ __forceinline__ void split_f16(float x, __half& hi, __half& lo) {
  hi = __float2half_rn(x);
  lo = __float2half_rn(x - __half2float(hi));
}

__global__ void exact_pack_kernel(const __grid_constant__ ExactPackArgs a) {
  const int l = blockIdx.y + 1, o = blockIdx.x;       // hidden layers 1 .. nh-1 (lin0 has no hidden-activation columns)
  if (o >= a.desc.out_dim[l]) return;
  const int in = a.desc.in_dim[l];
  const float* vrow = a.params.v[l] + size_t(o) * in;
  __shared__ float red[4];
  float ss = 0.f;
  for (int k = threadIdx.x; k < in; k += blockDim.x) { const float t = vrow[k]; ss += t * t; }
  for (int off = 16; off > 0; off >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, off);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  const float tot = (red[0] + red[1]) + (red[2] + red[3]);       // same order as pack_weights_kernel
  const float scale = a.params.g[l] != nullptr ? a.params.g[l][o] * (1.0f / sqrtf(tot)) : 1.0f;
  if (threadIdx.x == 0) {
    // aux B row of this output feature: the layer bias (hi, lo), once for each CTA of a pair (fused.cuh column layout)
    __half* ar = reinterpret_cast<__half*>(a.out + size_t(2) * a.lo.fwd_rows * FH) + (size_t(l) * FH + o) * AUX_SCENE_W;
    __half bh, bl;
    split_f16(a.params.b[l][o], bh, bl);
    ar[AUX_COL_ONE0] = bh; ar[AUX_COL_ONE0 + 1] = bl;
    ar[AUX_COL_ONE1] = bh; ar[AUX_COL_ONE1 + 1] = bl;
  }
  __nv_bfloat16* hi_row = a.out + size_t(a.lo.fwd_row0[l] + o) * FH;
  __nv_bfloat16* lo_row = a.out + size_t(a.lo.fwd_rows + a.lo.fwd_row0[l] + o) * FH;
  for (int k = threadIdx.x; k < a.lo.xw[l]; k += blockDim.x) {
    __half hi, lo2;
    split_f16(vrow[k] * scale, hi, lo2);
    hi_row[k] = __ushort_as_bfloat16(__half_as_ushort(hi));      // the pack is a 16-bit container; the MMAs read it as fp16
    lo_row[k] = __ushort_as_bfloat16(__half_as_ushort(lo2));
  }
}

// per-scene aux rows in fp16 (the role of scene_bias_kernel, fused_host.cu, at this mode's precision): for slot 0 (lin0) and
// slot 1 (the latent_in layer) and every output feature the scene's constant part (layer bias + latent columns . z, fp32) as
// (hi, lo) and the xyz columns as (w_hi, w_hi, w_lo) triples, in the K-column order of the aux step.
// grid (n_scenes, 2, 4), block 128; dynamic shared memory: L floats.
__global__ void exact_scene_aux_kernel(const float* __restrict__ lat_table, const int64_t* __restrict__ scene_ids, int L,
                                       const float* __restrict__ wlatT, const float* __restrict__ bias,
                                       const float* __restrict__ wxyz, int lat_layer, __half* __restrict__ out) {
  extern __shared__ float s_z[];
  const int s = blockIdx.x, slot = blockIdx.y, o = blockIdx.z * 128 + threadIdx.x;
  const float* z = lat_table + size_t(scene_ids[s]) * L;
  for (int d = threadIdx.x; d < L; d += blockDim.x) s_z[d] = z[d];
  __syncthreads();
  __align__(16) __half row[AUX_SCENE_W];
#pragma unroll
  for (int i = 0; i < AUX_SCENE_W; ++i) row[i] = __ushort_as_half(0);
  const int layer = slot == 0 ? 0 : lat_layer;
  if (layer >= 0) {
    float acc = 0.f;
    for (int d = 0; d < L; ++d) acc += wlatT[(size_t(slot) * L + d) * FH + o] * s_z[d];
    acc += bias[layer * FH + o];
    __half h, l2;
    split_f16(acc, h, l2);
    row[AUX_COL_ONE0] = h; row[AUX_COL_ONE0 + 1] = l2;
    row[AUX_COL_ONE1] = h; row[AUX_COL_ONE1 + 1] = l2;
#pragma unroll
    for (int dx = 0; dx < 3; ++dx) {
      split_f16(wxyz[(size_t(slot) * 3 + dx) * FH + o], h, l2);
      row[AUX_COL_XYZ + 3 * dx] = h; row[AUX_COL_XYZ + 3 * dx + 1] = h; row[AUX_COL_XYZ + 3 * dx + 2] = l2;
    }
  }
  uint4* dst = reinterpret_cast<uint4*>(out + ((size_t(s) * 2 + slot) * FH + o) * AUX_SCENE_W);
  const uint4* src = reinterpret_cast<const uint4*>(row);
  dst[0] = src[0];
  dst[1] = src[1];
}

__device__ __forceinline__ void xepi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(X_EPI_THREADS) : "memory"); }
// the two warps that own TMEM lanes r and r + 64 of the same 32-column slice (named barriers 2..9)
__device__ __forceinline__ void pair_bar(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }

struct XTri { uint32_t h, l; };
__device__ __forceinline__ XTri xsplit3(float x) {
  __half hi, lo;
  split_f16(x, hi, lo);
  return XTri{uint32_t(__half_as_ushort(hi)), uint32_t(__half_as_ushort(lo))};
}

}  // namespace

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(X_THREADS, 1)
chain_exact_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmAux,
                   const __grid_constant__ CUtensorMap tmScene, const __grid_constant__ ExactArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + X_OFF_B;
  const uint32_t smem_a = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint64_t bar_full[X_STAGES], bar_empty[X_STAGES], bar_tmem_full[2], bar_free_q0, bar_free_lo, bar_a_ready[4];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < X_STAGES; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
    mbar_init(&bar_tmem_full[0], 1); mbar_init(&bar_tmem_full[1], 1);
    mbar_init(&bar_free_q0, 1);
    mbar_init(&bar_free_lo, 1);
    for (int i = 0; i < 4; ++i) mbar_init(&bar_a_ready[i], 2 * X_EPI_WARPS);
    fence_barrier_init();
  }
  if (warp == X_EPI_WARPS + 1) tmem_alloc_pair<512>(&tmem_slot);
  if (warp == X_EPI_WARPS && lane == 0) { tma_prefetch_desc(&tmW); tma_prefetch_desc(&tmAux); tma_prefetch_desc(&tmScene); }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int nh = args.num_hidden;
  const int n_pairs = args.n_tiles;                  // one 128-point tile per CTA pair
  const int pair0 = blockIdx.x >> 1, pair_stride = gridDim.x >> 1;

  if (warp == X_EPI_WARPS) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(X_SVC_REGS));
    // ================================================================ TMA producer: per layer and accumulator half the aux
    // block, then this CTA's 128 rows of every K chunk of W_hi, then of W_lo
    const uint32_t full_leader = mapa_u32(smem_u32(&bar_full[0]), 0);
    const uint32_t full0 = smem_u32(&bar_full[0]), empty0 = smem_u32(&bar_empty[0]), sB_u = smem_u32(sB);
    uint32_t stage = 0, phase = 0;
    const uint64_t pol_w = l2_policy_evict_last();
    for (int tp = pair0; tp < n_pairs; tp += pair_stride) {
      const int scene = __ldg(args.tile_scene + tp);
      for (int l = 0; l < nh; ++l) {
        const int n_halves = args.prog[l].n_halves, k_chunks = args.prog[l].k_chunks;
        const int row_base = args.prog[l].w_row0 + int(rank) * 128;
        const int slot = args.prog[l].bias_slot;
        for (int h = 0; h < n_halves; ++h) {
          mbar_wait_a(empty0 + stage * 8, phase ^ 1);
          if (elect_one()) {
            if (leader) mbar_arrive_expect_tx_a(full0 + stage * 8, 4u * AUX_CORE_BYTES);
            const uint32_t dst = sB_u + stage * X_STAGE_BYTES, bar = full_leader + stage * 8;
            const int r = h * HALF_N + int(rank) * 128;
            if (slot >= 0) {
              tma_load_2d_pair_hint_a(dst, &tmScene, bar, 0, (scene * 2 + slot) * FH + r, pol_w);
              tma_load_2d_pair_hint_a(dst + AUX_CORE_BYTES, &tmScene, bar, 8, (scene * 2 + slot) * FH + r, pol_w);
            } else {
              tma_load_2d_pair_hint_a(dst, &tmAux, bar, 0, l * FH + r, pol_w);
              tma_load_2d_pair_hint_a(dst + AUX_CORE_BYTES, &tmAux, bar, 8, l * FH + r, pol_w);      // [nh x 512][16] fp16 rows
            }
          }
          __syncwarp();
          if (++stage == X_STAGES) { stage = 0; phase ^= 1; }
          for (int pass = 0; pass < 2; ++pass) {
            const int rows = row_base + h * HALF_N + pass * args.lo_row0;
            for (int kc = 0; kc < k_chunks; ++kc) {
              mbar_wait_a(empty0 + stage * 8, phase ^ 1);
              if (elect_one()) {
                if (leader) mbar_arrive_expect_tx_a(full0 + stage * 8, 2 * X_STAGE_BYTES);
                tma_load_2d_pair_hint_a(sB_u + stage * X_STAGE_BYTES, &tmW, full_leader + stage * 8, kc * 64, rows, pol_w);
              }
              __syncwarp();
              if (++stage == X_STAGES) { stage = 0; phase ^= 1; }
            }
          }
        }
      }
    }
  } else if (warp == X_EPI_WARPS + 1) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(X_SVC_REGS));
    // ================================================================ MMA issuer (leader CTA); the whole warp walks the loop
    if (leader) {
      const uint32_t idesc = umma_idesc_bf16(2 * TILE_M, HALF_N, 0, 0) & ~((1u << 7) | (1u << 10));   // a / b format F16
      const uint64_t desc_hi = umma_desc_hi_sw128(16, 1024);
      const uint64_t aux_hi = umma_desc_hi_noswz(AUX_CORE_BYTES, 128);
      const uint64_t a_base = desc_hi | uint64_t((smem_u32(sA) >> 4) & 0x3FFF);
      const uint64_t b_base = desc_hi | uint64_t((smem_u32(sB) >> 4) & 0x3FFF);
      const uint64_t a_aux = aux_hi | uint64_t(((smem_a + X_OFF_AUX) >> 4) & 0x3FFF);
      const uint64_t b_aux = aux_hi | uint64_t((smem_u32(sB) >> 4) & 0x3FFF);
      const uint32_t full0 = smem_u32(&bar_full[0]), empty0 = smem_u32(&bar_empty[0]);
      const uint32_t tfull0 = smem_u32(&bar_tmem_full[0]), fq0 = smem_u32(&bar_free_q0), flo = smem_u32(&bar_free_lo);
      const uint32_t ready0 = smem_u32(&bar_a_ready[0]);
      uint32_t stage = 0, phase = 0, a_cons = 0;
      for (int tp = pair0; tp < n_pairs; tp += pair_stride) {
        for (int l = 0; l < nh; ++l) {
          const int n_halves = args.prog[l].n_halves, k_chunks = args.prog[l].k_chunks;
          const uint32_t apar = a_cons & 1; ++a_cons;
          // the A tile's chunks 0..1 / 2..3 may be overwritten once the LAST half's W_lo pass has read them
          const int kc_lo_last = k_chunks < 4 ? k_chunks - 1 : 3;
          const int kc_q0_last = k_chunks < 2 ? k_chunks - 1 : 1;
          uint32_t waited = 0;
          for (int h = 0; h < n_halves; ++h) {
            const bool last_half = (h == n_halves - 1);
            const int n_steps = 1 + 2 * k_chunks;
            for (int i = 0; i < n_steps; ++i) {
              const int kc = i == 0 ? -1 : (i - 1) % k_chunks;       // -1: the aux step
              const bool lo_pass = i > k_chunks;
              uint32_t needm = (kc < 0) ? (3u << (2 * h)) : (1u << (kc >> 1));
              needm &= ~waited;
              while (needm) {
                const int qd = __ffs(int(needm)) - 1;
                mbar_wait_a(ready0 + qd * 8, apar);
                needm &= needm - 1;
                waited |= 1u << qd;
              }
              mbar_wait_a(full0 + stage * 8, phase);
              tc_fence_after();
              const uint64_t ad = a_base + uint64_t(kc * (A_CHUNK_BYTES >> 4));
              const uint64_t bd = b_base + uint64_t(stage * (X_STAGE_BYTES >> 4));
              const uint64_t bx = b_aux + uint64_t(stage * (X_STAGE_BYTES >> 4));
              const uint32_t d = tmem_base + h * HALF_N;
              const bool final_step = (i == n_steps - 1);
              if (elect_one()) {
                if (kc < 0) {
                  umma_bf16_ss_pair(d, a_aux, bx, idesc, 0u);
                } else {
                  umma_bf16_ss_pair(d, ad, bd, idesc, 1u);
                  umma_bf16_ss_pair(d, ad + 2, bd + 2, idesc, 1u);
                  umma_bf16_ss_pair(d, ad + 4, bd + 4, idesc, 1u);
                  umma_bf16_ss_pair(d, ad + 6, bd + 6, idesc, 1u);
                }
                umma_commit_pair_a(empty0 + stage * 8, 0b11);
                const bool last_reader = last_half && (lo_pass || k_chunks == 0);
                if (last_reader && kc == kc_q0_last) umma_commit_pair_a(fq0, 0b11);
                if (last_reader && kc == kc_lo_last) umma_commit_pair_a(flo, 0b11);
                if (final_step) {
                  umma_commit_pair_a(tfull0 + h * 8, 0b11);
                  if (n_halves == 1) umma_commit_pair_a(tfull0 + 8, 0b11);
                }
              }
              __syncwarp();
              if (++stage == X_STAGES) { stage = 0; phase ^= 1; }
            }
          }
          uint32_t rest = 0xFu & ~waited;
          while (rest) {
            const int qd = __ffs(int(rest)) - 1;
            mbar_wait_a(ready0 + qd * 8, apar);
            rest &= rest - 1;
          }
        }
      }
    }
  } else if (warp >= X_EPI_WARPS + 2) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(X_SVC_REGS));      // idle warps: complete the service warpgroup
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(X_EPI_REGS));
    // ================================================================ epilogue warps (both CTAs)
    const int q = warp & 3;                 // TMEM lane quadrant: 0, 1 = the x_hi rows, 2, 3 = the x_lo rows of the same points
    const int j = warp >> 2;                // 32-column slice of every 128-column quarter
    const int tid = threadIdx.x;            // 0..511
    const uint32_t rt = q * 32 + lane;      // TMEM lane == MMA row
    const uint32_t rp = rt & 63u;           // point within this CTA's 64
    const bool upper = q >= 2;
    const uint32_t tmem_row = tmem_base + (uint32_t(q * 32) << 16);
    const uint32_t ready_leader = mapa_u32(smem_u32(&bar_a_ready[0]), 0);
    sts_f32(smem_a + X_OFF_WLAST + uint32_t(tid) * 4u, __ldg(args.w_last + tid));
    const uint32_t swl = smem_a + X_OFF_WLAST;
    // exchange slots of this lane pair: value k of the 16-column pass at + k * 128
    const int pair_id = j * 2 + (q & 1);
    const uint32_t xslot = smem_a + X_OFF_XCHG + uint32_t(pair_id) * 2048u + uint32_t(lane) * 4u;
    // A-tile line of point rp (x_hi row; the x_lo row is 64 rows = 8192 bytes further) and the 16-byte pieces of this slice
    const uint32_t st_row = smem_a + rp * 128u + uint32_t(j >> 1) * A_CHUNK_BYTES;
    const uint32_t sw = (rp & 7u) << 4;
    const uint32_t aux_dst = smem_a + X_OFF_AUX + uint32_t(j) * AUX_CORE_BYTES + rt * 16u;
    uint32_t mma_iter = 0;
    const uint32_t one = 0x3C00u;           // fp16 1.0
    const uint32_t one0 = leader ? one : 0u, one1 = leader ? 0u : one;

    for (int tp = pair0; tp < n_pairs; tp += pair_stride) {
      const int64_t grow = int64_t(tp) * TILE_M + int64_t(rank) * X_ROWS + rp;
      const bool valid = grow < args.n_rows;
      // ---- aux A block of this tile (every MMA of the previous tile has completed: its last epilogue waited for tmem_full):
      //      x_hi rows carry the bias switch and the xyz triples, x_lo rows are zero (the bias must be added once)
      if (j < 2) {
        uint32_t w0 = 0, w1 = 0, w2 = 0, w3 = 0;
        if (!upper) {
          float px = 0.f, py = 0.f, pz = 0.f;
          if (valid) {
            const float* pt = args.xyz + grow * args.xyz_stride;
            px = __ldg(pt); py = __ldg(pt + 1); pz = __ldg(pt + 2);
          }
          if (j == 0) {
            const XTri x = xsplit3(px), y = xsplit3(py);
            w0 = one0 | (one0 << 16); w1 = x.h | (x.l << 16); w2 = x.h | (y.h << 16); w3 = y.l | (y.h << 16);
          } else {
            const XTri z = xsplit3(pz);
            w0 = z.h | (z.l << 16); w1 = z.h | (one1 << 16); w2 = one1; w3 = 0u;
          }
        }
        sts_u32x4(aux_dst, w0, w1, w2, w3);
        fence_proxy_async_smem();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
#pragma unroll
        for (int Q = 0; Q < 4; ++Q) mbar_arrive_remote(ready_leader + 8u * uint32_t(Q));
      }
      float ydot = 0.0f;

      for (int l = 0; l < nh; ++l) {
        const ChainLayer& L = args.prog[l];
        const bool last = (l == nh - 1);
        const uint32_t par = mma_iter & 1;
        ++mma_iter;

        // One 128-column quarter: two 16-column passes.  In pass 0 the x_lo lane ships its partial sums to the x_hi lane, in
        // pass 1 the other way round, so both warps of a pair post-process 16 of the 32 columns (and a thread only ever
        // overwrites exchange slots it has finished reading itself).
        auto quarter = [&](int Q, uint64_t* free_bar) {
          const uint32_t c0 = uint32_t(Q) * 128u + uint32_t(j) * 32u;
          const uint32_t qoff = uint32_t(Q) * 2u * A_CHUNK_BYTES;
          uint32_t hi[8], lo[8];
#pragma unroll
          for (int p = 0; p < 2; ++p) {
            uint32_t r[16];
            tmem_ld16(tmem_row + c0 + 16u * p, r);
            tmem_ld_wait();
            const bool finisher = (p == 0) ? !upper : upper;
            if (!finisher) {
#pragma unroll
              for (int k = 0; k < 16; ++k) sts_f32(xslot + uint32_t(k) * 128u, __uint_as_float(r[k]));
            }
            pair_bar(2 + pair_id);
            if (finisher) {
              float v[16];
#pragma unroll
              for (int k = 0; k < 16; ++k) v[k] = fmaxf(__uint_as_float(r[k]) + lds_f32(xslot + uint32_t(k) * 128u), 0.0f);
              if (last) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const float4 w = lds_f32x4(swl + (c0 + 16u * p) * 4u + 16u * k);
                  ydot += v[4 * k + 0] * w.x + v[4 * k + 1] * w.y + v[4 * k + 2] * w.z + v[4 * k + 3] * w.w;
                }
              } else {
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                  const float a0v = fminf(v[2 * k], 60000.0f), a1v = fminf(v[2 * k + 1], 60000.0f);
                  const __half2 h = __floats2half2_rn(a0v, a1v);
                  const float2 hf = __half22float2(h);
                  const __half2 lw = __floats2half2_rn(a0v - hf.x, a1v - hf.y);
                  hi[k] = *reinterpret_cast<const uint32_t*>(&h);
                  lo[k] = *reinterpret_cast<const uint32_t*>(&lw);
                }
              }
            }
          }
          if (!last) {
            // the stores wait for the tensor cores to be done with the old chunks of this quarter
            if (free_bar != nullptr) mbar_wait(free_bar, par);
            // this thread finished pass (upper ? 1 : 0): columns c0 + 16 p .. + 15 = pieces 2 p, 2 p + 1 of its slice
            const uint32_t piece = (uint32_t(j & 1) << 2) + (upper ? 2u : 0u);
            const uint32_t a0 = st_row + qoff + (((piece) << 4) ^ sw), a1 = st_row + qoff + (((piece + 1u) << 4) ^ sw);
            sts_u32x4(a0, hi[0], hi[1], hi[2], hi[3]);
            sts_u32x4(a1, hi[4], hi[5], hi[6], hi[7]);
            sts_u32x4(a0 + 8192u, lo[0], lo[1], lo[2], lo[3]);
            sts_u32x4(a1 + 8192u, lo[4], lo[5], lo[6], lo[7]);
            fence_proxy_async_smem();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_remote(ready_leader + 8u * uint32_t(Q));
          }
        };

        mbar_wait(&bar_tmem_full[0], par); tc_fence_after();
        quarter(0, &bar_free_q0);
        quarter(1, &bar_free_lo);
        mbar_wait(&bar_tmem_full[1], par); tc_fence_after();
        if (L.n_halves == 2) {
          quarter(2, nullptr);
          quarter(3, nullptr);
        } else {
          tc_fence_before();
          __syncwarp();
          if (!last && lane == 0) { mbar_arrive_remote(ready_leader + 16u); mbar_arrive_remote(ready_leader + 24u); }
        }
      }

      // ---- final layer: a point's 512 -> 1 product is spread over 4 slices x 2 lanes; fixed summation order
      sts_f32(smem_a + X_OFF_YPART + (uint32_t(j) * TILE_M + rt) * 4u, ydot);
      xepi_bar();
      if (j == 0 && !upper) {
        const uint32_t yp = smem_a + X_OFF_YPART + rp * 4u;
        float yv = 0.0f;
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) yv += lds_f32(yp + uint32_t(jj) * TILE_M * 4u) + lds_f32(yp + uint32_t(jj) * TILE_M * 4u + 64u * 4u);
        yv += __ldg(args.b_last);
        float t = tanhf(yv);
        if (args.use_tanh) t = tanhf(t);
        if (valid) args.sdf_out[grow] = t;
      }
      xepi_bar();
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == X_EPI_WARPS + 1) tmem_dealloc_pair<512>(tmem_base);
}

}  // namespace dsdf

extern "C" size_t dsdf_exact_pack_bytes(const dsdf_net_desc* desc) {
  using namespace dsdf;
  FusedLayout lo;
  if (desc == nullptr || fused_layout(desc, &lo) != 0) return 0;
  return (size_t(2) * size_t(lo.fwd_rows) * FH + size_t(lo.num_hidden) * FH * AUX_SCENE_W) * sizeof(__nv_bfloat16);
}

extern "C" int dsdf_exact_pack(const dsdf_net_desc* desc, const dsdf_params* params, void* exact_pack, void* stream) {
  using namespace dsdf;
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(desc != nullptr && params != nullptr && exact_pack != nullptr, "dsdf_exact_pack: null argument");
  ExactPackArgs a;
  a.desc = *desc;
  a.params = *params;
  if (int e = fused_layout(desc, &a.lo)) return e;
  a.out = static_cast<__nv_bfloat16*>(exact_pack);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DSDF_CHECK_CUDA(cudaMemsetAsync(exact_pack, 0, dsdf_exact_pack_bytes(desc), st));     // padded rows / columns are zero operands
  if (a.lo.num_hidden > 1) {
    DSDF_LAUNCHED;
    exact_pack_kernel<<<dim3(FH, a.lo.num_hidden - 1), 128, 0, st>>>(a);
  }
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_exact_scene_aux(const dsdf_net_desc* desc, const void* pack, const float* lat_table, const int64_t* scene_ids,
                                    int32_t n_scenes, void* scene_aux, void* stream) {
  using namespace dsdf;
  if (int e = require_sm100()) return e;
  FusedLayout lo;
  if (int e = fused_layout(desc, &lo)) return e;
  DSDF_REQUIRE(pack && lat_table && scene_ids && scene_aux && n_scenes > 0, "dsdf_exact_scene_aux: bad arguments");
  DSDF_REQUIRE((reinterpret_cast<uintptr_t>(scene_aux) & 15) == 0, "scene_aux must be 16-byte aligned");
  const uint8_t* p = static_cast<const uint8_t*>(pack);
  const int L = desc->latent_size;
  DSDF_LAUNCHED;
  exact_scene_aux_kernel<<<dim3(n_scenes, 2, FH / 128), 128, size_t(L) * sizeof(float), static_cast<cudaStream_t>(stream)>>>(
      lat_table, scene_ids, L, reinterpret_cast<const float*>(p + lo.off_wlatT), reinterpret_cast<const float*>(p + lo.off_bias),
      reinterpret_cast<const float*>(p + lo.off_wxyz), lo.lat_layer, static_cast<__half*>(scene_aux));
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_fused_forward_exact(const dsdf_net_desc* desc, const void* pack, const void* exact_pack, const void* scene_aux,
                                        const int32_t* tile_scene, const float* xyz, int32_t xyz_stride, int64_t n_rows,
                                        const float* b_last, float* sdf_out, void* stream) {
  using namespace dsdf;
  if (int e = require_sm100()) return e;
  FusedLayout lo;
  if (int e = fused_layout(desc, &lo)) return e;
  DSDF_REQUIRE(pack && exact_pack && scene_aux && tile_scene && xyz && b_last && sdf_out, "dsdf_fused_forward_exact: null argument");
  DSDF_REQUIRE(n_rows > 0 && xyz_stride >= 3, "dsdf_fused_forward_exact: n_rows must be positive and xyz_stride >= 3");
  DSDF_REQUIRE((reinterpret_cast<uintptr_t>(scene_aux) & 15) == 0 && (reinterpret_cast<uintptr_t>(exact_pack) & 15) == 0,
               "scene_aux and exact_pack must be 16-byte aligned");
  const uint8_t* p = static_cast<const uint8_t*>(pack);
  ExactArgs a;
  memset(&a, 0, sizeof(a));
  a.num_hidden = lo.num_hidden;
  a.n_rows = n_rows;
  a.n_tiles = int((n_rows + TILE_M - 1) / TILE_M);
  a.lo_row0 = lo.fwd_rows;
  for (int l = 0; l < lo.num_hidden; ++l) {
    ChainLayer& F = a.prog[l];
    F.has_mma = 1;
    F.w_row0 = lo.fwd_row0[l];
    F.n_halves = lo.fwd_halves[l];
    F.k_chunks = lo.fwd_kchunks[l];
    F.bias_slot = (l == 0) ? 0 : (l == lo.lat_layer ? 1 : -1);
    F.has_xyz = (l == 0 || l == lo.lat_layer);
    F.out_dim = desc->out_dim[l];
  }
  a.w_last = reinterpret_cast<const float*>(p + lo.off_wlast);
  a.b_last = b_last;
  a.tile_scene = tile_scene;
  a.xyz = xyz;
  a.xyz_stride = xyz_stride;
  a.use_tanh = desc->use_tanh;
  a.sdf_out = sdf_out;
  CUtensorMap tmW, tmAux, tmScene;
  if (int e = make_tmap_bf16_2d(&tmW, exact_pack, FH, uint64_t(2) * lo.fwd_rows, FH * 2, 64, 128)) return e;
  const __nv_bfloat16* xaux = static_cast<const __nv_bfloat16*>(exact_pack) + size_t(2) * lo.fwd_rows * FH;
  if (int e = make_tmap_bf16_2d_noswz(&tmAux, xaux, AUX_SCENE_W, uint64_t(lo.num_hidden) * FH, AUX_SCENE_W * 2, 8, 128)) return e;
  if (int e = make_tmap_bf16_2d_noswz(&tmScene, scene_aux, AUX_SCENE_W, uint64_t(1) << 30, AUX_SCENE_W * 2, 8, 128)) return e;
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(chain_exact_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, X_SMEM));
  int grid_pairs = sm_count() / 2;
  if (grid_pairs > a.n_tiles) grid_pairs = a.n_tiles;
  if (grid_pairs < 1) grid_pairs = 1;
  DSDF_LAUNCHED;
  chain_exact_kernel<<<2 * grid_pairs, X_THREADS, X_SMEM, static_cast<cudaStream_t>(stream)>>>(tmW, tmAux, tmScene, a);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}
