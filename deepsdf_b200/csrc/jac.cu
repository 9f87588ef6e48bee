// Input gradient of the fused decoder (sm_100a: tcgen05 + TMEM + TMA), the create_mesh sample grid, and the finish
// kernel of the frozen-decoder latent optimisation.
//
// input_grad_kernel:  G[r, :] = dY_0[r, :] . J_0 + dY_s[r, :] . J_s
//   dY_0 / dY_s   the bf16 pre-activation gradients of lin0 and of the latent_in layer that chain_dgrad_kernel stashed
//   J_0 / J_s     the [512 x (L+3)] latent and xyz columns of those layers' effective weights
// i.e. one GEMM  [N x 1024] . [1024 x (L+3)]  per 128-row tile: A tiles (dY) and the B operand (J^T, K-major, packed by
// pack_weights_kernel) are TMA-streamed through a 4-stage mbarrier ring, a single thread issues tcgen05.mma
// (M128 x N<=256 x K16, plus an N16 tail when L+3 > 256), the fp32 accumulators live in TMEM, and four epilogue warps
// transpose 32x32 blocks through shared memory so that every row of G is written in 128-byte segments.
// Replaces the L x n_cpt torch.autograd.functional.jvp re-evaluations of deep_sdf/mesh.py:405-422: one pass gives
// every row's full gradient, the consumers contract it with their tangents.
#include "common.cuh"
#include "fused.cuh"
#include "fused_host.h"
#include "host_util.h"

namespace dsdf {

namespace {

constexpr int J_MAX_STAGES = 8;                          // ring depth is chosen at run time from the stage size
constexpr int J_A_BYTES = TILE_M * 128;                 // [128 rows][64 K] bf16, SW128
constexpr int J_B_ROWS_MAX = 272;
constexpr int J_B_BYTES = J_B_ROWS_MAX * 128;           // [<=272 output columns][64 K] bf16, SW128
constexpr int J_RING_BYTES = 3 * (J_A_BYTES + J_B_BYTES);   // 153 600: 3 stages at L = 256, 8 stages of 18 KB at L <= 13
constexpr int J_OFF_T = J_RING_BYTES;
constexpr int J_T_FLOATS = 32 * 33;                     // one padded 32x32 transpose block per epilogue warp
constexpr int J_EPI_WARPS = 8;                           // two per TMEM lane quadrant, alternating 32-column chunks
constexpr int J_SMEM = J_OFF_T + J_EPI_WARPS * J_T_FLOATS * 4 + 1024;
constexpr int J_THREADS = (2 + J_EPI_WARPS) * 32;        // TMA warp, MMA warp, 8 epilogue warps

struct JacArgs {
  int32_t n_tiles;
  int32_t k_chunks;        // 8 (no latent_in layer) or 16
  int32_t lat_layer;
  int32_t jac_rows;        // padded output columns (multiple of 16)
  int32_t n_out;           // L + 3
  int32_t stage_bytes;     // 16 KB dY tile + jac_rows x 128 B of the operand, 1024-aligned
  int32_t n_stages;        // min(8, ring / stage_bytes)
  int32_t tmem_base1;      // TMEM column of the odd tiles' accumulators: 256, or 224 when jac_rows = 272 (64 shared columns)
  int64_t n_rows;
  int64_t n_rows_pad;
  float* G;
  int64_t ldg;
};

}  // namespace

__global__ void __launch_bounds__(J_THREADS, 1)
input_grad_kernel(const __grid_constant__ CUtensorMap tmDys, const __grid_constant__ CUtensorMap tmJ,
                  const __grid_constant__ JacArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* s_t = reinterpret_cast<float*>(smem + J_OFF_T);
  __shared__ uint64_t bar_full[J_MAX_STAGES], bar_empty[J_MAX_STAGES], bar_tmem_full, bar_tmem_empty;
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < J_MAX_STAGES; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
    mbar_init(&bar_tmem_full, 1);
    mbar_init(&bar_tmem_empty, J_EPI_WARPS);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(&tmem_slot);
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmDys); tma_prefetch_desc(&tmJ); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int n1 = args.jac_rows > 256 ? 256 : args.jac_rows;
  const int n2 = args.jac_rows - n1;                    // 0 or 16

  if (warp == 0) {
    // ================================================================ TMA producer
    const uint32_t full0 = smem_u32(&bar_full[0]), empty0 = smem_u32(&bar_empty[0]), s_u = smem_u32(smem);
    const uint32_t stage_tx = uint32_t(J_A_BYTES + args.jac_rows * 128);
    uint32_t stage = 0, phase = 0;
    for (int tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x) {
      for (int kc = 0; kc < args.k_chunks; ++kc) {
        mbar_wait_a(empty0 + stage * 8, phase ^ 1);
        if (elect_one()) {
          const uint32_t bar = full0 + stage * 8, base = s_u + stage * args.stage_bytes;
          mbar_arrive_expect_tx_a(bar, stage_tx);
          const int layer = kc < 8 ? 0 : args.lat_layer;
          tma_load_2d_a(base, &tmDys, bar, (kc & 7) * 64, int32_t(int64_t(layer) * args.n_rows_pad + int64_t(tile) * TILE_M));
          for (int rb = 0; rb < args.jac_rows; rb += 16)
            tma_load_2d_a(base + J_A_BYTES + rb * 128, &tmJ, bar, kc * 64, rb);
        }
        __syncwarp();
        if (++stage == uint32_t(args.n_stages)) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================================================================ MMA issuer (whole warp walks, one elected lane issues)
    const uint32_t idesc1 = umma_idesc_bf16(TILE_M, uint32_t(n1), 0, 0);
    const uint32_t idesc2 = umma_idesc_bf16(TILE_M, 16, 0, 0);
    const uint64_t desc_hi = umma_desc_hi_sw128(16, 1024);
    const uint64_t s_base = desc_hi | uint64_t((smem_u32(smem) >> 4) & 0x3FFF);
    const uint32_t full0 = smem_u32(&bar_full[0]), empty0 = smem_u32(&bar_empty[0]);
    const uint32_t tfull = smem_u32(&bar_tmem_full), tempty = smem_u32(&bar_tmem_empty);
    uint32_t stage = 0, phase = 0, it = 0;
    for (int tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x, ++it) {
      // Accumulators alternate between TMEM columns 0 and tmem_base1, so this tile's MMAs overlap the previous tile's
      // epilogue; the two regions may share 32 columns, which that epilogue drains first and then signals here.
      mbar_wait_a(tempty, (it & 1) ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + ((it & 1) ? uint32_t(args.tmem_base1) : 0u);
      for (int kc = 0; kc < args.k_chunks; ++kc) {
        mbar_wait_a(full0 + stage * 8, phase);
        tc_fence_after();
        const uint64_t ad = s_base + uint64_t(stage * uint32_t(args.stage_bytes >> 4));
        const uint64_t bd = ad + uint64_t(J_A_BYTES >> 4);
        const uint64_t bd2 = bd + uint64_t((256 * 128) >> 4);
        if (elect_one()) {
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4) {
            const uint32_t acc = (kc | k4) != 0 ? 1u : 0u;
            umma_bf16_ss(d_tmem, ad + 2 * k4, bd + 2 * k4, idesc1, acc);
            if (n2) umma_bf16_ss(d_tmem + 256, ad + 2 * k4, bd2 + 2 * k4, idesc2, acc);
          }
          umma_commit_a(empty0 + stage * 8);
          if (kc == args.k_chunks - 1) umma_commit_a(tfull);
        }
        __syncwarp();
        if (++stage == uint32_t(args.n_stages)) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ================================================================ epilogue warps: TMEM -> 32x32 transpose -> G
    const int q = warp & 3;                                  // TMEM lane quadrant this warp may read
    float* T = s_t + (warp - 2) * J_T_FLOATS;
    const uint32_t tmem_row = tmem_base + (uint32_t(q * 32) << 16);
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x, ++it) {
      mbar_wait_sleep(&bar_tmem_full, it & 1, 64);
      tc_fence_after();
      const int64_t row0 = int64_t(tile) * TILE_M + q * 32;
      const uint32_t t_row = tmem_row + ((it & 1) ? uint32_t(args.tmem_base1) : 0u);
      // this warp's 32-column chunks: c = hsel, hsel + 2, ... ; even tiles walk them downwards, odd tiles upwards, so the
      // FIRST chunk of every warp is the one that may overlap the other region - after it the next tile's MMAs may start
      const int hsel = (warp - 2) >> 2;
      const int n_chunks = (args.n_out + 31) >> 5;
      const int mine = (n_chunks - hsel + 1) >> 1;                 // chunks hsel, hsel+2, ... < n_chunks
      for (int k = 0; k < mine; ++k) {
        const int c0 = 32 * (hsel + 2 * ((it & 1) ? k : mine - 1 - k));
        uint32_t r[32];
        tmem_ld32(t_row + uint32_t(c0), r);
        tmem_ld_wait();
        if (k == 0) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bar_tmem_empty);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) T[lane * 33 + i] = __uint_as_float(r[i]);
        __syncwarp();
        const bool col_ok = c0 + lane < args.n_out;
#pragma unroll 8
        for (int rr = 0; rr < 32; ++rr) {
          const float v = T[rr * 33 + lane];
          if (col_ok && row0 + rr < args.n_rows) args.G[(row0 + rr) * args.ldg + c0 + lane] = v;
        }
        __syncwarp();
      }
      if (mine == 0) {                                             // a warp without chunks still takes part in the hand-over
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bar_tmem_empty);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// deep_sdf/mesh.py:37-53
__global__ void grid_points_kernel(int N, int64_t i0, int64_t n, float voxel, float* __restrict__ out) {
  const int64_t t = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int64_t idx = i0 + t;
  const int64_t iz = idx % N, iy = (idx / N) % N, ix = (idx / N / N) % N;
  // two roundings like the reference's `samples * voxel_size + voxel_origin` (mesh.py:50-52): no FMA contraction
  out[3 * t + 0] = __fadd_rn(__fmul_rn(float(ix), voxel), -1.0f);
  out[3 * t + 1] = __fadd_rn(__fmul_rn(float(iy), voxel), -1.0f);
  out[3 * t + 2] = __fadd_rn(__fmul_rn(float(iz), voxel), -1.0f);
}

// Frozen-decoder latent optimisation, one block (512 threads) per shape.  Tiles are shape-major with n_tiles / n_scenes
// tiles each.  Fixed summation order everywhere -> bit-identical run to run.
struct LatentFinishArgs {
  const float* tile_part;   // [n_tiles][n_slots][512]
  const float* tile_loss;   // [n_tiles]
  const float* wlatT;       // [2][L][512]
  const float* lat_table;
  const int64_t* scene_ids;
  int32_t n_slots, tiles_per_scene, L, lat_layer;
  float l2reg, inv_n;
  float* lat_grad;          // [n_scenes][L]
  float* shape_loss;        // [n_scenes] or null
};

__global__ void __launch_bounds__(FH) latent_finish_kernel(const __grid_constant__ LatentFinishArgs a) {
  const int s = blockIdx.x, c = threadIdx.x, warp = c >> 5, lane = c & 31;
  __shared__ float ss0[FH], ss1[FH];
  __shared__ float s_red[FH / 32];
  const int t0 = s * a.tiles_per_scene;
  float a0 = 0.f, a1 = 0.f;
  for (int t = 0; t < a.tiles_per_scene; ++t) {
    const float* part = a.tile_part + size_t(t0 + t) * a.n_slots * FH;
    a0 += __ldg(part + c);
    if (a.lat_layer > 0) a1 += __ldg(part + size_t(a.lat_layer) * FH + c);
  }
  ss0[c] = a0; ss1[c] = a1;
  const float* z = a.lat_table + size_t(a.scene_ids[s]) * a.L;
  // mean(z^2) and the data loss of this shape
  float q = 0.f;
  for (int k = c; k < a.L; k += FH) q += z[k] * z[k];
  for (int off = 16; off > 0; off >>= 1) q += __shfl_xor_sync(0xffffffffu, q, off);
  if (lane == 0) s_red[warp] = q;
  __syncthreads();
  if (c == 0 && a.shape_loss != nullptr) {
    float zz = 0.f;
    for (int w = 0; w < FH / 32; ++w) zz += s_red[w];
    float data = 0.f;
    for (int t = 0; t < a.tiles_per_scene; ++t) data += a.tile_loss[t0 + t];
    a.shape_loss[s] = data * a.inv_n + a.l2reg * zz / float(a.L);
  }
  for (int k = warp; k < a.L; k += FH / 32) {
    float acc = 0.f;
    for (int o = lane; o < FH; o += 32) {
      acc += ss0[o] * __ldg(a.wlatT + size_t(k) * FH + o);
      if (a.lat_layer > 0) acc += ss1[o] * __ldg(a.wlatT + (size_t(a.L) + k) * FH + o);
    }
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
    if (lane == 0) a.lat_grad[size_t(s) * a.L + k] = acc + a.l2reg * 2.0f * z[k] / float(a.L);
  }
}

}  // namespace dsdf

using namespace dsdf;

extern "C" int dsdf_fused_input_grad(const dsdf_net_desc* desc, const void* pack, const dsdf_train_bufs* bufs, int64_t n_rows,
                                     float* G, int64_t ldg, void* stream) {
  if (int e = require_sm100()) return e;
  FusedLayout lo;
  if (int e = fused_layout(desc, &lo)) return e;
  DSDF_REQUIRE(n_rows > 0 && n_rows % TILE_M == 0, "input gradient needs n_rows to be a positive multiple of 128");
  DSDF_REQUIRE(bufs && bufs->dys && G, "input gradient needs the dY stash and an output buffer");
  const int n_out = desc->latent_size + 3;
  DSDF_REQUIRE(ldg >= n_out, "ldg must be >= latent_size + 3");
  DSDF_REQUIRE(lo.jac_rows <= J_B_ROWS_MAX, "latent_size too large for the input-gradient kernel");
  JacArgs a;
  memset(&a, 0, sizeof(a));
  a.n_tiles = int(n_rows / TILE_M);
  a.k_chunks = lo.lat_layer > 0 ? 16 : 8;
  a.lat_layer = lo.lat_layer;
  a.jac_rows = lo.jac_rows;
  a.n_out = n_out;
  a.stage_bytes = (J_A_BYTES + lo.jac_rows * 128 + 1023) / 1024 * 1024;
  a.n_stages = J_RING_BYTES / a.stage_bytes < J_MAX_STAGES ? J_RING_BYTES / a.stage_bytes : J_MAX_STAGES;
  const int cols32 = (lo.jac_rows + 31) / 32 * 32;              // the epilogue reads whole 32-column chunks
  DSDF_REQUIRE(cols32 <= 288, "latent_size too large for the input-gradient kernel");
  a.tmem_base1 = cols32 <= 256 ? 256 : 512 - cols32;            // <= 64 shared columns = the first chunk of each warp set
  a.n_rows = n_rows;
  a.n_rows_pad = n_rows;
  a.G = G;
  a.ldg = ldg;
  const uint8_t* p = static_cast<const uint8_t*>(pack);
  CUtensorMap tmDys, tmJ;
  if (int e = make_tmap_bf16_2d(&tmDys, bufs->dys, FH, uint64_t(lo.num_hidden) * n_rows, FH * 2, 64, TILE_M)) return e;
  if (int e = make_tmap_bf16_2d(&tmJ, p + lo.off_wjac, 2 * FH, lo.jac_rows, 2 * FH * 2, 64, 16)) return e;
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(input_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, J_SMEM));
  int grid = sm_count();
  if (grid > a.n_tiles) grid = a.n_tiles;
  DSDF_LAUNCHED;
  input_grad_kernel<<<grid, J_THREADS, J_SMEM, static_cast<cudaStream_t>(stream)>>>(tmDys, tmJ, a);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_grid_points(int32_t N, int64_t i0, int64_t n, float* xyz_out, void* stream) {
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(N >= 2 && i0 >= 0 && n >= 0 && i0 + n <= int64_t(N) * N * N, "grid range out of bounds");
  if (n == 0) return 0;
  const float voxel = float(2.0 / double(N - 1));      // Python double, rounded once when it meets the fp32 tensor
  DSDF_LAUNCHED;
  grid_points_kernel<<<unsigned((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(N, i0, n, voxel, xyz_out);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_fused_latent_finish(const dsdf_net_desc* desc, const void* pack, const dsdf_train_bufs* bufs,
                                        const int32_t* tile_scene, int64_t n_rows, const float* lat_table,
                                        const int64_t* scene_ids, int32_t n_scenes, float l2reg, float inv_n, float* lat_grad,
                                        float* shape_loss, void* stream) {
  if (int e = require_sm100()) return e;
  FusedLayout lo;
  if (int e = fused_layout(desc, &lo)) return e;
  (void)tile_scene;
  DSDF_REQUIRE(n_rows > 0 && n_rows % TILE_M == 0 && n_scenes > 0, "latent finish: bad sizes");
  const int64_t n_tiles = n_rows / TILE_M;
  DSDF_REQUIRE(n_tiles % n_scenes == 0, "latent finish needs the same number of 128-row tiles for every shape (shape-major)");
  LatentFinishArgs a;
  a.tile_part = bufs->tile_part;
  a.tile_loss = bufs->tile_loss;
  a.wlatT = reinterpret_cast<const float*>(static_cast<const uint8_t*>(pack) + lo.off_wlatT);
  a.lat_table = lat_table;
  a.scene_ids = scene_ids;
  a.n_slots = lo.num_hidden + 6;
  a.tiles_per_scene = int(n_tiles / n_scenes);
  a.L = desc->latent_size;
  a.lat_layer = lo.lat_layer;
  a.l2reg = l2reg;
  a.inv_n = inv_n;
  a.lat_grad = lat_grad;
  a.shape_loss = shape_loss;
  DSDF_LAUNCHED;
  latent_finish_kernel<<<n_scenes, FH, 0, static_cast<cudaStream_t>(stream)>>>(a);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}
