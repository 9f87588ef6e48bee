// Fused DeepSDF decoder backward, data-gradient chain (sm_100a, tcgen05 + TMEM + TMA).
//
// Mirror image of fused_fwd.cu.  Per 128-row tile, for t = last hidden layer .. 0 the kernel produces
//   dY_t = dX_{t+1} (.) relu'/dropout mask of layer t        (dX_{t+1} = dY_{t+1} . W_{t+1}, tensor cores)
// with dY of the last hidden layer seeded by the rank-1 product dL/dy * w_last.  dY tiles live in shared memory as
// bf16 (SW128, the A operand of the next UMMA chain link), W^T streams through a 3-stage TMA ring, accumulators fill
// TMEM, and the 8 epilogue warps apply the 1-bit keep masks written by the forward.  Per tile it also emits
//   * column sums of every dY_t (bias gradients; per-scene sums of dY_0 / dY_skip ARE the latent-code gradient
//     up to a tiny [S x 512] x [512 x L] product, so the latent-gradient reduction is deterministic by construction),
//   * xyz-weighted column sums of dY_0 / dY_skip (gradients of the xyz weight columns),
//   * the bf16 dY tiles (TMA store) that wgrad.cu contracts with the stashed activations.
// The data gradient w.r.t. lin0's input is never formed: training only needs it reduced per scene.
// Replaces the autograd graph of train_deep_sdf.py:533 (data-gradient part).
#include "common.cuh"
#include "fused.cuh"
#include "host_util.h"

namespace dsdf {

namespace {

constexpr uint32_t kBarEpiB = 1;
__device__ __forceinline__ void epi_bar_sync_b() { asm volatile("bar.sync %0, %1;" ::"n"(kBarEpiB), "n"(EPI_WARPS * 32) : "memory"); }

__device__ __forceinline__ void store_group_b(uint8_t* sA, uint32_t row, uint32_t c0, const uint32_t* p) {
  uint8_t* base = sA + (c0 >> 6) * A_CHUNK_BYTES + row * 128u;
  const uint32_t piece0 = (c0 & 63u) >> 3;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    uint32_t piece = (piece0 + t) ^ (row & 7u);
    *reinterpret_cast<uint4*>(base + piece * 16u) = make_uint4(p[4 * t], p[4 * t + 1], p[4 * t + 2], p[4 * t + 3]);
  }
}

// v (32 columns of one row) -> masked, scaled, packed
__device__ __forceinline__ void bwd_group(float (&v)[32], uint32_t bits, float scale, uint32_t (&p)[16]) {
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = ((bits >> i) & 1u) ? v[i] * scale : 0.0f;
#pragma unroll
  for (int i = 0; i < 16; ++i) p[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
}

}  // namespace

__global__ void __launch_bounds__(CHAIN_THREADS, 1)
fused_dgrad_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmDys,
                   const __grid_constant__ BwdArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + A_BYTES;
  __shared__ uint64_t bar_full[NUM_STAGES], bar_empty[NUM_STAGES], bar_tmem_full[2], bar_a_ready[2];
  __shared__ uint32_t tmem_slot;
  __shared__ float s_xyz[3][TILE_M];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NUM_STAGES; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
    mbar_init(&bar_tmem_full[0], 1); mbar_init(&bar_tmem_full[1], 1);
    mbar_init(&bar_a_ready[0], EPI_WARPS); mbar_init(&bar_a_ready[1], EPI_WARPS);
    fence_barrier_init();
  }
  if (warp == EPI_WARPS + 1) tmem_alloc<512>(&tmem_slot);
  if (warp == EPI_WARPS && lane == 0) { tma_prefetch_desc(&tmW); tma_prefetch_desc(&tmDys); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int nh = args.num_hidden;

  if (warp == EPI_WARPS) {
    // ================================================================ TMA producer (W^T)
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x) {
        for (int t = nh - 2; t >= 0; --t) {
          const ChainLayer& L = args.prog[t + 1];
          for (int h = 0; h < L.n_halves; ++h)
            for (int kc = 0; kc < L.k_chunks; ++kc) {
              mbar_wait(&bar_empty[stage], phase ^ 1);
              mbar_arrive_expect_tx(&bar_full[stage], STAGE_BYTES);
              tma_load_2d(sB + stage * STAGE_BYTES, &tmW, &bar_full[stage], kc * 64, L.w_row0 + h * HALF_N);
              if (++stage == NUM_STAGES) { stage = 0; phase ^= 1; }
            }
        }
      }
    }
  } else if (warp == EPI_WARPS + 1) {
    // ================================================================ MMA issuer
    const uint32_t idesc = umma_idesc_bf16(TILE_M, HALF_N, 0, 0);
    const uint64_t desc_hi = umma_desc_hi_sw128(16, 1024);
    const uint32_t sA_u = smem_u32(sA), sB_u = smem_u32(sB);
    int stage = 0; uint32_t phase = 0; uint32_t mma_iter = 0;
    for (int tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x) {
      for (int t = nh - 2; t >= 0; --t) {
        const ChainLayer& L = args.prog[t + 1];
        const uint32_t par = mma_iter & 1; ++mma_iter;
        bool waited1 = false;
        for (int h = 0; h < L.n_halves; ++h) {
          for (int kc = 0; kc < L.k_chunks; ++kc) {
            if (h == 0 && kc == 0) { mbar_wait(&bar_a_ready[0], par); }
            if ((h == 1 || kc >= 4) && !waited1) { mbar_wait(&bar_a_ready[1], par); waited1 = true; }
            mbar_wait(&bar_full[stage], phase);
            tc_fence_after();
            if (lane == 0) {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                uint64_t ad = umma_desc(sA_u + kc * A_CHUNK_BYTES + k * 32, desc_hi);
                uint64_t bd = umma_desc(sB_u + stage * STAGE_BYTES + k * 32, desc_hi);
                umma_bf16_ss(tmem_base + h * HALF_N, ad, bd, idesc, (kc | k) != 0 ? 1u : 0u);
              }
              umma_commit(&bar_empty[stage]);
            }
            __syncwarp();
            if (++stage == NUM_STAGES) { stage = 0; phase ^= 1; }
          }
          if (lane == 0) umma_commit(&bar_tmem_full[h]);
          __syncwarp();
        }
        if (L.n_halves == 1) {
          if (!waited1) mbar_wait(&bar_a_ready[1], par);
          if (lane == 0) umma_commit(&bar_tmem_full[1]);
          __syncwarp();
        }
      }
    }
  } else {
    // ================================================================ epilogue warps
    const int q = warp & 3;
    const int j = warp >> 2;
    const uint32_t rt = q * 32 + lane;
    const uint32_t tmem_row = tmem_base + (uint32_t(q * 32) << 16);
    uint32_t mma_iter = 0;
    bool stores_pending = false;
    for (int tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x) {
      const int64_t lrow = int64_t(tile) * TILE_M + rt;
      const bool valid = lrow < args.n_rows;
      const float dyv = __ldg(args.dy + lrow);     // dy buffer is padded to n_rows_pad and zero beyond n_rows
      if (j == 0) {
        const float* pt = args.xyz + lrow * args.xyz_stride;
        s_xyz[0][rt] = valid ? __ldg(pt + 0) : 0.0f;
        s_xyz[1][rt] = valid ? __ldg(pt + 1) : 0.0f;
        s_xyz[2][rt] = valid ? __ldg(pt + 2) : 0.0f;
      }
      for (int t = nh - 1; t >= 0; --t) {
        const bool has_mma = (t < nh - 1);
        const int n_halves = has_mma ? args.prog[t + 1].n_halves : 2;
        const float scale = args.scale[t];
        const bool produce_a = (t > 0);
        // gate bits: [layer][tile][16 words][128 rows]
        const uint32_t* mrow = args.bitmask + (size_t(t) * (args.n_rows_pad / TILE_M) + tile) * MASK_WORDS * TILE_M + rt;
        uint32_t par = 0;
        if (has_mma) { par = mma_iter & 1; ++mma_iter; }
        uint32_t hold[64];

        // ---- half 0 -> registers
        if (has_mma) { mbar_wait(&bar_tmem_full[0], par); tc_fence_after(); }
        {
          const uint32_t mwords[4] = {__ldg(mrow + (j * 4 + 0) * TILE_M), __ldg(mrow + (j * 4 + 1) * TILE_M),
                                      __ldg(mrow + (j * 4 + 2) * TILE_M), __ldg(mrow + (j * 4 + 3) * TILE_M)};
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t c0 = j * 128 + g * 32;
            float v[32];
            if (has_mma) {
              uint32_t raw[32];
              tmem_ld32(tmem_row + c0, raw);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]);
            } else {
#pragma unroll
              for (int jj = 0; jj < 8; ++jj) {
                float4 w = __ldg(reinterpret_cast<const float4*>(args.w_last + c0) + jj);
                v[4 * jj + 0] = dyv * w.x; v[4 * jj + 1] = dyv * w.y; v[4 * jj + 2] = dyv * w.z; v[4 * jj + 3] = dyv * w.w;
              }
            }
            uint32_t p[16];
            bwd_group(v, mwords[g], scale, p);
#pragma unroll
            for (int i = 0; i < 16; ++i) hold[g * 16 + i] = p[i];
          }
        }
        if (has_mma) { mbar_wait(&bar_tmem_full[1], par); tc_fence_after(); }
        // previous dY tile: its column-sum pass and TMA store must be done before the tile is overwritten
        if (stores_pending && threadIdx.x == 0) tma_store_wait_read0();
        stores_pending = false;
        epi_bar_sync_b();
#pragma unroll
        for (int g = 0; g < 4; ++g) store_group_b(sA, rt, j * 128 + g * 32, &hold[g * 16]);
        fence_proxy_async_smem();
        tc_fence_before();
        __syncwarp();
        if (produce_a && lane == 0) mbar_arrive(&bar_a_ready[0]);

        // ---- half 1
        if (n_halves == 2) {
          const uint32_t mwords[4] = {__ldg(mrow + (8 + j * 4 + 0) * TILE_M), __ldg(mrow + (8 + j * 4 + 1) * TILE_M),
                                      __ldg(mrow + (8 + j * 4 + 2) * TILE_M), __ldg(mrow + (8 + j * 4 + 3) * TILE_M)};
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t c0 = HALF_N + j * 128 + g * 32;
            float v[32];
            if (has_mma) {
              uint32_t raw[32];
              tmem_ld32(tmem_row + c0, raw);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]);
            } else {
#pragma unroll
              for (int jj = 0; jj < 8; ++jj) {
                float4 w = __ldg(reinterpret_cast<const float4*>(args.w_last + c0) + jj);
                v[4 * jj + 0] = dyv * w.x; v[4 * jj + 1] = dyv * w.y; v[4 * jj + 2] = dyv * w.z; v[4 * jj + 3] = dyv * w.w;
              }
            }
            uint32_t p[16];
            bwd_group(v, mwords[g], scale, p);
            store_group_b(sA, rt, c0, p);
          }
          fence_proxy_async_smem();
        }
        tc_fence_before();
        __syncwarp();
        if (produce_a && lane == 0) mbar_arrive(&bar_a_ready[1]);

        // ---- tile of dY_t complete in shared memory: column sums + stash
        epi_bar_sync_b();
        {
          // warp w owns 64-column chunk w; lane owns one bf16x2 word (2 adjacent columns) of each row
          const int chunk = warp;
          const bool chunk_valid = chunk < n_halves * 4;
          const bool want_xyz = args.xyzsum_slot[t] >= 0;
          float s0 = 0.f, s1 = 0.f, sx0 = 0.f, sx1 = 0.f, sy0 = 0.f, sy1 = 0.f, sz0 = 0.f, sz1 = 0.f;
          if (chunk_valid) {
            const uint8_t* cb = sA + chunk * A_CHUNK_BYTES;
#pragma unroll 4
            for (int r = 0; r < TILE_M; ++r) {
              const uint32_t w = *reinterpret_cast<const uint32_t*>(cb + r * 128 + (((lane >> 2) ^ (r & 7)) << 4) + ((lane & 3) << 2));
              const float lo = bf16_lo_to_f32(w), hi = bf16_hi_to_f32(w);
              s0 += lo; s1 += hi;
              if (want_xyz) {
                const float x = s_xyz[0][r], y = s_xyz[1][r], z = s_xyz[2][r];
                sx0 += x * lo; sx1 += x * hi; sy0 += y * lo; sy1 += y * hi; sz0 += z * lo; sz1 += z * hi;
              }
            }
          }
          float* part = args.tile_part + size_t(tile) * args.part_stride;
          const int col = chunk * 64 + lane * 2;
          if (args.colsum_slot[t] >= 0)
            *reinterpret_cast<float2*>(part + args.colsum_slot[t] * FH + col) = make_float2(s0, s1);
          if (want_xyz) {
            const int sl = args.xyzsum_slot[t];
            *reinterpret_cast<float2*>(part + (sl + 0) * FH + col) = make_float2(sx0, sx1);
            *reinterpret_cast<float2*>(part + (sl + 1) * FH + col) = make_float2(sy0, sy1);
            *reinterpret_cast<float2*>(part + (sl + 2) * FH + col) = make_float2(sz0, sz1);
          }
        }
        if (args.store_dys && t > 0) {
          if (threadIdx.x == 0) {
            const int nchunks = n_halves * 4;
            for (int c = 0; c < nchunks; ++c)
              tma_store_2d(&tmDys, sA + c * A_CHUNK_BYTES, c * 64, int32_t(int64_t(t) * args.n_rows_pad + int64_t(tile) * TILE_M));
            tma_store_commit();
          }
          stores_pending = true;
        }
      }
    }
    if (stores_pending && threadIdx.x == 0) tma_store_wait_all0();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == EPI_WARPS + 1) tmem_dealloc<512>(tmem_base);
}

int launch_fused_dgrad(const CUtensorMap& tmW, const CUtensorMap& tmDys, const BwdArgs& args, int grid, cudaStream_t stream) {
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(fused_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CHAIN_SMEM));
  fused_dgrad_kernel<<<grid, CHAIN_THREADS, CHAIN_SMEM, stream>>>(tmW, tmDys, args);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace dsdf
