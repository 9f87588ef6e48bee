// Fused DeepSDF decoder forward for hidden width 512 (sm_100a, tcgen05 + TMEM + TMA).
//
// One persistent CTA per SM walks 128-row point tiles.  For a tile, ALL hidden layers run back to back:
//   * the activation tile [128 x 512] lives in shared memory as bf16 in the SW128 K-major UMMA layout (128 KB),
//   * weights are TMA-streamed from L2 as [256 out x 64 K] bf16 stages (3-stage mbarrier ring),
//   * one thread issues tcgen05.mma (M128 x N256 x K16); the fp32 accumulators [128 x 512] fill the SM's TMEM,
//   * 8 epilogue warps read TMEM (tcgen05.ld), add bias (or the per-scene bias that carries the latent code),
//     add the xyz columns in fp32, ReLU, Philox dropout, pack to bf16 and write the next layer's A operand.
//     Accumulator half 0 is post-processed into REGISTERS while the tensor core still computes half 1, and is
//     dumped into the A buffer only when the layer's last MMA has retired; half 1 is post-processed while the
//     tensor core already runs the first K-chunks of the next layer (which only need half 0's columns).
//   * the final 512 -> 1 layer, tanh, clamp, L1 loss and dL/dy are fused into the last epilogue,
//   * in training, each layer's activation tile is TMA-stored (bf16) for wgrad, plus 1 bit/element relu*dropout mask.
// Replaces Decoder.forward (deep_sdf/networks/deep_sdf_decoder.py:76-111) and train_deep_sdf.py:511-519.
#include "common.cuh"
#include "fused.cuh"
#include "host_util.h"

namespace dsdf {

namespace {

constexpr uint32_t kBarEpi = 1;   // named barrier id for the 256 epilogue threads

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync %0, %1;" ::"n"(kBarEpi), "n"(EPI_WARPS * 32) : "memory"); }

struct EpiCtx {
  const float* bias;     // [512] for this layer (layer bias or scene bias), never null
  const float* wxyz;     // [3][512] or null
  float x, y, z;         // this thread's point
  bool dropout;
  uint32_t thr;
  float scale;
  uint64_t seed;
  uint64_t grow;         // global row id
  uint32_t layer;
  const float* w_last;   // non-null on the last hidden layer
};

// Post-process 32 consecutive columns [c0, c0+32) of one row.  v: fp32 pre-activations (accumulators or zeros).
// Returns 16 packed bf16x2 words in p, the relu*dropout keep bits, and accumulates the final-layer dot product.
template <bool kTrain>
__device__ __forceinline__ void epi_group(float (&v)[32], uint32_t c0, const EpiCtx& cx, uint32_t (&p)[16], uint32_t& bits,
                                          float& ydot) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    float4 b = __ldg(reinterpret_cast<const float4*>(cx.bias + c0) + j);
    v[4 * j + 0] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
  }
  if (cx.wxyz != nullptr) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 wx = __ldg(reinterpret_cast<const float4*>(cx.wxyz + c0) + j);
      float4 wy = __ldg(reinterpret_cast<const float4*>(cx.wxyz + FH + c0) + j);
      float4 wz = __ldg(reinterpret_cast<const float4*>(cx.wxyz + 2 * FH + c0) + j);
      v[4 * j + 0] += cx.x * wx.x + cx.y * wy.x + cx.z * wz.x;
      v[4 * j + 1] += cx.x * wx.y + cx.y * wy.y + cx.z * wz.y;
      v[4 * j + 2] += cx.x * wx.z + cx.y * wy.z + cx.z * wz.z;
      v[4 * j + 3] += cx.x * wx.w + cx.y * wy.w + cx.z * wz.w;
    }
  }
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.0f);
  if (kTrain && cx.dropout) {
    const uint32_t keep = dropout_keep_word(cx.seed, cx.layer, cx.grow, c0 >> 5, cx.thr);
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = ((keep >> i) & 1u) ? v[i] * cx.scale : 0.0f;
  }
  bits = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) bits |= (v[i] > 0.0f ? 1u : 0u) << i;
  if (cx.w_last != nullptr) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 w = __ldg(reinterpret_cast<const float4*>(cx.w_last + c0) + j);
      ydot += v[4 * j + 0] * w.x + v[4 * j + 1] * w.y + v[4 * j + 2] * w.z + v[4 * j + 3] * w.w;
    }
  }
#pragma unroll
  for (int i = 0; i < 16; ++i) p[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
}

// write 32 packed columns [c0, c0+32) of row `row` into the SW128 A tile
__device__ __forceinline__ void store_group(uint8_t* sA, uint32_t row, uint32_t c0, const uint32_t* p) {
  uint8_t* base = sA + (c0 >> 6) * A_CHUNK_BYTES + row * 128u;
  const uint32_t piece0 = (c0 & 63u) >> 3;
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    uint32_t piece = (piece0 + t) ^ (row & 7u);
    *reinterpret_cast<uint4*>(base + piece * 16u) = make_uint4(p[4 * t], p[4 * t + 1], p[4 * t + 2], p[4 * t + 3]);
  }
}

}  // namespace

template <bool kTrain>
__global__ void __launch_bounds__(CHAIN_THREADS, 1)
fused_forward_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmActs,
                     const __grid_constant__ FwdArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + A_BYTES;
  __shared__ uint64_t bar_full[NUM_STAGES], bar_empty[NUM_STAGES], bar_tmem_full[2], bar_a_ready[2];
  __shared__ uint32_t tmem_slot;
  __shared__ float s_ypart[TILE_M];
  __shared__ float s_loss[4];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < NUM_STAGES; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
    mbar_init(&bar_tmem_full[0], 1); mbar_init(&bar_tmem_full[1], 1);
    mbar_init(&bar_a_ready[0], EPI_WARPS); mbar_init(&bar_a_ready[1], EPI_WARPS);
    fence_barrier_init();
  }
  if (warp == EPI_WARPS + 1) tmem_alloc<512>(&tmem_slot);
  if (warp == EPI_WARPS && lane == 0) { tma_prefetch_desc(&tmW); tma_prefetch_desc(&tmActs); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int nh = args.num_hidden;

  if (warp == EPI_WARPS) {
    // ================================================================ TMA producer (weights)
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x) {
        for (int l = 0; l < nh; ++l) {
          const ChainLayer& L = args.prog[l];
          if (!L.has_mma) continue;
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
      for (int l = 0; l < nh; ++l) {
        const ChainLayer& L = args.prog[l];
        if (!L.has_mma) continue;
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
          // keep both barriers in lock step; half 1 of TMEM is untouched by this layer
          if (!waited1) mbar_wait(&bar_a_ready[1], par);
          if (lane == 0) umma_commit(&bar_tmem_full[1]);
          __syncwarp();
        }
      }
    }
  } else {
    // ================================================================ epilogue warps
    const int q = warp & 3;            // TMEM lane quadrant
    const int j = warp >> 2;           // which 128-column slice of each accumulator half
    const uint32_t rt = q * 32 + lane; // row within tile == TMEM lane
    const uint32_t tmem_row = tmem_base + (uint32_t(q * 32) << 16);
    uint32_t mma_iter = 0;
    bool stores_pending = false;
    for (int tile = blockIdx.x; tile < args.n_tiles; tile += gridDim.x) {
      const int64_t lrow = int64_t(tile) * TILE_M + rt;
      const bool valid = lrow < args.n_rows;
      EpiCtx cx;
      cx.x = cx.y = cx.z = 0.0f;
      if (valid) { const float* pt = args.xyz + lrow * args.xyz_stride; cx.x = __ldg(pt); cx.y = __ldg(pt + 1); cx.z = __ldg(pt + 2); }
      cx.thr = args.drop_thr; cx.scale = args.drop_scale; cx.seed = args.seed; cx.grow = args.row0 + uint64_t(lrow);
      const float* cb = args.cbias + size_t(__ldg(args.tile_scene + tile)) * 2 * FH;
      float ydot = 0.0f;

      for (int l = 0; l < nh; ++l) {
        const ChainLayer& L = args.prog[l];
        cx.bias = (L.bias_slot >= 0) ? (cb + L.bias_slot * FH) : args.bias[l];
        cx.wxyz = L.has_xyz ? args.wxyz[l] : nullptr;
        cx.dropout = L.dropout != 0;
        cx.layer = uint32_t(l);
        cx.w_last = (l == nh - 1) ? args.w_last : nullptr;
        const bool produce_a = (l < nh - 1);       // the last hidden layer feeds no MMA
        uint32_t mask_words[4];
        uint32_t hold[64];
        uint32_t par = 0;
        if (L.has_mma) { par = mma_iter & 1; ++mma_iter; }

        // ---- half 0 -> registers (overlaps the tensor core working on half 1)
        if (L.has_mma) { mbar_wait(&bar_tmem_full[0], par); tc_fence_after(); }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const uint32_t c0 = j * 128 + g * 32;
          float v[32];
          if (L.has_mma) {
            uint32_t raw[32];
            tmem_ld32(tmem_row + c0, raw);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]);
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = 0.0f;
          }
          uint32_t p[16];
          epi_group<kTrain>(v, c0, cx, p, mask_words[g], ydot);
#pragma unroll
          for (int i = 0; i < 16; ++i) hold[g * 16 + i] = p[i];
        }
        if (kTrain && args.bitmask != nullptr) {
          uint32_t* dst = args.bitmask + ((size_t(l) * (args.n_rows_pad / TILE_M) + tile) * MASK_WORDS + j * 4) * TILE_M + rt;
#pragma unroll
          for (int g = 0; g < 4; ++g) dst[size_t(g) * TILE_M] = mask_words[g];
        }
        // ---- all MMAs of this layer retired -> the A tile may be overwritten
        if (L.has_mma) { mbar_wait(&bar_tmem_full[1], par); tc_fence_after(); }
        if (kTrain && stores_pending) {
          if (threadIdx.x == 0) tma_store_wait_read0();
          epi_bar_sync();
          stores_pending = false;
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) store_group(sA, rt, j * 128 + g * 32, &hold[g * 16]);
        fence_proxy_async_smem();
        tc_fence_before();
        __syncwarp();
        if (produce_a && lane == 0) mbar_arrive(&bar_a_ready[0]);

        // ---- half 1 straight into the A tile (overlaps the next layer's first K chunks)
        if (L.n_halves == 2) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const uint32_t c0 = HALF_N + j * 128 + g * 32;
            float v[32];
            if (L.has_mma) {
              uint32_t raw[32];
              tmem_ld32(tmem_row + c0, raw);
              tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]);
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = 0.0f;
            }
            uint32_t p[16];
            epi_group<kTrain>(v, c0, cx, p, mask_words[g], ydot);
            store_group(sA, rt, c0, p);
          }
          if (kTrain && args.bitmask != nullptr) {
            uint32_t* dst = args.bitmask + ((size_t(l) * (args.n_rows_pad / TILE_M) + tile) * MASK_WORDS + 8 + j * 4) * TILE_M + rt;
#pragma unroll
            for (int g = 0; g < 4; ++g) dst[size_t(g) * TILE_M] = mask_words[g];
          }
          fence_proxy_async_smem();
        } else if (kTrain && args.bitmask != nullptr) {
          uint32_t* dst = args.bitmask + ((size_t(l) * (args.n_rows_pad / TILE_M) + tile) * MASK_WORDS + 8 + j * 4) * TILE_M + rt;
#pragma unroll
          for (int g = 0; g < 4; ++g) dst[size_t(g) * TILE_M] = 0u;
        }
        tc_fence_before();
        __syncwarp();
        if (produce_a && lane == 0) mbar_arrive(&bar_a_ready[1]);

        // ---- stash the activation tile for wgrad
        if (kTrain && args.store_acts) {
          epi_bar_sync();
          if (threadIdx.x == 0) {
            const int nchunks = L.n_halves * 4;
            for (int c = 0; c < nchunks; ++c)
              tma_store_2d(&tmActs, sA + c * A_CHUNK_BYTES, c * 64, int32_t(int64_t(l) * args.n_rows_pad + int64_t(tile) * TILE_M));
            tma_store_commit();
          }
          stores_pending = true;
        }
      }

      // ---- final layer: combine the two column slices, tanh, loss
      if (j == 1) s_ypart[rt] = ydot;
      epi_bar_sync();
      if (j == 0) {
        float yv = ydot + s_ypart[rt] + __ldg(args.b_last);
        float t1 = tanhf(yv);
        float t = t1, dt = 1.0f - t1 * t1;
        if (args.use_tanh) { t = tanhf(t1); dt *= (1.0f - t * t); }
        if (valid) args.sdf_out[lrow] = t;
        if (kTrain && args.gt != nullptr) {
          float g = valid ? __ldg(args.gt + lrow * args.gt_stride) : 0.0f;
          float pc = t, gc = g, inside = 1.0f;
          if (args.clamp > 0.0f) {
            pc = fminf(fmaxf(t, -args.clamp), args.clamp);
            gc = fminf(fmaxf(g, -args.clamp), args.clamp);
            inside = (t >= -args.clamp && t <= args.clamp) ? 1.0f : 0.0f;
          }
          float diff = pc - gc;
          float sgn = (diff > 0.0f) ? 1.0f : ((diff < 0.0f) ? -1.0f : 0.0f);
          float dyv = valid ? sgn * inside * dt * args.inv_n : 0.0f;
          args.dy_out[lrow] = dyv;
          float a = valid ? fabsf(diff) : 0.0f;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
          if (lane == 0) s_loss[q] = a;
        }
      }
      epi_bar_sync();
      if (kTrain && args.gt != nullptr && threadIdx.x == 0)
        args.tile_loss[tile] = ((s_loss[0] + s_loss[1]) + s_loss[2]) + s_loss[3];
    }
    if (kTrain && stores_pending && threadIdx.x == 0) tma_store_wait_all0();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == EPI_WARPS + 1) tmem_dealloc<512>(tmem_base);
}

int launch_fused_forward(const CUtensorMap& tmW, const CUtensorMap& tmActs, const FwdArgs& args, int grid,
                         cudaStream_t stream) {
  auto kern = args.training ? fused_forward_kernel<true> : fused_forward_kernel<false>;
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, CHAIN_SMEM));
  kern<<<grid, CHAIN_THREADS, CHAIN_SMEM, stream>>>(tmW, tmActs, args);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace dsdf
