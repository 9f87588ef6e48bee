// Weight-gradient contraction over the point axis (sm_100a, tcgen05 + TMEM + TMA):
//   dW_l[o, i] = sum_rows dY_l[row, o] * X_l[row, i]        X_l = stashed bf16 activations, dY_l = stashed bf16 dY
// Both operands are row-major [rows][512] in HBM, i.e. MN-major for this GEMM (the contraction index is the row), so the
// same SW128 TMA boxes {64 features x 32 rows} serve as UMMA operands through MN-major descriptors
// (LBO = 4096 between 64-feature groups, SBO = 1024 between 8-row groups).
// Work unit = slab (layer, 128 output features) x split of the row range; one CTA per unit keeps the slab's
// [128 x N<=512] fp32 accumulator in TMEM for its whole row range and writes one partial, which dsdf_fused_finish
// reduces in a fixed order (deterministic split-K).
// Replaces the weight-gradient part of chunk_loss.backward() (train_deep_sdf.py:533).
#include "common.cuh"
#include "fused.cuh"
#include "host_util.h"

namespace dsdf {

namespace {
constexpr int WG_ROWS = 32;                       // rows (K) per pipeline stage
constexpr int WG_BOX_BYTES = WG_ROWS * 128;       // one {64 x 32} bf16 box = 4 KB
constexpr int WG_STAGE_BYTES = 10 * WG_BOX_BYTES; // A: 2 boxes (128 out features), B: up to 8 boxes (512 in features)
constexpr int WG_STAGES = 5;
constexpr int WG_THREADS = 192;                   // warp 0 TMA, warp 1 MMA + TMEM alloc, warps 2..5 epilogue
constexpr int WG_SMEM = WG_STAGES * WG_STAGE_BYTES + 1024;
}  // namespace

__global__ void __launch_bounds__(WG_THREADS, 1)
wgrad_kernel(const __grid_constant__ CUtensorMap tmDys, const __grid_constant__ CUtensorMap tmActs,
             const __grid_constant__ WgradArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_full[WG_STAGES], bar_empty[WG_STAGES], bar_done;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- which unit is mine
  int slab = -1, split = 0;
  for (int s = 0; s < args.n_slabs; ++s) {
    const int first = args.slab[s].first_cta;
    if (int(blockIdx.x) >= first && int(blockIdx.x) < first + args.slab[s].splits) { slab = s; split = blockIdx.x - first; }
  }
  if (slab < 0) return;
  const WgradSlab S = args.slab[slab];
  const int64_t total_steps = args.n_rows_pad / WG_ROWS;
  const int64_t step_beg = total_steps * split / S.splits;
  const int64_t step_end = total_steps * (split + 1) / S.splits;
  const int n_boxes_b = S.n / 64;
  const uint32_t stage_tx = uint32_t(2 + n_boxes_b) * WG_BOX_BYTES;

  if (threadIdx.x == 0) {
    for (int s = 0; s < WG_STAGES; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
    mbar_init(&bar_done, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(&tmem_slot);
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmDys); tma_prefetch_desc(&tmActs); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      const int64_t dy_row0 = int64_t(S.dy_layer) * args.n_rows_pad;
      const int64_t x_row0 = int64_t(S.x_layer) * args.n_rows_pad;
      for (int64_t st = step_beg; st < step_end; ++st) {
        mbar_wait(&bar_empty[stage], phase ^ 1);
        mbar_arrive_expect_tx(&bar_full[stage], stage_tx);
        uint8_t* base = smem + stage * WG_STAGE_BYTES;
        const int32_t r = int32_t(st * WG_ROWS);
        tma_load_2d(base, &tmDys, &bar_full[stage], S.m0, int32_t(dy_row0) + r);
        tma_load_2d(base + WG_BOX_BYTES, &tmDys, &bar_full[stage], S.m0 + 64, int32_t(dy_row0) + r);
        for (int b = 0; b < n_boxes_b; ++b)
          tma_load_2d(base + (2 + b) * WG_BOX_BYTES, &tmActs, &bar_full[stage], b * 64, int32_t(x_row0) + r);
        if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = umma_idesc_bf16(128, 256, 1, 1);            // both operands MN-major
    const uint64_t desc_hi = umma_desc_hi_sw128(WG_BOX_BYTES, 1024);   // LBO = next 64-feature group, SBO = next 8 rows
    const uint32_t s_u = smem_u32(smem);
    int stage = 0; uint32_t phase = 0;
    const int n_halves = S.n / 256;
    for (int64_t st = step_beg; st < step_end; ++st) {
      mbar_wait(&bar_full[stage], phase);
      tc_fence_after();
      if (lane == 0) {
        const uint32_t a0 = s_u + stage * WG_STAGE_BYTES;
        const uint32_t b0 = a0 + 2 * WG_BOX_BYTES;
#pragma unroll
        for (int ks = 0; ks < WG_ROWS / 16; ++ks) {
          const uint64_t ad = umma_desc(a0 + ks * 2048, desc_hi);
          for (int h = 0; h < n_halves; ++h) {
            const uint64_t bd = umma_desc(b0 + h * 4 * WG_BOX_BYTES + ks * 2048, desc_hi);
            umma_bf16_ss(tmem_base + h * 256, ad, bd, idesc, (st > step_beg || ks > 0) ? 1u : 0u);
          }
        }
        umma_commit(&bar_empty[stage]);
      }
      __syncwarp();
      if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
    }
    if (lane == 0) umma_commit(&bar_done);
    __syncwarp();
  } else {
    // epilogue: TMEM -> fp32 partial [128][n]
    const int q = warp & 3;
    const int row = q * 32 + lane;
    float* out = args.partials + S.part_offset + size_t(split) * 128 * S.n + size_t(row) * S.n;
    if (step_end > step_beg) {
      mbar_wait(&bar_done, 0);
      tc_fence_after();
      for (int c0 = 0; c0 < S.n; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_base + (uint32_t(q * 32) << 16) + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 8; ++i)
          reinterpret_cast<float4*>(out + c0)[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                                                 __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
      }
    } else {
      for (int c0 = 0; c0 < S.n; c0 += 4) *reinterpret_cast<float4*>(out + c0) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// Last layer (512 -> 1): dW_last[c] = sum_r dy[r] * X_last[r, c];  db_last = sum_r dy[r].
// grid = n_blocks, 256 threads; thread t owns columns 2t, 2t+1.  part[block][513].
__global__ void __launch_bounds__(256) wgrad_last_kernel(const __nv_bfloat16* __restrict__ x_last, const float* __restrict__ dy,
                                                         int64_t n_rows_pad, float* __restrict__ part) {
  const int64_t rows_per = (n_rows_pad + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = rows_per * blockIdx.x;
  int64_t r1 = r0 + rows_per;
  if (r1 > n_rows_pad) r1 = n_rows_pad;
  float s0 = 0.f, s1 = 0.f, sd = 0.f;
  const uint32_t* xw = reinterpret_cast<const uint32_t*>(x_last);
  for (int64_t r = r0; r < r1; ++r) {
    const float d = __ldg(dy + r);
    const uint32_t w = __ldg(xw + r * (FH / 2) + threadIdx.x);
    s0 += d * bf16_lo_to_f32(w);
    s1 += d * bf16_hi_to_f32(w);
    sd += d;
  }
  float* o = part + size_t(blockIdx.x) * (FH + 1);
  o[2 * threadIdx.x] = s0;
  o[2 * threadIdx.x + 1] = s1;
  if (threadIdx.x == 0) o[FH] = sd;
}

int launch_wgrad(const CUtensorMap& tmDys, const CUtensorMap& tmActs, const WgradArgs& args, int grid, cudaStream_t stream) {
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM));
  wgrad_kernel<<<grid, WG_THREADS, WG_SMEM, stream>>>(tmDys, tmActs, args);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int launch_wgrad_last(const void* x_last, const float* dy, int64_t n_rows_pad, float* part, int n_blocks, cudaStream_t stream) {
  wgrad_last_kernel<<<n_blocks, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(x_last), dy, n_rows_pad, part);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace dsdf
