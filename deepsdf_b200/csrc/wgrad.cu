// Weight-gradient contraction over the point axis (sm_100a, tcgen05 cta_group::2 + TMEM + TMA multicast, 4-CTA clusters):
//   dW_l[o, i] = sum_rows dY_l[row, o] * X_{l-1}[row, i]     X = stashed bf16 activations, dY = stashed bf16 dY
// Both operands are row-major [rows][512] in HBM, i.e. MN-major for this GEMM (the contraction index is the row), so the
// same SW128 TMA boxes {64 features x 32 rows} serve as UMMA operands through MN-major descriptors
// (LBO = 4096 between 64-feature groups, SBO = 1024 between 8-row groups).
// A cluster of four CTAs owns a row segment of one layer.  CTA r computes output features [128 r, 128 r + 128) into a
// [128 x n<=512] fp32 TMEM accumulator.  The cluster is two CTA PAIRS (ranks {0,1} and {2,3}): a pair runs M = 256
// cta_group::2 MMAs and splits the X operand between its two CTAs (even rank: the lower 128 features of every 256-wide
// N block, odd rank: the upper 128), so a CTA stages only 24 KB per 32 rows (own dY 8 KB + half of X 16 KB) instead of
// 40 KB: 8 stages fit, and the per-SM fill rate drops from 80 to 48 B/clk (the 40 KB version sat at 48 % tensor
// utilisation with neither HBM nor L2 busy).  The same X half is needed by the same-parity CTA of the other pair:
// each of the two fetches one 64-feature box and multicasts it to both, so X is still read from L2 once per cluster.
// Barriers: every CTA has its own full barrier (own TMA + the partner's multicast); the odd CTA's spare warp relays its
// full phases to the pair leader; MMA commits arrive (multicast) on the empty barrier of all four CTAs.
// (An L2 prefetch of the producer's boxes 8-48 steps ahead was measured SLOWER, 3.9 -> 4.7-5.7 ms: more requests in flight
// only add DRAM row conflicts for this 128-byte-piece access pattern.)
// One partial per (segment, CTA), reduced by dsdf_fused_finish in plan order (deterministic split-K).
// Replaces the weight-gradient part of chunk_loss.backward() (train_deep_sdf.py:533).
#include "common.cuh"
#include "fused.cuh"
#include "host_util.h"

namespace dsdf {

namespace {
constexpr int WG_ROWS = 32;                       // rows (K) per pipeline stage
constexpr int WG_BOX_BYTES = WG_ROWS * 128;       // one {64 x 32} bf16 box = 4 KB
constexpr int WG_STAGE_BYTES = 6 * WG_BOX_BYTES;  // A: 2 boxes (128 out features), B: 2 boxes per 256-wide N block (this CTA's half)
constexpr int WG_STAGES = 9;
constexpr int WG_THREADS = 192;                   // warp 0 TMA, warp 1 MMA (even rank) / relay (odd rank) + TMEM alloc, warps 2..5 epilogue
constexpr int WG_SMEM = WG_STAGES * WG_STAGE_BYTES + 1024;
}  // namespace

__global__ void __cluster_dims__(4, 1, 1) __launch_bounds__(WG_THREADS, 1)
wgrad_kernel(const __grid_constant__ CUtensorMap tmDys, const __grid_constant__ CUtensorMap tmActs,
             const __grid_constant__ WgradArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_full[WG_STAGES], bar_peer_full[WG_STAGES], bar_empty[WG_STAGES], bar_done, bar_epi_free;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();          // = 128-feature output block of the layer
  const uint32_t par_r = rank & 1u, pair = rank >> 1;
  const uint32_t pair_leader = rank & ~1u;
  const int cluster = blockIdx.x >> 2;
  const int seg0 = args.plan.cluster_seg0[cluster], seg1 = args.plan.cluster_seg0[cluster + 1];

  if (threadIdx.x == 0) {
    for (int s = 0; s < WG_STAGES; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_peer_full[s], 1); mbar_init(&bar_empty[s], 2); }
    mbar_init(&bar_done, 1);
    mbar_init(&bar_epi_free, 8);                    // 4 epilogue warps of each CTA of the pair
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_pair<512>(&tmem_slot);
  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmDys); tma_prefetch_desc(&tmActs); }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    // ================================================================ TMA producer (whole warp walks, one lane issues)
    const uint32_t full0 = smem_u32(&bar_full[0]), empty0 = smem_u32(&bar_empty[0]), s_u = smem_u32(smem);
    const uint16_t mask = uint16_t(0b0101u << par_r);                  // the same-parity CTA of both pairs
    uint32_t stage = 0, phase = 0;
    for (int sg = seg0; sg < seg1; ++sg) {
      const int layer = args.plan.seg[sg].layer, n_halves = args.plan.seg[sg].n >> 8;
      const int st_beg = args.plan.seg[sg].step_beg, st_end = args.plan.seg[sg].step_end;
      const uint32_t stage_tx = uint32_t(2 + 2 * n_halves) * WG_BOX_BYTES;
      const int32_t dy_row0 = int32_t(int64_t(layer) * args.n_rows_pad);
      const int32_t x_row0 = int32_t(int64_t(layer - 1) * args.n_rows_pad);
      for (int st = st_beg; st < st_end; ++st) {
        mbar_wait_a(empty0 + stage * 8, phase ^ 1);
        if (elect_one()) {
          const uint32_t base = s_u + stage * WG_STAGE_BYTES, bar = full0 + stage * 8;
          const int32_t r = st * WG_ROWS;
          mbar_arrive_expect_tx_a(bar, stage_tx);
          tma_load_2d_a(base, &tmDys, bar, int32_t(rank) * 128, dy_row0 + r);
          tma_load_2d_a(base + WG_BOX_BYTES, &tmDys, bar, int32_t(rank) * 128 + 64, dy_row0 + r);
          for (int h = 0; h < n_halves; ++h)       // this CTA fetches box `pair` of its parity's half for both same-parity CTAs
            tma_load_2d_mcast_a(base + (2 + 2 * h + pair) * WG_BOX_BYTES, &tmActs, bar, 256 * h + 128 * int(par_r) + 64 * int(pair),
                                x_row0 + r, mask);
        }
        __syncwarp();
        if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 && par_r == 1) {
    // ================================================================ relay (odd CTA): tell the pair leader that this CTA's
    // stage has landed
    const uint32_t full0 = smem_u32(&bar_full[0]);
    const uint32_t peer0 = mapa_u32(smem_u32(&bar_peer_full[0]), pair_leader);
    uint32_t stage = 0, phase = 0;
    for (int sg = seg0; sg < seg1; ++sg) {
      const int st_beg = args.plan.seg[sg].step_beg, st_end = args.plan.seg[sg].step_end;
      for (int st = st_beg; st < st_end; ++st) {
        mbar_wait_a(full0 + stage * 8, phase);
        if (elect_one()) mbar_arrive_remote(peer0 + stage * 8);
        __syncwarp();
        if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ================================================================ MMA issuer (pair leader; warp-uniform loop, elected issue)
    const uint32_t idesc = umma_idesc_bf16(256, 256, 1, 1);            // M = 256 over the pair, both operands MN-major
    const uint64_t desc_hi = umma_desc_hi_sw128(WG_BOX_BYTES, 1024);   // LBO = next 64-feature group, SBO = next 8 rows
    const uint64_t d_base = desc_hi | uint64_t((smem_u32(smem) >> 4) & 0x3FFF);
    const uint32_t full0 = smem_u32(&bar_full[0]), peer0 = smem_u32(&bar_peer_full[0]), empty0 = smem_u32(&bar_empty[0]);
    const uint32_t done_a = smem_u32(&bar_done), free_a = smem_u32(&bar_epi_free);
    const uint16_t pair_mask = uint16_t(0b11u << (2 * pair));
    uint32_t stage = 0, phase = 0;
    for (int sg = seg0; sg < seg1; ++sg) {
      const int n_halves = args.plan.seg[sg].n >> 8;
      const int st_beg = args.plan.seg[sg].step_beg, st_end = args.plan.seg[sg].step_end;
      if (sg > seg0) { mbar_wait_a(free_a, uint32_t(sg - seg0 - 1) & 1u); tc_fence_after(); }   // accumulators drained
      for (int st = st_beg; st < st_end; ++st) {
        mbar_wait_a(full0 + stage * 8, phase);
        mbar_wait_a(peer0 + stage * 8, phase);
        tc_fence_after();
        const uint64_t a0 = d_base + uint64_t(stage * (WG_STAGE_BYTES >> 4));
        const uint64_t b0 = a0 + uint64_t((2 * WG_BOX_BYTES) >> 4);
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < WG_ROWS / 16; ++ks) {
            const uint32_t acc = (st > st_beg || ks > 0) ? 1u : 0u;
            umma_bf16_ss_pair(tmem_base, a0 + ks * 128, b0 + ks * 128, idesc, acc);
            if (n_halves == 2) umma_bf16_ss_pair(tmem_base + 256, a0 + ks * 128, b0 + uint64_t((2 * WG_BOX_BYTES) >> 4) + ks * 128, idesc, acc);
          }
          umma_commit_pair_a(empty0 + stage * 8, 0xF);
          if (st == st_end - 1) umma_commit_pair_a(done_a, pair_mask);
        }
        __syncwarp();
        if (++stage == WG_STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp >= 2) {
    // ================================================================ epilogue: TMEM -> fp32 partial [128][n]
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t free_leader = mapa_u32(smem_u32(&bar_epi_free), pair_leader);
    for (int sg = seg0; sg < seg1; ++sg) {
      const int n = args.plan.seg[sg].n;
      float* out = args.partials + args.plan.seg[sg].part_offset + (size_t(rank) * 128 + row) * n;
      mbar_wait_sleep(&bar_done, uint32_t(sg - seg0) & 1u, 2000);
      tc_fence_after();
      for (int c0 = 0; c0 < n; c0 += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_base + (uint32_t(q * 32) << 16) + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 8; ++i)
          reinterpret_cast<float4*>(out + c0)[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]),
                                                                 __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_remote(free_leader);
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_pair<512>(tmem_base);
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

int launch_wgrad(const CUtensorMap& tmDys, const CUtensorMap& tmActs, const WgradArgs& args, cudaStream_t stream) {
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM));
  DSDF_LAUNCHED;
  wgrad_kernel<<<4 * args.plan.n_clusters, WG_THREADS, WG_SMEM, stream>>>(tmDys, tmActs, args);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int wgrad_max_active_clusters() {
  cudaFuncSetAttribute(wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(4 * WG_CLUSTERS);
  cfg.blockDim = dim3(WG_THREADS);
  cfg.dynamicSmemBytes = WG_SMEM;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 4; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, wgrad_kernel, &cfg) != cudaSuccess) { cudaGetLastError(); n = -1; }
  return n;
}

int launch_wgrad_last(const void* x_last, const float* dy, int64_t n_rows_pad, float* part, int n_blocks, cudaStream_t stream) {
  DSDF_LAUNCHED;
  wgrad_last_kernel<<<n_blocks, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(x_last), dy, n_rows_pad, part);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace dsdf
