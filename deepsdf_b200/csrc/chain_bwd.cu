// Fused DeepSDF decoder backward, data-gradient chain, CTA-pair version (sm_100a: tcgen05 cta_group::2 + TMEM + TMA).
//
// Mirror image of chain_fwd.cu (same pair / pipeline structure, see there).  Per 128-row tile, for t = last hidden layer
// .. 0 the kernel produces
//   dY_t = dX_{t+1} (.) gate_t        dX_{t+1} = dY_{t+1} . W_{t+1}   (tensor cores, W^T streamed through the TMA ring)
// with dY of the last hidden layer seeded by the rank-1 product dL/dy * w_last.  The dY tile lives in shared memory as
// bf16 (SW128: the A operand of the next chain link), the gates are the 1-bit relu*dropout words written by the forward;
// the inverted-dropout scale lives in the packed W^T blocks (training variant of the pack, fused_host.cu), so the
// epilogue only gates.  Once a tile of dY_t is complete the 16 epilogue warps also emit
//   * its column sums (bias gradients; the per-scene sums of dY_0 / dY_skip ARE the latent-code gradient up to a tiny
//     [S x 512] x [512 x L] product, so the latent-gradient reduction is a fixed-order sum of per-tile partials),
//   * the xyz-weighted column sums of dY_0 / dY_skip (gradients of the xyz weight columns),
//   * the bf16 dY tile itself (TMA store) that wgrad.cu contracts with the stashed activations.
// The data gradient w.r.t. lin0's input is never formed: training only needs it reduced per scene.
// Replaces the autograd graph of train_deep_sdf.py:533 (data-gradient part).
#include "common.cuh"
#include "fused.cuh"
#include "host_util.h"

namespace dsdf {

namespace {

constexpr int B_EPI_WARPS = 16;
constexpr int B_EPI_THREADS = B_EPI_WARPS * 32;
constexpr int B_THREADS = (B_EPI_WARPS + 4) * 32;   // + TMA warp + MMA warp + stash warp + one idle warp (completes the service
                                                    //   warpgroup: setmaxnreg works on whole warpgroups)
constexpr int B_EPI_REGS = 112, B_SVC_REGS = 24;    // register re-split: 16 x 112 + 4 x 24 <= 20 x 96 per thread-slot
constexpr int B_STAGES = 5;
constexpr int B_STAGE_BYTES = 128 * 128;
constexpr int B_OFF_B = A_BYTES;
constexpr int B_OFF_WLAST = B_OFF_B + B_STAGES * B_STAGE_BYTES;   // f32 [512]
constexpr int B_OFF_XYZ = B_OFF_WLAST + FH * 4;                   // float4 (x, y, z, -) of row r at slot (r & 7) * 17 + (r >> 3):
                                                                  //   the column sums walk rows r0 + 8 i with ONE LDS.128 per row,
                                                                  //   and the eight row classes of a warp fall on distinct banks
constexpr int B_SMEM = B_OFF_XYZ + 8 * 17 * 16 + 1024;

__device__ __forceinline__ void bwd_bar() { asm volatile("bar.sync 1, %0;" ::"n"(B_EPI_THREADS) : "memory"); }

// packed fp32 pairs (sm_100 FFMA2)
__device__ __forceinline__ uint64_t pack_f32x2(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void unpack_f32x2(uint64_t v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ void fma_f32x2(uint64_t& d, uint64_t a, uint64_t b) { asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b)); }

// 16 columns [c0, c0+16) of one row: v = accumulators (or dy * w_last), gated and scaled, packed to bf16 into the A tile
__device__ __forceinline__ void bwd_pass16(uint32_t taddr, bool has_acc, float dyv, uint32_t swl, uint32_t gate16,
                                           uint32_t (&pk)[8]) {
  // swl: SHARED-space byte address (explicit ld.shared).  pk: the 16 gated values as packed bf16 pairs - the caller stores
  // them (two 16-byte pieces of this row's SW128 line) once the tile quarter may be overwritten
  float v[16];
  if (has_acc) {
    uint32_t r[16];
    tmem_ld16(taddr, r);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4 w = lds_f32x4(swl + 16 * k);
      const float d = dyv;
      v[4 * k + 0] = d * w.x; v[4 * k + 1] = d * w.y; v[4 * k + 2] = d * w.z; v[4 * k + 3] = d * w.w;
    }
  }
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = (gate16 & (1u << i)) ? v[i] : 0.0f;
#pragma unroll
  for (int i = 0; i < 8; ++i) pk[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
}

}  // namespace

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(B_THREADS, 1)
chain_dgrad_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmDys,
                   const __grid_constant__ BwdArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + B_OFF_B;
  // 32-bit shared-space base for everything the epilogue touches (see chain_fwd.cu)
  const uint32_t smem_a = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint64_t bar_full[B_STAGES], bar_empty[B_STAGES], bar_tmem_full[2], bar_free_q0, bar_free_lo, bar_a_ready[4];
  // per 128-column quarter of the dY tile: written by all 16 epilogue warps -> read by the stash warp (TMA store) and by
  // the four warps that own its column sums -> released for the next chain link
  __shared__ uint64_t bar_q_written[4], bar_q_released[4];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < B_STAGES; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
    mbar_init(&bar_tmem_full[0], 1); mbar_init(&bar_tmem_full[1], 1);
    mbar_init(&bar_free_q0, 1);
    mbar_init(&bar_free_lo, 1);
    for (int i = 0; i < 4; ++i) mbar_init(&bar_a_ready[i], 2 * B_EPI_WARPS);
    for (int i = 0; i < 4; ++i) { mbar_init(&bar_q_written[i], B_EPI_WARPS); mbar_init(&bar_q_released[i], 5); }
    fence_barrier_init();
  }
  if (warp == B_EPI_WARPS + 1) tmem_alloc_pair<512>(&tmem_slot);
  if (warp == B_EPI_WARPS && lane == 0) { tma_prefetch_desc(&tmW); tma_prefetch_desc(&tmDys); }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int nh = args.num_hidden;
  const int n_pairs = (args.n_tiles + 1) >> 1;
  const int pair0 = blockIdx.x >> 1, pair_stride = gridDim.x >> 1;

  // Register re-split (each role issues its own setmaxnreg first): 96 registers per thread at launch (20 warps); the four
  // single-lane service warps keep 24, the sixteen epilogue warps get 112.
  if (warp == B_EPI_WARPS) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(B_SVC_REGS));
    // ================================================================ TMA producer (this CTA's half of every W^T block)
    const uint32_t full_leader = mapa_u32(smem_u32(&bar_full[0]), 0);
    const uint32_t full0 = smem_u32(&bar_full[0]), empty0 = smem_u32(&bar_empty[0]), sB_u = smem_u32(sB);
    uint32_t stage = 0, phase = 0;
    const uint64_t pol_w = l2_policy_evict_last();
    for (int tp = pair0; tp < n_pairs; tp += pair_stride) {
      for (int t = nh - 2; t >= 0; --t) {
        const int n_halves = args.prog[t + 1].n_halves, k_chunks = args.prog[t + 1].k_chunks;
        const int row_base = args.prog[t + 1].w_row0 + int(rank) * 128;
        for (int h = 0; h < n_halves; ++h)
          for (int kc = 0; kc < k_chunks; ++kc) {
            mbar_wait_a(empty0 + stage * 8, phase ^ 1);
            if (elect_one()) {
              if (leader) mbar_arrive_expect_tx_a(full0 + stage * 8, 2 * B_STAGE_BYTES);
              tma_load_2d_pair_hint_a(sB_u + stage * B_STAGE_BYTES, &tmW, full_leader + stage * 8, kc * 64, row_base + h * HALF_N, pol_w);
            }
            __syncwarp();
            if (++stage == B_STAGES) { stage = 0; phase ^= 1; }
          }
      }
    }
  } else if (warp == B_EPI_WARPS + 1) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(B_SVC_REGS));
    // ================================================================ MMA issuer (leader CTA; whole warp, elected issue)
    if (leader) {
      const uint32_t idesc = umma_idesc_bf16(2 * TILE_M, HALF_N, 0, 0);
      const uint64_t desc_hi = umma_desc_hi_sw128(16, 1024);
      const uint64_t a_base = desc_hi | uint64_t((smem_u32(sA) >> 4) & 0x3FFF);
      const uint64_t b_base = desc_hi | uint64_t((smem_u32(sB) >> 4) & 0x3FFF);
      const uint32_t full0 = smem_u32(&bar_full[0]), empty0 = smem_u32(&bar_empty[0]);
      const uint32_t tfull0 = smem_u32(&bar_tmem_full[0]), fq0 = smem_u32(&bar_free_q0), flo = smem_u32(&bar_free_lo);
      const uint32_t ready0 = smem_u32(&bar_a_ready[0]);
      uint32_t stage = 0, phase = 0, a_cons = 0;
      for (int tp = pair0; tp < n_pairs; tp += pair_stride) {
        for (int t = nh - 2; t >= 0; --t) {
          const int n_halves = args.prog[t + 1].n_halves, k_chunks = args.prog[t + 1].k_chunks;
          const uint32_t apar = a_cons & 1; ++a_cons;
          const int kc_lo_last = k_chunks < 4 ? k_chunks - 1 : 3;
          const int kc_q0_last = k_chunks < 2 ? k_chunks - 1 : 1;
          uint32_t waited = 0;
          for (int h = 0; h < n_halves; ++h) {
            const bool last_half = (h == n_halves - 1);
            for (int kc = 0; kc < k_chunks; ++kc) {
              uint32_t needm = 1u << (kc >> 1);
              if (kc == 0) needm |= 3u << (2 * h);      // the first MMA of a half overwrites accumulators the epilogue reads
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
              const uint64_t bd = b_base + uint64_t(stage * (B_STAGE_BYTES >> 4));
              const uint32_t d = tmem_base + h * HALF_N;
              if (elect_one()) {
                umma_bf16_ss_pair(d, ad, bd, idesc, kc != 0 ? 1u : 0u);
                umma_bf16_ss_pair(d, ad + 2, bd + 2, idesc, 1u);
                umma_bf16_ss_pair(d, ad + 4, bd + 4, idesc, 1u);
                umma_bf16_ss_pair(d, ad + 6, bd + 6, idesc, 1u);
                umma_commit_pair_a(empty0 + stage * 8, 0b11);
                if (last_half && kc == kc_q0_last) umma_commit_pair_a(fq0, 0b11);
                if (last_half && kc == kc_lo_last) umma_commit_pair_a(flo, 0b11);
                if (kc == k_chunks - 1) {
                  umma_commit_pair_a(tfull0 + h * 8, 0b11);
                  if (n_halves == 1) umma_commit_pair_a(tfull0 + 8, 0b11);
                }
              }
              __syncwarp();
              if (++stage == B_STAGES) { stage = 0; phase ^= 1; }
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
  } else if (warp == B_EPI_WARPS + 3) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(B_SVC_REGS));      // idle warp: completes the service warpgroup
  } else if (warp == B_EPI_WARPS + 2) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(B_SVC_REGS));
    // ================================================================ stash warp: TMA-store every completed quarter of
    // dY_t (t > 0) for wgrad and release it once the store has read the shared memory (one store kept in flight)
    const uint32_t wr0 = smem_u32(&bar_q_written[0]), rl0 = smem_u32(&bar_q_released[0]), sA_u = smem_u32(sA);
    uint32_t cnt = 0;
    int pending = -1;
    const uint64_t pol_s = l2_policy_evict_first();
    for (int tp = pair0; tp < n_pairs; tp += pair_stride) {
      const int tile = 2 * tp + int(rank);
      const bool tile_ok = tile < args.n_tiles;
      for (int t = nh - 1; t >= 0; --t) {
        const int nq = (t < nh - 1 ? args.prog[t + 1].n_halves : 2) * 2;
        const bool do_store = tile_ok && ((args.store_dys >> t) & 1);
        const int32_t grow0 = int32_t(int64_t(t) * args.n_rows_pad + int64_t(tile) * TILE_M);
        for (int Q = 0; Q < nq; ++Q) {
          mbar_wait_a(wr0 + Q * 8, (cnt >> Q) & 1u);
          if (elect_one()) {
            if (do_store) {
              tma_store_2d_hint_a(&tmDys, sA_u + (2 * Q) * A_CHUNK_BYTES, Q * 128, grow0, pol_s);
              tma_store_2d_hint_a(&tmDys, sA_u + (2 * Q + 1) * A_CHUNK_BYTES, Q * 128 + 64, grow0, pol_s);
              tma_store_commit();
              asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
              if (pending >= 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(rl0 + pending * 8) : "memory");
            } else {
              tma_store_wait_read0();
              if (pending >= 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(rl0 + pending * 8) : "memory");
              asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(rl0 + Q * 8) : "memory");
            }
          }
          pending = do_store ? Q : -1;
          __syncwarp();
          cnt ^= 1u << Q;
        }
      }
    }
    if (elect_one()) {
      tma_store_wait_all0();
      if (pending >= 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(rl0 + pending * 8) : "memory");
    }
    __syncwarp();
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(B_EPI_REGS));
    // ================================================================ epilogue warps (both CTAs)
    const int q = warp & 3;
    const int j = warp >> 2;
    const int tid = threadIdx.x;
    const uint32_t rt = q * 32 + lane;
    const uint32_t tmem_row = tmem_base + (uint32_t(q * 32) << 16);
    const uint32_t ready_leader = mapa_u32(smem_u32(&bar_a_ready[0]), 0);
    sts_f32(smem_a + B_OFF_WLAST + uint32_t(tid) * 4u, __ldg(args.w_last + tid));
    const uint32_t swl_a = smem_a + B_OFF_WLAST, xyz_a = smem_a + B_OFF_XYZ;
    const uint32_t st_row = smem_a + rt * 128u + uint32_t(j >> 1) * A_CHUNK_BYTES;
    const uint32_t st_ps = (uint32_t(j & 1) << 6) ^ ((rt & 7u) << 4);
    const uint32_t st_a0 = st_row + st_ps, st_a1 = st_row + (st_ps ^ 16u), st_a2 = st_row + (st_ps ^ 32u), st_a3 = st_row + (st_ps ^ 48u);
    uint32_t mma_iter = 0;
    uint32_t wpar = 0, wany = 0;        // per quarter: parity of / any hand-overs so far
    const int tiles_alloc = int(args.n_rows_pad / TILE_M);

    for (int tp = pair0; tp < n_pairs; tp += pair_stride) {
      const int tile = 2 * tp + int(rank);
      const bool tile_ok = tile < args.n_tiles;
      const int tile_c = tile_ok ? tile : args.n_tiles - 1;
      const int64_t lrow = int64_t(tile_c) * TILE_M + rt;
      const bool valid = tile_ok && lrow < args.n_rows;
      const float dyv = valid ? __ldg(args.dy + lrow) : 0.0f;
      if (j == 0) {
        const float* pt = args.xyz + lrow * args.xyz_stride;
        sts_u32x4(xyz_a + ((rt & 7u) * 17u + (rt >> 3)) * 16u, __float_as_uint(valid ? __ldg(pt + 0) : 0.0f),
                  __float_as_uint(valid ? __ldg(pt + 1) : 0.0f), __float_as_uint(valid ? __ldg(pt + 2) : 0.0f), 0u);
      }
      bwd_bar();      // s_xyz visible to the column-sum passes of this tile

      // gate words of this row for hidden layer t: word 4*Q + j of [layer][tile][16][128].  Fetched one layer ahead (before
      // the column-sum pass of the layer above), so that their L2 latency is never on the path to the first quarter.
      uint32_t gn0 = 0, gn1 = 0, gn2 = 0, gn3 = 0;
      auto fetch_gates = [&](int t) {
        gn0 = gn1 = gn2 = gn3 = 0;
        if (tile_ok && t >= 0) {
          const uint32_t* mrow = args.bitmask + ((size_t(t) * tiles_alloc + tile_c) * MASK_WORDS + j) * TILE_M + rt;
          gn0 = __ldg(mrow); gn1 = __ldg(mrow + 4 * TILE_M); gn2 = __ldg(mrow + 8 * TILE_M); gn3 = __ldg(mrow + 12 * TILE_M);
        }
      };
      fetch_gates(nh - 1);
      for (int t = nh - 1; t >= 0; --t) {
        const bool has_acc = (t < nh - 1);
        const int n_halves = has_acc ? args.prog[t + 1].n_halves : 2;
        const bool produce_a = (t > 0);
        const uint32_t gw0 = gn0, gw1 = gn1, gw2 = gn2, gw3 = gn3;
        uint32_t par = 0;
        if (has_acc) { par = mma_iter & 1; ++mma_iter; }

        // `free_bar` (or null): the mbarrier that says the MMAs no longer read the dY chunks of this quarter
        auto quarter = [&](int Q, uint32_t gate, uint64_t* free_bar) {
          const uint32_t c0 = uint32_t(Q) * 128u + uint32_t(j) * 32u;
          const uint32_t qoff = uint32_t(Q) * 2u * A_CHUNK_BYTES;
          uint32_t p0[8], p1[8];
          bwd_pass16(tmem_row + c0, has_acc, dyv, swl_a + c0 * 4, gate, p0);
          bwd_pass16(tmem_row + c0 + 16, has_acc, dyv, swl_a + (c0 + 16) * 4, gate >> 16, p1);
          // only the stores wait: for the tensor cores to have read the old chunks, and for the previous dY tile's quarter to
          // have been stashed and column-summed
          if (free_bar != nullptr) mbar_wait(free_bar, par);
          if ((wany >> Q) & 1u) mbar_wait(&bar_q_released[Q], ((wpar >> Q) & 1u) ^ 1u);
          sts_u32x4(st_a0 + qoff, p0[0], p0[1], p0[2], p0[3]);
          sts_u32x4(st_a1 + qoff, p0[4], p0[5], p0[6], p0[7]);
          sts_u32x4(st_a2 + qoff, p1[0], p1[1], p1[2], p1[3]);
          sts_u32x4(st_a3 + qoff, p1[4], p1[5], p1[6], p1[7]);
          fence_proxy_async_smem();
          tc_fence_before();
          __syncwarp();
          if (produce_a && lane == 0) mbar_arrive_remote(ready_leader + 8u * uint32_t(Q));
          if (lane == 0) mbar_arrive(&bar_q_written[Q]);
          wpar ^= 1u << Q; wany |= 1u << Q;
        };

        if (has_acc) { mbar_wait(&bar_tmem_full[0], par); tc_fence_after(); }
        quarter(0, gw0, has_acc ? &bar_free_q0 : nullptr);
        quarter(1, gw1, has_acc ? &bar_free_lo : nullptr);
        if (has_acc) { mbar_wait(&bar_tmem_full[1], par); tc_fence_after(); }
        if (n_halves == 2) {
          quarter(2, gw2, nullptr);
          quarter(3, gw3, nullptr);
        } else {
          tc_fence_before();
          __syncwarp();
          if (produce_a && lane == 0) { mbar_arrive_remote(ready_leader + 16u); mbar_arrive_remote(ready_leader + 24u); }
        }

        fetch_gates(t - 1);      // in flight under the column sums
        // ---- column sums of this warp's 32 columns, once all 16 warps have written that quarter
        {
          // warp w owns columns [32w, 32w+32) = four 16-byte pieces of one 64-column chunk.  lane = (piece, row class r0):
          // it walks the 16 rows r0 + 8i of its piece with LDS.128 at constant offsets (the SW128 XOR depends on r & 7 = r0
          // only; the eight row classes of a piece cover all 32 banks, so the loads are conflict-free), then the eight
          // row classes are folded with a halving butterfly (fixed order).
          const int cw = warp * 32;
          const bool cols_valid = cw < n_halves * 256;
          const bool want_xyz = args.xyzsum_slot[t] >= 0;
          const bool do_sums = want_xyz || args.colsum_slot[t] >= 0;
          const uint32_t r0 = uint32_t(lane) & 7u, pl = uint32_t(lane) >> 3;
          float s[8], sx[8], sy[8], sz[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) { s[k] = 0.f; sx[k] = 0.f; sy[k] = 0.f; sz[k] = 0.f; }
          if (cols_valid) {
            mbar_wait(&bar_q_written[warp >> 2], ((wpar >> (warp >> 2)) & 1u) ^ 1u);
          }
          if (cols_valid && do_sums) {
            const uint32_t piece = uint32_t(warp & 1) * 4u + pl;
            const uint32_t a0 = smem_a + uint32_t(cw >> 6) * A_CHUNK_BYTES + r0 * 128u + ((piece ^ r0) << 4);
            const uint32_t xb = xyz_a + r0 * (17u * 16u);
            if (!want_xyz) {
#pragma unroll
              for (int i = 0; i < TILE_M / 8; ++i) {
                uint32_t w0, w1, w2, w3;
                asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(w0), "=r"(w1), "=r"(w2), "=r"(w3) : "r"(a0 + uint32_t(i) * 1024u));
                s[0] += bf16_lo_to_f32(w0); s[1] += bf16_hi_to_f32(w0); s[2] += bf16_lo_to_f32(w1); s[3] += bf16_hi_to_f32(w1);
                s[4] += bf16_lo_to_f32(w2); s[5] += bf16_hi_to_f32(w2); s[6] += bf16_lo_to_f32(w3); s[7] += bf16_hi_to_f32(w3);
              }
            } else {
              // packed fp32 pairs (FFMA2: one instruction per (lo, hi) column pair and sum - half the FP issue slots of the
              // scalar form, same roundings: fma(v, 1, s) == s + v)
              uint64_t s2[4], sx2[4], sy2[4], sz2[4];
#pragma unroll
              for (int k = 0; k < 4; ++k) { s2[k] = 0ull; sx2[k] = 0ull; sy2[k] = 0ull; sz2[k] = 0ull; }
#pragma unroll 4
              for (int i = 0; i < TILE_M / 8; ++i) {
                uint32_t w[4];
                asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]) : "r"(a0 + uint32_t(i) * 1024u));
                const float4 p = lds_f32x4(xb + uint32_t(i) * 16u);
                const float x = p.x, y = p.y, z = p.z;
                const uint64_t one2 = pack_f32x2(1.0f, 1.0f), x2 = pack_f32x2(x, x), y2 = pack_f32x2(y, y), z2 = pack_f32x2(z, z);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const uint64_t v2 = pack_f32x2(bf16_lo_to_f32(w[k]), bf16_hi_to_f32(w[k]));
                  fma_f32x2(s2[k], v2, one2);
                  fma_f32x2(sx2[k], v2, x2);
                  fma_f32x2(sy2[k], v2, y2);
                  fma_f32x2(sz2[k], v2, z2);
                }
              }
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                unpack_f32x2(s2[k], s[2 * k], s[2 * k + 1]);
                unpack_f32x2(sx2[k], sx[2 * k], sx[2 * k + 1]);
                unpack_f32x2(sy2[k], sy[2 * k], sy[2 * k + 1]);
                unpack_f32x2(sz2[k], sz[2 * k], sz[2 * k + 1]);
              }
            }
          }
          if (cols_valid) { __syncwarp(); if (lane == 0) mbar_arrive(&bar_q_released[warp >> 2]); }
          // fold the eight row classes: after the three steps lane (piece, r0) holds column 8*piece + 4*b0 + 2*b1 + b2 of r0
          auto fold8 = [&](float (&v)[8]) -> float {
            const bool b0 = lane & 1, b1 = lane & 2, b2 = lane & 4;
            float t4[4], t2[2];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float keep = b0 ? v[k + 4] : v[k], send = b0 ? v[k] : v[k + 4];
              t4[k] = keep + __shfl_xor_sync(0xffffffffu, send, 1);
            }
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              const float keep = b1 ? t4[k + 2] : t4[k], send = b1 ? t4[k] : t4[k + 2];
              t2[k] = keep + __shfl_xor_sync(0xffffffffu, send, 2);
            }
            const float keep = b2 ? t2[1] : t2[0], send = b2 ? t2[0] : t2[1];
            return keep + __shfl_xor_sync(0xffffffffu, send, 4);
          };
          const float cs = fold8(s);
          float cx = 0.f, cy = 0.f, cz = 0.f;
          if (want_xyz) { cx = fold8(sx); cy = fold8(sy); cz = fold8(sz); }
          if (tile_ok && cols_valid && do_sums) {
            float* part = args.tile_part + size_t(tile) * args.part_stride;
            const int col = cw + int(pl) * 8 + ((lane & 1) << 2) + (lane & 2) + ((lane & 4) >> 2);
            if (args.colsum_slot[t] >= 0) part[args.colsum_slot[t] * FH + col] = cs;
            if (want_xyz) {
              const int sl = args.xyzsum_slot[t];
              part[(sl + 0) * FH + col] = cx;
              part[(sl + 1) * FH + col] = cy;
              part[(sl + 2) * FH + col] = cz;
            }
          }
        }
      }
      // s_xyz of this tile must not be overwritten while a slower warp is still in the last column-sum pass
      bwd_bar();
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == B_EPI_WARPS + 1) tmem_dealloc_pair<512>(tmem_base);
}

int launch_chain_dgrad(const CUtensorMap& tmW, const CUtensorMap& tmDys, const BwdArgs& args, int n_sms, cudaStream_t stream) {
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(chain_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, B_SMEM));
  static int max_pairs = 0;
  if (max_pairs == 0) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(unsigned(n_sms & ~1));
    cfg.blockDim = dim3(B_THREADS);
    cfg.dynamicSmemBytes = B_SMEM;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, chain_dgrad_kernel, &cfg) != cudaSuccess || n <= 0) { cudaGetLastError(); n = n_sms / 2; }
    max_pairs = n;
  }
  const int n_pairs = (args.n_tiles + 1) / 2;
  int grid_pairs = max_pairs < n_pairs ? max_pairs : n_pairs;
  if (grid_pairs < 1) grid_pairs = 1;
  DSDF_LAUNCHED;
  chain_dgrad_kernel<<<2 * grid_pairs, B_THREADS, B_SMEM, stream>>>(tmW, tmDys, args);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace dsdf
