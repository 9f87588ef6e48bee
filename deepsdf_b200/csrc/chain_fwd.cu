// Fused DeepSDF decoder forward, CTA-pair version (sm_100a: tcgen05 cta_group::2 + TMEM + TMA, 2-CTA clusters).
//
// Two CTAs on the two SMs of a TPC form a pair.  Each CTA owns one 128-row point tile (its activation tile [128 x 512]
// bf16 in its own shared memory, its accumulators [128 x 512] fp32 in its own TMEM) and the pair shares every weight
// tile: for a [256 out x 64 K] weight block each CTA TMA-loads only its 128 output rows, and one tcgen05.mma
// (cta_group::2, M = 256 over the pair, N = 256, K = 16) issued by the leader CTA multiplies both activation tiles with
// both halves.  That halves the L2 -> SM weight traffic and the shared-memory space of the weight ring per row, which
// is what makes room for a 5-stage ring, bias / xyz-weight staging and 16 epilogue warps.
//
// Per layer, per pair:   MMA order  (half 0: K chunks 0..7), (half 1: K chunks 0..7)
//   * tmem_full[0] (multicast commit) -> both CTAs' epilogues post-process accumulator half 0 into REGISTERS while the
//     tensor cores compute half 1;  free_lo (commit after half 1's K chunk 3) says activation chunks 0..3 are no longer
//     read, so half 0 is written into the A tile before the layer's MMAs have even finished;
//   * tmem_full[1] -> half 1 is post-processed and written quarter by quarter (128 columns); the next layer's MMAs start
//     on the K chunks that are already there (a_ready[0..3], one mbarrier per quarter, arrivals from both CTAs).
// Epilogue per element: FFMA (bias and dropout scale folded, bias staged in shared memory), sign-bit funnel -> ReLU
// gate word, AND with the bit-sliced Philox keep word (computed before the accumulators are ready), select, bf16 pack.
// The final 512 -> 1 layer, tanh, clamp, L1 loss and dL/dy are fused into the last epilogue; in training every layer's
// activation tile is TMA-stored (bf16) for wgrad together with its 1-bit gates.
// Replaces Decoder.forward (deep_sdf/networks/deep_sdf_decoder.py:76-111) and train_deep_sdf.py:511-519.
#include <type_traits>

#include "common.cuh"
#include "fused.cuh"
#include "host_util.h"

namespace dsdf {

namespace {

constexpr int P_EPI_WARPS = 16;
constexpr int P_EPI_THREADS = P_EPI_WARPS * 32;        // 512
constexpr int P_THREADS = (P_EPI_WARPS + 4) * 32;      // + TMA warp + MMA warp + stash warp + one idle warp that completes the
                                                       //   service warpgroup (setmaxnreg works on whole warpgroups)
constexpr int P_EPI_REGS = 112, P_SVC_REGS = 24;       // register re-split: 16 x 112 + 4 x 24 <= 20 x 96 per thread-slot
constexpr int P_STAGES = 5;
constexpr int P_STAGE_BYTES = 128 * 128;               // this CTA's half ([128 out rows][64 K] bf16) of a weight block
constexpr int P_OFF_B = A_BYTES;
constexpr int P_OFF_BIAS = P_OFF_B + P_STAGES * P_STAGE_BYTES;   // f32 [2][512]  (double-buffered per layer)
constexpr int P_OFF_WXYZ = P_OFF_BIAS + 2 * FH * 4;              // f32 [3][512]
constexpr int P_OFF_WLAST = P_OFF_WXYZ + 3 * FH * 4;             // f32 [512]
constexpr int P_OFF_YPART = P_OFF_WLAST + FH * 4;                // f32 [4][128]
constexpr int P_SMEM = P_OFF_YPART + 4 * TILE_M * 4 + 1024;

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(P_EPI_THREADS) : "memory"); }

// Post-process 16 consecutive columns [c0, c0+16) of one row and write them (bf16) into the SW128 activation tile.
//   has_acc  read the fp32 accumulators from TMEM (taddr) - false for the MMA-less first layer
//   st0/st1  shared addresses of the two 16-byte pieces of the activation tile that receive the 16 columns
//   sb       shared-memory bias (already multiplied by the dropout scale); swx xyz weight rows (same scaling) or null
//   keep16   low 16 bits: Philox keep bits of these columns (all ones when the layer has no dropout)
// Returns the 16 gate bits (relu AND keep); adds this row's share of the last-layer dot product to ydot.
// kPlain: the common hidden layer (accumulators present, no per-row xyz / latent columns, not the last layer) as a
// compile-time variant, so that its passes carry no flag handling (R2P takes all seven predicates; layer-invariant
// conditions kept as run-time flags end up in a bit-field register that every pass re-tests).
template <bool kTrain, bool kRowLat, bool kPlain>
__device__ __forceinline__ uint32_t epi_pass16(uint32_t taddr, bool has_acc_rt, float ascale, uint32_t sb, bool has_xyz_rt, uint32_t swx,
                                               float x, float y, float z, uint32_t keep16, bool is_last_rt, uint32_t swl,
                                               uint32_t st0, uint32_t st1, float& ydot, const float* wl, int n_rl,
                                               const float (&rl)[4]) {
  // sb / swx / swl / sA are SHARED-space byte addresses (of this pass's first column where they are per column)
  const bool has_acc = kPlain ? true : has_acc_rt, has_xyz = kPlain ? false : has_xyz_rt, is_last = kPlain ? false : is_last_rt;
  float v[16];
  if (has_acc) {
    uint32_t r[16];
    tmem_ld16(taddr, r);
    tmem_ld_wait();
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4 b = lds_f32x4(sb + 16 * k);
      v[4 * k + 0] = fmaf(__uint_as_float(r[4 * k + 0]), ascale, b.x);
      v[4 * k + 1] = fmaf(__uint_as_float(r[4 * k + 1]), ascale, b.y);
      v[4 * k + 2] = fmaf(__uint_as_float(r[4 * k + 2]), ascale, b.z);
      v[4 * k + 3] = fmaf(__uint_as_float(r[4 * k + 3]), ascale, b.w);
    }
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4 b = lds_f32x4(sb + 16 * k);
      v[4 * k + 0] = b.x; v[4 * k + 1] = b.y; v[4 * k + 2] = b.z; v[4 * k + 3] = b.w;
    }
  }
  if (has_xyz) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4 wx = lds_f32x4(swx + 16 * k);
      const float4 wy = lds_f32x4(swx + FH * 4 + 16 * k);
      const float4 wz = lds_f32x4(swx + 2 * FH * 4 + 16 * k);
      v[4 * k + 0] += x * wx.x + y * wy.x + z * wz.x;
      v[4 * k + 1] += x * wx.y + y * wy.y + z * wz.y;
      v[4 * k + 2] += x * wx.z + y * wy.z + z * wz.z;
      v[4 * k + 3] += x * wx.w + y * wy.w + z * wz.w;
    }
  }
  if (kRowLat && wl != nullptr) {
    // row-latent mode: this row's own code times the latent weight columns (fp32, read through L1: the mode serves a few
    // hundred thousand rows per call, shared memory has no room for the table, and only two layers come here)
#pragma unroll
    for (int d = 0; d < 4; ++d) {
      if (d < n_rl) {
        const float s = rl[d] * ascale;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float4 w = __ldg(reinterpret_cast<const float4*>(wl + size_t(d) * FH) + k);
          v[4 * k + 0] += s * w.x; v[4 * k + 1] += s * w.y; v[4 * k + 2] += s * w.z; v[4 * k + 3] += s * w.w;
        }
      }
    }
  }
  uint32_t gate = 0;
  if (kTrain) {
    uint32_t sg = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) sg = __funnelshift_l(__float_as_uint(v[i]), sg, 1);   // bit (15-i) = sign of v[i]
    gate = (~(__brev(sg) >> 16)) & keep16 & 0xFFFFu;
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = (gate & (1u << i)) ? v[i] : 0.0f;
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.0f);
  }
  if (is_last) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4 w = lds_f32x4(swl + 16 * k);
      ydot += v[4 * k + 0] * w.x + v[4 * k + 1] * w.y + v[4 * k + 2] * w.z + v[4 * k + 3] * w.w;
    }
  }
  // st0 / st1: the two 16-byte pieces of this row's SW128 line that hold these 16 columns (addresses formed once per thread)
  sts_u32x4(st0, pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
  sts_u32x4(st1, pack_bf16x2(v[8], v[9]), pack_bf16x2(v[10], v[11]), pack_bf16x2(v[12], v[13]), pack_bf16x2(v[14], v[15]));
  return gate;
}

}  // namespace

template <bool kTrain, bool kRowLat>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(P_THREADS, 1)
chain_forward_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmActs,
                     const __grid_constant__ FwdArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + P_OFF_B;
  // the same base as a 32-bit shared-space address: everything the epilogue touches is addressed from it with explicit
  // ld/st.shared, so no generic pointer (and no re-derivation of its alignment) lives in the hot loop
  const uint32_t smem_a = (smem_u32(smem_raw) + 1023u) & ~1023u;
  __shared__ uint64_t bar_full[P_STAGES], bar_empty[P_STAGES], bar_tmem_full[2], bar_free_q0, bar_free_lo, bar_a_ready[4];
  __shared__ uint64_t bar_q_written[4], bar_q_stashed[4];   // activation stash hand-over (training), per 128-column quarter
  __shared__ uint32_t tmem_slot;
  __shared__ float s_loss[4];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if (threadIdx.x == 0) {
    for (int s = 0; s < P_STAGES; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
    mbar_init(&bar_tmem_full[0], 1); mbar_init(&bar_tmem_full[1], 1);
    mbar_init(&bar_free_q0, 1);
    mbar_init(&bar_free_lo, 1);
    for (int i = 0; i < 4; ++i) mbar_init(&bar_a_ready[i], 2 * P_EPI_WARPS);
    for (int i = 0; i < 4; ++i) { mbar_init(&bar_q_written[i], P_EPI_WARPS); mbar_init(&bar_q_stashed[i], 1); }
    fence_barrier_init();
  }
  if (warp == P_EPI_WARPS + 1) tmem_alloc_pair<512>(&tmem_slot);
  if (warp == P_EPI_WARPS && lane == 0) { tma_prefetch_desc(&tmW); tma_prefetch_desc(&tmActs); }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int nh = args.num_hidden;
  const int n_pairs = (args.n_tiles + 1) >> 1;
  const int pair0 = blockIdx.x >> 1, pair_stride = gridDim.x >> 1;

  // Register re-split (each role issues its own setmaxnreg first): the kernel launches with 96 registers per thread (20
  // warps), the four single-lane service warps keep 24, the sixteen epilogue warps get 112.
  if (warp == P_EPI_WARPS) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(P_SVC_REGS));
    // ================================================================ TMA producer (this CTA's half of every weight block)
    {
      // the whole warp walks the loop (warp-uniform values stay in uniform registers); one elected lane issues
      const uint32_t full_leader = mapa_u32(smem_u32(&bar_full[0]), 0);
      const uint32_t full0 = smem_u32(&bar_full[0]), empty0 = smem_u32(&bar_empty[0]), sB_u = smem_u32(sB);
      uint32_t stage = 0, phase = 0;
      const uint64_t pol_w = l2_policy_evict_last();    // 3.4 MB of weights, re-read by every tile
      for (int tp = pair0; tp < n_pairs; tp += pair_stride) {
        for (int l = 0; l < nh; ++l) {
          if (!args.prog[l].has_mma) continue;
          const int n_halves = args.prog[l].n_halves, k_chunks = args.prog[l].k_chunks;
          const int row_base = args.prog[l].w_row0 + int(rank) * 128;
          for (int h = 0; h < n_halves; ++h)
            for (int kc = 0; kc < k_chunks; ++kc) {
              mbar_wait_a(empty0 + stage * 8, phase ^ 1);
              if (elect_one()) {
                if (leader) mbar_arrive_expect_tx_a(full0 + stage * 8, 2 * P_STAGE_BYTES);
                tma_load_2d_pair_hint_a(sB_u + stage * P_STAGE_BYTES, &tmW, full_leader + stage * 8, kc * 64, row_base + h * HALF_N, pol_w);
              }
              __syncwarp();
              if (++stage == P_STAGES) { stage = 0; phase ^= 1; }
            }
        }
      }
    }
  } else if (warp == P_EPI_WARPS + 1) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(P_SVC_REGS));
    // ================================================================ MMA issuer (leader CTA).  The whole warp walks the
    // loop so that every operand is warp-uniform (uniform registers feed UTCHMMA directly; issuing from a divergent
    // `lane == 0` branch makes the compiler wrap each MMA in an ELECT / R2UR waterfall that costs more than the 128
    // tensor-core cycles the MMA itself takes); one elected lane issues.  Barrier addresses, descriptor bases and the
    // layer fields are precomputed: the warp must issue four MMAs (= 512 tensor cycles) per stage without being the bottleneck.
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
        for (int l = 0; l < nh; ++l) {
          if (!args.prog[l].has_mma) continue;
          const int n_halves = args.prog[l].n_halves, k_chunks = args.prog[l].k_chunks;
          const uint32_t apar = a_cons & 1; ++a_cons;
          const int kc_lo_last = k_chunks < 4 ? k_chunks - 1 : 3;
          const int kc_q0_last = k_chunks < 2 ? k_chunks - 1 : 1;
          uint32_t waited = 0;     // quarters of the previous layer's output whose a_ready has been consumed
          for (int h = 0; h < n_halves; ++h) {
            const bool last_half = (h == n_halves - 1);
            for (int kc = 0; kc < k_chunks; ++kc) {
              // K chunk kc reads quarter kc/2 of the A tile.  Besides, the first MMA of half h overwrites that half's
              // accumulators, which the previous layer's epilogue reads quarter by quarter: half 0 needs quarters 0,1
              // drained, half 1 needs quarters 2,3.
              uint32_t needm = 1u << (kc >> 1);
              if (kc == 0) needm |= 3u << (2 * h);
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
              const uint64_t bd = b_base + uint64_t(stage * (P_STAGE_BYTES >> 4));
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
              if (++stage == P_STAGES) { stage = 0; phase ^= 1; }
            }
          }
          // keep the per-quarter barriers in lock step when this layer's K did not need all of them
          uint32_t rest = 0xFu & ~waited;
          while (rest) {
            const int qd = __ffs(int(rest)) - 1;
            mbar_wait_a(ready0 + qd * 8, apar);
            rest &= rest - 1;
          }
        }
      }
    }
  } else if (warp == P_EPI_WARPS + 3) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(P_SVC_REGS));      // idle warp: completes the service warpgroup
  } else if (warp == P_EPI_WARPS + 2) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(P_SVC_REGS));
    // ================================================================ stash warp (training): as soon as all 16 epilogue
    // warps have written a 128-column quarter of the activation tile, TMA-store its two 64-column chunks (bf16) for
    // wgrad and tell the epilogue when the store has read the shared memory, so that the quarter may be overwritten by
    // the next layer.  No CTA-wide barrier is involved: the epilogue warps never wait for each other because of the stash.
    if (kTrain && args.store_acts) {
      const uint32_t wr0 = smem_u32(&bar_q_written[0]), st0 = smem_u32(&bar_q_stashed[0]), sA_u = smem_u32(sA);
      uint32_t cnt = 0;            // bit Q: parity of the number of quarter-Q hand-overs so far
      int pending = -1;            // quarter whose store is still in flight (its q_stashed arrive is owed)
      const uint64_t pol_s = l2_policy_evict_first();   // the stash streams out: it must not push the weight blocks out of L2
      for (int tp = pair0; tp < n_pairs; tp += pair_stride) {
        const int tile = 2 * tp + int(rank);
        const bool tile_ok = tile < args.n_tiles;
        for (int l = 0; l < nh; ++l) {
          const int nq = args.prog[l].n_halves * 2;
          const int32_t grow0 = int32_t(int64_t(l) * args.n_rows_pad + int64_t(tile) * TILE_M);
          for (int Q = 0; Q < nq; ++Q) {
            mbar_wait_a(wr0 + Q * 8, (cnt >> Q) & 1u);
            if (elect_one()) {
              if (tile_ok) {
                tma_store_2d_hint_a(&tmActs, sA_u + (2 * Q) * A_CHUNK_BYTES, Q * 128, grow0, pol_s);
                tma_store_2d_hint_a(&tmActs, sA_u + (2 * Q + 1) * A_CHUNK_BYTES, Q * 128 + 64, grow0, pol_s);
                tma_store_commit();
                // keep one store in flight: everything but the group just committed has been read out of shared memory
                asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                if (pending >= 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(st0 + pending * 8) : "memory");
              } else {
                tma_store_wait_read0();
                if (pending >= 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(st0 + pending * 8) : "memory");
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(st0 + Q * 8) : "memory");
              }
            }
            pending = tile_ok ? Q : -1;
            __syncwarp();
            cnt ^= 1u << Q;
          }
        }
      }
      if (elect_one()) {
        tma_store_wait_all0();
        if (pending >= 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(st0 + pending * 8) : "memory");
      }
      __syncwarp();
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(P_EPI_REGS));
    // ================================================================ epilogue warps (both CTAs)
    const int q = warp & 3;            // TMEM lane quadrant
    const int j = warp >> 2;           // 32-column slice of every 128-column quarter
    const int tid = threadIdx.x;       // 0..511
    const uint32_t rt = q * 32 + lane; // row within the tile == TMEM lane
    const uint32_t tmem_row = tmem_base + (uint32_t(q * 32) << 16);
    const uint32_t ready_leader = mapa_u32(smem_u32(&bar_a_ready[0]), 0);   // + 8*quarter
    sts_f32(smem_a + P_OFF_WLAST + uint32_t(tid) * 4u, __ldg(args.w_last + tid));
    // this thread's store addresses inside a quarter: row line + 64-column chunk of its slice + swizzled 16-byte piece
    const uint32_t st_row = smem_a + rt * 128u + uint32_t(j >> 1) * A_CHUNK_BYTES;
    const uint32_t st_ps = (uint32_t(j & 1) << 6) ^ ((rt & 7u) << 4);
    const uint32_t st_a0 = st_row + st_ps, st_a1 = st_row + (st_ps ^ 16u), st_a2 = st_row + (st_ps ^ 32u), st_a3 = st_row + (st_ps ^ 48u);
    uint32_t mma_iter = 0;
    const bool stash = kTrain && args.store_acts != 0;
    uint32_t wpar = 0, wany = 0;        // per quarter: parity of / any hand-overs to the stash warp so far
    const int tiles_alloc = int(args.n_rows_pad / TILE_M);

    for (int tp = pair0; tp < n_pairs; tp += pair_stride) {
      const int tile = 2 * tp + int(rank);
      const bool tile_ok = tile < args.n_tiles;
      const int tile_c = tile_ok ? tile : args.n_tiles - 1;
      const bool valid = tile_ok && (int64_t(tile_c) * TILE_M + rt) < args.n_rows &&
                         (args.tile_rows == nullptr || int(rt) < __ldg(args.tile_rows + tile_c));
      float px = 0.f, py = 0.f, pz = 0.f;
      if (valid) {
        const float* pt = args.xyz + (int64_t(tile_c) * TILE_M + rt) * args.xyz_stride;
        px = __ldg(pt); py = __ldg(pt + 1); pz = __ldg(pt + 2);
      }
      const int scene = kRowLat ? 0 : __ldg(args.tile_scene + tile_c);
      float ydot = 0.0f;
      float rl[4] = {0.f, 0.f, 0.f, 0.f};
      if (kRowLat && valid) {
        const float* pl = args.row_lat + (int64_t(tile_c) * TILE_M + rt) * args.row_lat_stride;
#pragma unroll
        for (int d = 0; d < 4; ++d) if (d < args.row_lat_dim) rl[d] = __ldg(pl + d);
      }

      for (int l = 0; l < nh; ++l) {
        const ChainLayer& L = args.prog[l];
        const bool last = (l == nh - 1);
        const bool produce_a = !last;
        const bool drop = kTrain && L.dropout != 0 && args.dropout_on != 0;
        const bool has_acc = L.has_mma != 0;
        const float ascale = drop ? args.drop_scale : 1.0f;
        const uint32_t sb = smem_a + P_OFF_BIAS + uint32_t(l & 1) * FH * 4;
        // ---- before the accumulators are ready: stage this layer's bias (+ xyz weights), draw the first keep words
        {
          const float* src = (L.bias_slot >= 0 && !kRowLat) ? (args.cbias + (size_t(scene) * 2 + L.bias_slot) * FH) : args.bias[l];
          sts_f32(sb + uint32_t(tid) * 4u, __ldg(src + tid) * ascale);
          if (L.has_xyz) {
            const float* wsrc = args.wxyz[l];
            const uint32_t wx = smem_a + P_OFF_WXYZ + uint32_t(tid) * 4u;
            sts_f32(wx, __ldg(wsrc + tid) * ascale);
            sts_f32(wx + FH * 4, __ldg(wsrc + FH + tid) * ascale);
            sts_f32(wx + 2 * FH * 4, __ldg(wsrc + 2 * FH + tid) * ascale);
          }
        }
        const uint64_t grow = args.row0 + uint64_t(int64_t(tile_c) * TILE_M + rt);
        uint32_t kwa = 0xFFFFFFFFu, kwb = 0xFFFFFFFFu;
        if (drop) {
          kwa = dropout_keep_word_rk(args.philox_rk, uint32_t(l), grow, uint32_t(j), args.drop_thr);
          kwb = dropout_keep_word_rk(args.philox_rk, uint32_t(l), grow, uint32_t(4 + j), args.drop_thr);
        }
        epi_bar();
        const bool has_xyz = L.has_xyz != 0;
        const uint32_t swx = smem_a + P_OFF_WXYZ, swl = smem_a + P_OFF_WLAST;
        uint32_t par = 0;
        if (has_acc) { par = mma_iter & 1; ++mma_iter; }
        uint32_t* gdst = nullptr;
        if (kTrain && args.bitmask != nullptr && tile_ok)
          gdst = args.bitmask + ((size_t(l) * tiles_alloc + tile) * MASK_WORDS + j) * TILE_M + rt;   // word 4*Q + j at + 4*Q*128

        // One 128-column quarter: two 16-column passes straight from TMEM into the A tile.
        const float* wl_layer = (kRowLat && L.bias_slot >= 0) ? args.wlat + size_t(L.bias_slot) * args.row_lat_dim * FH : nullptr;
        auto quarter_v = [&](int Q, uint32_t keep, auto PLAIN) {
          constexpr bool kPlain = decltype(PLAIN)::value;
          const uint32_t c0 = uint32_t(Q) * 128u + uint32_t(j) * 32u;
          // the previous layer's stash of this quarter must have been read out of the tile
          if (stash && ((wany >> Q) & 1u)) mbar_wait(&bar_q_stashed[Q], ((wpar >> Q) & 1u) ^ 1u);
          const uint32_t qoff = uint32_t(Q) * 2u * A_CHUNK_BYTES;        // a quarter = two 64-column chunks
          const uint32_t g0 = epi_pass16<kTrain, kRowLat, kPlain>(tmem_row + c0, has_acc, ascale, sb + c0 * 4, has_xyz, swx + c0 * 4, px,
                                                                  py, pz, keep, last, swl + c0 * 4, st_a0 + qoff, st_a1 + qoff, ydot,
                                                                  wl_layer ? wl_layer + c0 : nullptr, args.row_lat_dim, rl);
          const uint32_t g1 = epi_pass16<kTrain, kRowLat, kPlain>(tmem_row + c0 + 16, has_acc, ascale, sb + (c0 + 16) * 4, has_xyz,
                                                                  swx + (c0 + 16) * 4, px, py, pz, keep >> 16, last, swl + (c0 + 16) * 4,
                                                                  st_a2 + qoff, st_a3 + qoff, ydot,
                                                                  wl_layer ? wl_layer + c0 + 16 : nullptr, args.row_lat_dim, rl);
          if (kTrain && gdst != nullptr) gdst[size_t(Q) * 4 * TILE_M] = g0 | (g1 << 16);
          fence_proxy_async_smem();
          tc_fence_before();
          __syncwarp();
          if (produce_a && lane == 0) mbar_arrive_remote(ready_leader + 8u * uint32_t(Q));
          if (stash) {
            if (lane == 0) mbar_arrive(&bar_q_written[Q]);
            wpar ^= 1u << Q; wany |= 1u << Q;
          }
        };
        // one run-time dispatch per quarter instead of flag tests in every 16-column pass
        const bool plain = has_acc && !has_xyz && !last;
        auto quarter = [&](int Q, uint32_t keep) {
          if (plain) quarter_v(Q, keep, std::true_type{});
          else quarter_v(Q, keep, std::false_type{});
        };

        // ---- half 0: ready while the tensor cores still compute half 1.  Activation chunks 0,1 may be overwritten once
        //      half 1's MMAs have passed K chunk 1 (free_q0), chunks 2,3 after K chunk 3 (free_lo); the previous layer's
        //      stash (TMA store) must have been read out of the tile.
        if (has_acc) { mbar_wait(&bar_tmem_full[0], par); tc_fence_after(); mbar_wait(&bar_free_q0, par); }
        quarter(0, kwa);
        if (has_acc) mbar_wait(&bar_free_lo, par);
        quarter(1, kwb);
        // ---- half 1: after all of this layer's MMAs; the next layer starts on the K chunks that are already there
        if (drop && L.n_halves == 2) {
          kwa = dropout_keep_word_rk(args.philox_rk, uint32_t(l), grow, uint32_t(8 + j), args.drop_thr);
          kwb = dropout_keep_word_rk(args.philox_rk, uint32_t(l), grow, uint32_t(12 + j), args.drop_thr);
        }
        if (has_acc) { mbar_wait(&bar_tmem_full[1], par); tc_fence_after(); }
        if (L.n_halves == 2) {
          quarter(2, kwa);
          quarter(3, kwb);
        } else {
          if (kTrain && gdst != nullptr) { gdst[size_t(2) * 4 * TILE_M] = 0u; gdst[size_t(3) * 4 * TILE_M] = 0u; }
          tc_fence_before();
          __syncwarp();
          if (produce_a && lane == 0) { mbar_arrive_remote(ready_leader + 16u); mbar_arrive_remote(ready_leader + 24u); }
        }
      }

      const int64_t lrow = int64_t(tile_c) * TILE_M + rt;
      // ---- final layer: combine the four column slices, tanh, loss
      sts_f32(smem_a + P_OFF_YPART + (uint32_t(j) * TILE_M + rt) * 4u, ydot);
      epi_bar();
      if (j == 0) {
        const uint32_t yp = smem_a + P_OFF_YPART + rt * 4u;
        const float yv = ((lds_f32(yp) + lds_f32(yp + TILE_M * 4)) + (lds_f32(yp + 2 * TILE_M * 4) + lds_f32(yp + 3 * TILE_M * 4))) + __ldg(args.b_last);
        const float t1 = tanhf(yv);
        float t = t1, dt = 1.0f - t1 * t1;
        if (args.use_tanh) { t = tanhf(t1); dt *= (1.0f - t * t); }
        if (valid) args.sdf_out[lrow] = t;
        if (kTrain && args.gt != nullptr) {
          const float g = valid ? __ldg(args.gt + lrow * args.gt_stride) : 0.0f;
          float pc = t, gc = g, inside = 1.0f;
          if (args.clamp > 0.0f) {
            pc = fminf(fmaxf(t, -args.clamp), args.clamp);
            gc = fminf(fmaxf(g, -args.clamp), args.clamp);
            inside = (t >= -args.clamp && t <= args.clamp) ? 1.0f : 0.0f;
          }
          const float diff = pc - gc;
          const float sgn = (diff > 0.0f) ? 1.0f : ((diff < 0.0f) ? -1.0f : 0.0f);
          if (tile_ok) args.dy_out[lrow] = valid ? sgn * inside * dt * args.inv_n : 0.0f;
          float a = valid ? fabsf(diff) : 0.0f;
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
          if (lane == 0) s_loss[q] = a;
        } else if (kTrain && args.dy_unit != 0) {
          if (tile_ok) args.dy_out[lrow] = valid ? dt : 0.0f;
        }
      }
      epi_bar();
      if (kTrain && args.gt != nullptr && tid == 0 && tile_ok)
        args.tile_loss[tile] = ((s_loss[0] + s_loss[1]) + s_loss[2]) + s_loss[3];
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == P_EPI_WARPS + 1) tmem_dealloc_pair<512>(tmem_base);
}

// Number of CTA pairs that can be co-resident (one CTA per SM, both CTAs of a pair on one TPC); cached per kernel.
template <typename K>
static int resident_pairs(K kern, int n_sms) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(n_sms & ~1));
  cfg.blockDim = dim3(P_THREADS);
  cfg.dynamicSmemBytes = P_SMEM;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) { cudaGetLastError(); n = n_sms / 2; }
  return n;
}

int launch_chain_forward(const CUtensorMap& tmW, const CUtensorMap& tmActs, const FwdArgs& args, int n_sms, cudaStream_t stream) {
  const bool rows = args.row_lat != nullptr;
  auto kern = rows ? (args.training ? chain_forward_kernel<true, true> : chain_forward_kernel<false, true>)
                   : (args.training ? chain_forward_kernel<true, false> : chain_forward_kernel<false, false>);
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, P_SMEM));
  static int max_pairs[4] = {0, 0, 0, 0};
  int& mp = max_pairs[(args.training ? 1 : 0) + (rows ? 2 : 0)];
  if (mp == 0) mp = resident_pairs(kern, n_sms);
  const int n_pairs = (args.n_tiles + 1) / 2;
  int grid_pairs = mp;
  if (grid_pairs > n_pairs) grid_pairs = n_pairs;
  if (grid_pairs < 1) grid_pairs = 1;
  kern<<<2 * grid_pairs, P_THREADS, P_SMEM, stream>>>(tmW, tmActs, args);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int chain_forward_resident_pairs(int n_sms) {
  cudaFuncSetAttribute(chain_forward_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, P_SMEM);
  return resident_pairs(chain_forward_kernel<true, false>, n_sms);
}

}  // namespace dsdf
