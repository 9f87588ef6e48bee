// Fused DeepSDF decoder forward, CTA-pair version (sm_100a: tcgen05 cta_group::2 + TMEM + TMA, 2-CTA clusters).
//
// Two CTAs on the two SMs of a TPC form a pair.  Each CTA owns one 128-row point tile (its activation tile [128 x 512]
// bf16 in its own shared memory, its accumulators [128 x 512] fp32 in its own TMEM) and the pair shares every weight
// tile: for a [256 out x 64 K] weight block each CTA TMA-loads only its 128 output rows, and one tcgen05.mma
// (cta_group::2, M = 256 over the pair, N = 256, K = 16) issued by the leader CTA multiplies both activation tiles with
// both halves.  That halves the L2 -> SM weight traffic and the shared-memory space of the weight ring per row.
//
// Everything that is NOT a hidden-activation product enters through one extra K = 16 "aux" step per accumulator half
// (fused.cuh): the layer bias, the per-scene bias of lin0 / the latent_in layer (bias + latent columns . z, one vector per
// scene), the xyz columns and - in row-latent mode - the per-row latent columns.  Their values are split hi + lo into
// bf16 pairs so the products keep ~16 mantissa bits.  The aux A block [128 x 32] is written once per tile by the
// epilogue threads, the aux B blocks ride in the weight ring.  So lin0 is a (tiny) tensor-core layer like the others, every
// layer's epilogue is the same code, and the dropout scale 1/(1-p) is folded into the NEXT layer's packed weights
// (fused_host.cu), which leaves the epilogue without a single multiply-add:
//   TMEM load -> sign-bit funnel -> ReLU gate word AND bit-sliced Philox keep word -> select -> bf16 pack -> STS.128
//
// Per layer, per pair:   MMA order  (half 0: aux, K chunks 0..7), (half 1: aux, K chunks 0..7)
//   * tmem_full[0] (multicast commit) -> both CTAs' epilogues post-process accumulator half 0 while the tensor cores
//     compute half 1;  free_q0 / free_lo (commits after half 1's K chunks 1 / 3) say activation chunks 0..1 / 2..3 are no
//     longer read, so half 0 is written into the A tile before the layer's MMAs have even finished;
//   * tmem_full[1] -> half 1 is post-processed and written quarter by quarter (128 columns); the next layer's MMAs start
//     on the K chunks that are already there (a_ready[0..3], one mbarrier per quarter, arrivals from both CTAs).
// The final 512 -> 1 layer, tanh, clamp, L1 loss and dL/dy are fused into the last epilogue; in training every layer's
// activation tile is TMA-stored (bf16, UNSCALED: relu(h) * keep) for wgrad together with its 1-bit gates.
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
constexpr int P_OFF_AUX = P_OFF_B + P_STAGES * P_STAGE_BYTES;    // aux A block: 4 k-cores x [128 rows][8 bf16]
constexpr int P_OFF_WLAST = P_OFF_AUX + 4 * AUX_CORE_BYTES;      // f32 [512]
constexpr int P_OFF_YPART = P_OFF_WLAST + FH * 4;                // f32 [4][128]
constexpr int P_SMEM = P_OFF_YPART + 4 * TILE_M * 4 + 1024;

// Timeline instrumentation (build with -DDSDF_TRACE; never in the shipped library - tools/trace_fwd.py): SM clock at the
// hand-over points of the 4th tile of CTA pair 0, read back with dsdf_debug_trace.  [role 0: epilogue warp 0, 1: MMA warp (2 per
// step: a_ready seen, B stage full), 2: epilogue warp 15, 3: TMA warp (stage empty seen, per step), 4: epilogue warp 0 of CTA 1]
#ifdef DSDF_TRACE
__device__ unsigned long long g_trace[6][16][40];
#define TR(role, l, e) do { if (trace_on) g_trace[role][l][e] = clock64(); } while (0)
#else
#define TR(role, l, e) do { } while (0)
#endif

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(P_EPI_THREADS) : "memory"); }

// Post-process 16 consecutive columns [c0, c0+16) of one row and write them (bf16) into the SW128 activation tile.
//   taddr    TMEM address of the 16 fp32 accumulators (bias, xyz and latent parts already inside: aux step)
//   keep16   low 16 bits: Philox keep bits of these columns (all ones when the layer has no dropout)
//   pk       the 16 results, packed bf16 pairs; the caller stores them (two 16-byte pieces of the SW128 activation tile)
//            once the tile chunk may be overwritten - everything up to there overlaps that wait
//   kLast    last hidden layer: also accumulate this row's share of the 512 -> 1 product into ydot (swl: w_last in smem)
// Returns the 16 gate bits (relu AND keep).
template <bool kTrain, bool kLast>
__device__ __forceinline__ uint32_t epi_pass16(uint32_t taddr, uint32_t keep16, uint32_t swl, uint32_t (&pk)[8], float& ydot) {
  float v[16];
  {
    uint32_t r[16];
    tmem_ld16(taddr, r);
    tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
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
  if (kLast) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float4 w = lds_f32x4(swl + 16 * k);
      ydot += v[4 * k + 0] * w.x + v[4 * k + 1] * w.y + v[4 * k + 2] * w.z + v[4 * k + 3] * w.w;
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) pk[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
  return gate;
}

// (hi, lo, hi) of x as bf16 bit patterns: the A-side triple that meets the B-side (w_hi, w_hi, w_lo)
struct Tri { uint32_t h, l; };
__device__ __forceinline__ Tri split3(float x) {
  __nv_bfloat16 hi, lo;
  split_bf16(x, hi, lo);
  return Tri{uint32_t(__bfloat16_as_ushort(hi)), uint32_t(__bfloat16_as_ushort(lo))};
}

}  // namespace

template <bool kTrain>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(P_THREADS, 1)
chain_forward_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmActs,
                     const __grid_constant__ CUtensorMap tmAux, const __grid_constant__ CUtensorMap tmScene,
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
  if (warp == P_EPI_WARPS && lane == 0) {
    tma_prefetch_desc(&tmW); tma_prefetch_desc(&tmActs); tma_prefetch_desc(&tmAux); tma_prefetch_desc(&tmScene);
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int nh = args.num_hidden;
  const int n_pairs = (args.n_tiles + 1) >> 1;
  const int pair0 = blockIdx.x >> 1, pair_stride = gridDim.x >> 1;
  const bool row_mode = args.row_lat != nullptr;

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
#ifdef DSDF_TRACE
        const bool trace_on = blockIdx.x == 0 && tp == pair0 + 3 * pair_stride && lane == 0;
#endif
        // the two tiles of the pair may belong to different scenes: k-core 0 of a per-scene aux block is taken from
        // rank 0's scene, k-core 1 from rank 1's (fused.cuh)
        int scene0 = 0, scene1 = 0;
        if (!row_mode) {
          const int t0 = 2 * tp, t1 = (2 * tp + 1 < args.n_tiles) ? 2 * tp + 1 : args.n_tiles - 1;
          scene0 = __ldg(args.tile_scene + t0);
          scene1 = __ldg(args.tile_scene + t1);
        }
        for (int l = 0; l < nh; ++l) {
          const int n_halves = args.prog[l].n_halves, k_chunks = args.prog[l].k_chunks;
          const int row_base = args.prog[l].w_row0 + int(rank) * 128;
          const int slot = row_mode ? -1 : args.prog[l].bias_slot;
          const int aux_cores = 2 * args.aux_steps[l];
          for (int h = 0; h < n_halves; ++h) {
            // ---- aux stage: bias (+ xyz / latent) columns of these 256 output features
            mbar_wait_a(empty0 + stage * 8, phase ^ 1);
            TR(3, l, h * 9);
            if (elect_one()) {
              if (leader) mbar_arrive_expect_tx_a(full0 + stage * 8, uint32_t(2 * aux_cores) * AUX_CORE_BYTES);
              const uint32_t dst = sB_u + stage * P_STAGE_BYTES, bar = full_leader + stage * 8;
              const int r = h * HALF_N + int(rank) * 128;
              if (slot >= 0) {
                tma_load_2d_pair_hint_a(dst, &tmScene, bar, 0, (scene0 * 2 + slot) * FH + r, pol_w);
                tma_load_2d_pair_hint_a(dst + AUX_CORE_BYTES, &tmScene, bar, 8, (scene1 * 2 + slot) * FH + r, pol_w);
              } else {
                for (int c = 0; c < aux_cores; ++c)
                  tma_load_2d_pair_hint_a(dst + c * AUX_CORE_BYTES, &tmAux, bar, 8 * c, l * FH + r, pol_w);
              }
            }
            __syncwarp();
            if (++stage == P_STAGES) { stage = 0; phase ^= 1; }
            for (int kc = 0; kc < k_chunks; ++kc) {
              mbar_wait_a(empty0 + stage * 8, phase ^ 1);
              TR(3, l, h * 9 + kc + 1);
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
      const uint64_t aux_hi = umma_desc_hi_noswz(AUX_CORE_BYTES, 128);      // LBO = next k-core, SBO = next 8 rows
      const uint64_t a_base = desc_hi | uint64_t((smem_u32(sA) >> 4) & 0x3FFF);
      const uint64_t b_base = desc_hi | uint64_t((smem_u32(sB) >> 4) & 0x3FFF);
      const uint64_t a_aux = aux_hi | uint64_t(((smem_a + P_OFF_AUX) >> 4) & 0x3FFF);
      const uint64_t b_aux = aux_hi | uint64_t((smem_u32(sB) >> 4) & 0x3FFF);
      const uint32_t full0 = smem_u32(&bar_full[0]), empty0 = smem_u32(&bar_empty[0]);
      const uint32_t tfull0 = smem_u32(&bar_tmem_full[0]), fq0 = smem_u32(&bar_free_q0), flo = smem_u32(&bar_free_lo);
      const uint32_t ready0 = smem_u32(&bar_a_ready[0]);
      uint32_t stage = 0, phase = 0, a_cons = 0;
      for (int tp = pair0; tp < n_pairs; tp += pair_stride) {
#ifdef DSDF_TRACE
        const bool trace_on = blockIdx.x == 0 && tp == pair0 + 3 * pair_stride && lane == 0;
#endif
        for (int l = 0; l < nh; ++l) {
          const int n_halves = args.prog[l].n_halves, k_chunks = args.prog[l].k_chunks;
          const int aux_steps = args.aux_steps[l];
          const uint32_t apar = a_cons & 1; ++a_cons;
          const int kc_lo_last = k_chunks < 4 ? k_chunks - 1 : 3;      // -1: the aux step is the last reader
          const int kc_q0_last = k_chunks < 2 ? k_chunks - 1 : 1;
          uint32_t waited = 0;     // quarters of the previous layer's output whose a_ready has been consumed
          for (int h = 0; h < n_halves; ++h) {
            const bool last_half = (h == n_halves - 1);
            for (int kc = -1; kc < k_chunks; ++kc) {
              // kc = -1 is the aux step: it reads the aux A block (written at the start of the tile, before the a_ready
              // arrivals layer 0 waits for) and OVERWRITES this half's accumulators, which the previous layer's epilogue
              // reads quarter by quarter: half 0 needs quarters 0,1 drained, half 1 needs quarters 2,3.
              // K chunk kc >= 0 reads quarter kc/2 of the A tile.
              uint32_t needm = (kc < 0) ? (3u << (2 * h)) : (1u << (kc >> 1));
              needm &= ~waited;
              while (needm) {
                const int qd = __ffs(int(needm)) - 1;
                mbar_wait_a(ready0 + qd * 8, apar);
                needm &= needm - 1;
                waited |= 1u << qd;
              }
              TR(1, l, 2 * (h * 9 + kc + 1));
              mbar_wait_a(full0 + stage * 8, phase);
              tc_fence_after();
              TR(1, l, 2 * (h * 9 + kc + 1) + 1);
              const uint64_t ad = a_base + uint64_t(kc * (A_CHUNK_BYTES >> 4));
              const uint64_t bd = b_base + uint64_t(stage * (P_STAGE_BYTES >> 4));
              const uint64_t bx = b_aux + uint64_t(stage * (P_STAGE_BYTES >> 4));
              const uint32_t d = tmem_base + h * HALF_N;
              if (elect_one()) {
                if (kc < 0) {
                  umma_bf16_ss_pair(d, a_aux, bx, idesc, 0u);
                  if (aux_steps == 2) umma_bf16_ss_pair(d, a_aux + ((2 * AUX_CORE_BYTES) >> 4), bx + ((2 * AUX_CORE_BYTES) >> 4), idesc, 1u);
                } else {
                  umma_bf16_ss_pair(d, ad, bd, idesc, 1u);
                  umma_bf16_ss_pair(d, ad + 2, bd + 2, idesc, 1u);
                  umma_bf16_ss_pair(d, ad + 4, bd + 4, idesc, 1u);
                  umma_bf16_ss_pair(d, ad + 6, bd + 6, idesc, 1u);
                }
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
    const uint32_t aux_dst = smem_a + P_OFF_AUX + uint32_t(j) * AUX_CORE_BYTES + rt * 16u;   // k-core j of this row
    const uint32_t swl = smem_a + P_OFF_WLAST;
    uint32_t mma_iter = 0;
    const bool stash = kTrain && args.store_acts != 0;
    uint32_t wpar = 0, wany = 0;        // per quarter: parity of / any hand-overs to the stash warp so far
    const int tiles_alloc = int(args.n_rows_pad / TILE_M);
    const uint32_t one = 0x3F80u;       // bf16 1.0
    const uint32_t one0 = leader ? one : 0u, one1 = leader ? 0u : one;

    const int dense_chunks = args.dense_chunks;
    // per-step dropout streams without per-step kernel parameters: the optimiser step counter lives in device memory
    const uint64_t step_key = (kTrain && args.step_dev != nullptr) ? (uint64_t(__ldg(args.step_dev)) << 40) : 0ull;
    for (int tp = pair0; tp < n_pairs; tp += pair_stride) {
      const int tile = 2 * tp + int(rank);
      const bool tile_ok = tile < args.n_tiles;
      const int tile_c = tile_ok ? tile : args.n_tiles - 1;
      const bool valid = tile_ok && (int64_t(tile_c) * TILE_M + rt) < args.n_rows &&
                         (args.tile_rows == nullptr || int(rt) < __ldg(args.tile_rows + tile_c));
#ifdef DSDF_TRACE
      const bool trace_on = blockIdx.x <= 1 && tp == pair0 + 3 * pair_stride && lane == 0 && (warp == 0 || (warp == 15 && blockIdx.x == 0));
      const int trole = blockIdx.x == 1 ? 4 : (warp == 0 ? 0 : 2);
#endif
      // dense per-row codes: this row's code as bf16 K chunks [chunk0, chunk0 + dense_chunks) of the activation tile
      // (thread j of the row writes chunks j, j + 4, ...; columns beyond L are zero)
      auto write_code_chunks = [&](int chunk0) {
        const float* pl = args.row_lat + (int64_t(tile_c) * TILE_M + rt) * args.row_lat_stride;
        for (int c = j; c < dense_chunks; c += 4) {
          const uint32_t line = smem_a + uint32_t(chunk0 + c) * A_CHUNK_BYTES + rt * 128u;
#pragma unroll
          for (int p = 0; p < 8; ++p) {
            float f[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int col = c * 64 + p * 8 + i;
              f[i] = (valid && col < args.row_lat_dim) ? __ldg(pl + col) : 0.0f;
            }
            sts_u32x4(line + ((uint32_t(p) ^ (rt & 7u)) << 4), pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]),
                      pack_bf16x2(f[6], f[7]));
          }
        }
      };
      if (dense_chunks > 0) write_code_chunks(0);      // lin0's K operand (all earlier MMAs of this pair have completed, see below)
      // ---- aux A block of this tile: k-core j of row rt (every previous MMA of this CTA pair has completed: the last
      //      epilogue of the previous tile waited for its tmem_full)
      {
        uint32_t w0 = 0, w1 = 0, w2 = 0, w3 = 0;
        if (j < 2) {
          float px = 0.f, py = 0.f, pz = 0.f;
          if (valid) {
            const float* pt = args.xyz + (int64_t(tile_c) * TILE_M + rt) * args.xyz_stride;
            px = __ldg(pt); py = __ldg(pt + 1); pz = __ldg(pt + 2);
          }
          if (j == 0) {       // cols 0..7: one0, one0, x(h,l,h), y(h,l,h)
            const Tri x = split3(px), y = split3(py);
            w0 = one0 | (one0 << 16); w1 = x.h | (x.l << 16); w2 = x.h | (y.h << 16); w3 = y.l | (y.h << 16);
          } else {            // cols 8..15: z(h,l,h), one1, one1, 0, 0, 0
            const Tri z = split3(pz);
            w0 = z.h | (z.l << 16); w1 = z.h | (one1 << 16); w2 = one1; w3 = 0u;
          }
        } else if (row_mode && dense_chunks == 0) {
          // cols 16 + 3d ..: (hi, lo, hi) of latent component d < 4 -> 12 bf16 over k-cores 2 and 3
          uint32_t e[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) e[i] = 0u;
          if (valid) {
            const float* pl = args.row_lat + (int64_t(tile_c) * TILE_M + rt) * args.row_lat_stride;
            if (args.row_lat_dim <= AUX_MAX_ROW_LAT) {
#pragma unroll
              for (int d = 0; d < AUX_MAX_ROW_LAT; ++d)
                if (d < args.row_lat_dim) {
                  const Tri t = split3(__ldg(pl + d));
                  e[3 * d] = t.h; e[3 * d + 1] = t.l; e[3 * d + 2] = t.h;
                }
            } else {
              // 4 < L <= 16: one bf16 term per component (this thread's k-core only: components 8 (j - 2) .. 8 (j - 2) + 7)
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const int d = (j - 2) * 8 + i;
                const uint32_t v = (d < args.row_lat_dim) ? uint32_t(__bfloat16_as_ushort(__float2bfloat16_rn(__ldg(pl + d)))) : 0u;
                if (j == 2) e[i] = v; else e[8 + i] = v;
              }
            }
          }
          if (j == 2) { w0 = e[0] | (e[1] << 16); w1 = e[2] | (e[3] << 16); w2 = e[4] | (e[5] << 16); w3 = e[6] | (e[7] << 16); }
          else { w0 = e[8] | (e[9] << 16); w1 = e[10] | (e[11] << 16); w2 = e[12] | (e[13] << 16); w3 = e[14] | (e[15] << 16); }
        }
        sts_u32x4(aux_dst, w0, w1, w2, w3);
        fence_proxy_async_smem();
        tc_fence_before();
        __syncwarp();
        // the "layer -1" hand-over: layer 0's MMAs wait for these like any layer waits for its predecessor's quarters
        if (lane == 0) {
#pragma unroll
          for (int Q = 0; Q < 4; ++Q) mbar_arrive_remote(ready_leader + 8u * uint32_t(Q));
        }
      }
      float ydot = 0.0f;

      for (int l = 0; l < nh; ++l) {
        const ChainLayer& L = args.prog[l];
        const bool last = (l == nh - 1);
        const bool produce_a = !last;
        const bool drop = kTrain && L.dropout != 0 && args.dropout_on != 0;
        // ---- before the accumulators are ready: draw the first keep words
        const uint64_t grow = args.row0 + step_key + uint64_t(int64_t(tile_c) * TILE_M + rt);
        // all four keep words of the layer now: the warps idle here anyway (the tensor cores are still on half 0), and later
        // every cycle between tmem_full[1] and the last a_ready arrive is on the critical path of the next layer
        uint32_t kwa = 0xFFFFFFFFu, kwb = 0xFFFFFFFFu, kwc = 0xFFFFFFFFu, kwd = 0xFFFFFFFFu;
        TR(trole, l, 0);
        if (drop) {
          kwa = dropout_keep_word_rk(args.philox_rk, uint32_t(l), grow, uint32_t(j), args.drop_thr);
          kwb = dropout_keep_word_rk(args.philox_rk, uint32_t(l), grow, uint32_t(4 + j), args.drop_thr);
          if (L.n_halves == 2) {
            kwc = dropout_keep_word_rk(args.philox_rk, uint32_t(l), grow, uint32_t(8 + j), args.drop_thr);
            kwd = dropout_keep_word_rk(args.philox_rk, uint32_t(l), grow, uint32_t(12 + j), args.drop_thr);
          }
        }
        TR(trole, l, 1);      // keep words drawn
        const uint32_t par = mma_iter & 1;
        ++mma_iter;
        uint32_t* gdst = nullptr;
        if (kTrain && args.bitmask != nullptr && tile_ok)
          gdst = args.bitmask + ((size_t(l) * tiles_alloc + tile) * MASK_WORDS + j) * TILE_M + rt;   // word 4*Q + j at + 4*Q*128

        // One 128-column quarter: two 16-column passes straight from TMEM into the A tile.
        // `free_bar` (or null): the mbarrier that says the MMAs no longer read the activation chunks of this quarter
        auto quarter_v = [&](int Q, uint32_t keep, uint64_t* free_bar, auto LAST) {
          constexpr bool kLast = decltype(LAST)::value;
          const uint32_t c0 = uint32_t(Q) * 128u + uint32_t(j) * 32u;
          const uint32_t qoff = uint32_t(Q) * 2u * A_CHUNK_BYTES;        // a quarter = two 64-column chunks
          uint32_t p0[8], p1[8];
          const uint32_t g0 = epi_pass16<kTrain, kLast>(tmem_row + c0, keep, swl + c0 * 4, p0, ydot);
          const uint32_t g1 = epi_pass16<kTrain, kLast>(tmem_row + c0 + 16, keep >> 16, swl + (c0 + 16) * 4, p1, ydot);
          if (kTrain && gdst != nullptr) gdst[size_t(Q) * 4 * TILE_M] = g0 | (g1 << 16);
          // only the stores have to wait: for the tensor cores to be done with the old chunks, and for the previous layer's
          // stash of this quarter to have been read out of the tile
          TR(trole, l, 3 + 3 * Q + (Q >= 2 ? 1 : 0));                         // quarter computed
          if (free_bar != nullptr) mbar_wait(free_bar, par);
          if (stash && ((wany >> Q) & 1u)) mbar_wait(&bar_q_stashed[Q], ((wpar >> Q) & 1u) ^ 1u);
          TR(trole, l, 4 + 3 * Q + (Q >= 2 ? 1 : 0));                         // store slot free
          sts_u32x4(st_a0 + qoff, p0[0], p0[1], p0[2], p0[3]);
          sts_u32x4(st_a1 + qoff, p0[4], p0[5], p0[6], p0[7]);
          sts_u32x4(st_a2 + qoff, p1[0], p1[1], p1[2], p1[3]);
          sts_u32x4(st_a3 + qoff, p1[4], p1[5], p1[6], p1[7]);
          fence_proxy_async_smem();
          tc_fence_before();
          __syncwarp();
          if (produce_a && lane == 0) mbar_arrive_remote(ready_leader + 8u * uint32_t(Q));
          if (stash) {
            if (lane == 0) mbar_arrive(&bar_q_written[Q]);
            wpar ^= 1u << Q; wany |= 1u << Q;
          }
          TR(trole, l, 5 + 3 * Q + (Q >= 2 ? 1 : 0));                         // stored + signalled
        };
        // one run-time dispatch per quarter instead of flag tests in every 16-column pass
        auto quarter = [&](int Q, uint32_t keep, uint64_t* free_bar) {
          if (last) quarter_v(Q, keep, free_bar, std::true_type{});
          else quarter_v(Q, keep, free_bar, std::false_type{});
        };

        // ---- half 0: ready while the tensor cores still compute half 1.  Activation chunks 0,1 may be overwritten once
        //      half 1's MMAs have passed K chunk 1 (free_q0), chunks 2,3 after K chunk 3 (free_lo); the previous layer's
        //      stash (TMA store) must have been read out of the tile.
        mbar_wait(&bar_tmem_full[0], par); tc_fence_after();
        TR(trole, l, 2);
        quarter(0, kwa, &bar_free_q0);
        quarter(1, kwb, &bar_free_lo);
        // ---- half 1: after all of this layer's MMAs; the next layer starts on the K chunks that are already there
        mbar_wait(&bar_tmem_full[1], par); tc_fence_after();
        TR(trole, l, 9);
        if (L.n_halves == 2) {
          quarter(2, kwc, nullptr);
          quarter(3, kwd, nullptr);
        } else {
          if (kTrain && gdst != nullptr) { gdst[size_t(2) * 4 * TILE_M] = 0u; gdst[size_t(3) * 4 * TILE_M] = 0u; }
          if (dense_chunks > 0 && l + 1 == args.dense_layer) {
            // the upper half of the tile is free (this layer's MMAs are done: tmem_full[1]; its output fills chunks 0..3 only):
            // the next layer's latent K chunks go there
            if (stash) {
              if ((wany >> 2) & 1u) mbar_wait(&bar_q_stashed[2], ((wpar >> 2) & 1u) ^ 1u);
              if ((wany >> 3) & 1u) mbar_wait(&bar_q_stashed[3], ((wpar >> 3) & 1u) ^ 1u);
            }
            write_code_chunks(4);
            fence_proxy_async_smem();
          }
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

#ifdef DSDF_TRACE
extern "C" int dsdf_debug_trace(void* host_out) {
  return cudaMemcpyFromSymbol(host_out, g_trace, sizeof(g_trace)) == cudaSuccess ? 0 : 1;
}
#endif

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

int launch_chain_forward(const CUtensorMap& tmW, const CUtensorMap& tmActs, const CUtensorMap& tmAux, const CUtensorMap& tmScene,
                         const FwdArgs& args, int n_sms, cudaStream_t stream) {
  auto kern = args.training ? chain_forward_kernel<true> : chain_forward_kernel<false>;
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, P_SMEM));
  static int max_pairs[2] = {0, 0};
  int& mp = max_pairs[args.training ? 1 : 0];
  if (mp == 0) mp = resident_pairs(kern, n_sms);
  const int n_pairs = (args.n_tiles + 1) / 2;
  int grid_pairs = mp;
  if (grid_pairs > n_pairs) grid_pairs = n_pairs;
  if (grid_pairs < 1) grid_pairs = 1;
  DSDF_LAUNCHED;
  kern<<<2 * grid_pairs, P_THREADS, P_SMEM, stream>>>(tmW, tmActs, tmAux, tmScene, args);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int chain_forward_resident_pairs(int n_sms) {
  cudaFuncSetAttribute(chain_forward_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, P_SMEM);
  return resident_pairs(chain_forward_kernel<true>, n_sms);
}

}  // namespace dsdf
