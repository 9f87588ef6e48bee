// tcgen05 / TMA bring-up micro-tests.  One CTA computes C[128 x N] = A * B^T on the tensor cores with the operand
// layouts the fused kernels rely on, so that every descriptor convention is checked against a PyTorch matmul before
// the big kernels depend on it:
//   mode bit0: A is MN-major (stored [K rows][128 M] in global)  - wgrad's dY^T operand
//   mode bit1: B is MN-major (stored [K rows][N]   in global)    - wgrad's X^T operand
//   mode bit2: A tile is written to shared memory by threads with sw128_offset() instead of TMA (the epilogue's path)
//   mode bit3: additionally TMA-store the A tile back to global (activation / dY stash path)
// Descriptor LBO/SBO are runtime arguments so the host can sweep conventions without rebuilding.
#include "common.cuh"
#include "host_util.h"
#include "../../include/dsdf.h"

namespace dsdf {

struct SelftestArgs {
  int K, N, mode;
  int reps;                     // > 1: repeat the whole K loop (throughput measurement; C is then reps x the product)
  long long* cycles;            // [1] clock64 ticks from the first MMA issue to the completion of the last, or null
  uint32_t a_lbo, a_sbo, b_lbo, b_sbo;
  const __nv_bfloat16* a_raw;   // for manual writes: K-major [128][K]
  float* C;                     // [128][N]
};

constexpr int ST_THREADS = 192;   // warps 0-3 epilogue (TMEM lane quadrants), warp 4 TMA, warp 5 MMA/alloc

__global__ void __launch_bounds__(ST_THREADS, 1)
umma_selftest_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ CUtensorMap tmStore, SelftestArgs args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int K = args.K, N = args.N;
  const bool a_mn = args.mode & 1, b_mn = args.mode & 2, manual_a = args.mode & 4, do_store = args.mode & 8;
  uint8_t* sA = smem;                         // up to 64 KB
  uint8_t* sB = smem + 65536;                 // up to 128 KB
  __shared__ uint64_t bar_full, bar_done;
  __shared__ uint32_t tmem_base_slot;
  const int warp = threadIdx.x >> 5;

  if (threadIdx.x == 0) {
    mbar_init(&bar_full, 1);
    mbar_init(&bar_done, 1);
    fence_barrier_init();
  }
  if (warp == 5) tmem_alloc<256>(&tmem_base_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;

  // ---- stage operands
  uint32_t a_bytes = 128u * K * 2u, b_bytes = uint32_t(N) * K * 2u;
  if (manual_a) {
    // K-major A written by threads: chunk kc = [128 rows][64 cols]
    for (int idx = threadIdx.x; idx < 128 * K; idx += blockDim.x) {
      int r = idx / K, c = idx % K;
      uint32_t off = (c >> 6) * (128u * 128u) + sw128_offset(r, c & 63);
      *reinterpret_cast<__nv_bfloat16*>(sA + off) = args.a_raw[idx];
    }
    fence_proxy_async_smem();
  }
  __syncthreads();
  if (warp == 4 && elect_one()) {
    uint32_t tx = b_bytes + (manual_a ? 0u : a_bytes);
    mbar_arrive_expect_tx(&bar_full, tx);
    if (!manual_a) {
      if (!a_mn) {
        for (int kc = 0; kc < K / 64; ++kc) tma_load_2d(sA + kc * 16384, &tmA, &bar_full, kc * 64, 0);   // box {64, 128}
      } else {
        for (int mg = 0; mg < 2; ++mg) tma_load_2d(sA + mg * (K * 128), &tmA, &bar_full, mg * 64, 0);    // box {64, K}
      }
    }
    if (!b_mn) {
      for (int kc = 0; kc < K / 64; ++kc) tma_load_2d(sB + kc * (N * 128), &tmB, &bar_full, kc * 64, 0);   // box {64, N}
    } else {
      for (int ng = 0; ng < N / 64; ++ng) tma_load_2d(sB + ng * (K * 128), &tmB, &bar_full, ng * 64, 0);   // box {64, K}
    }
  }
  if (warp == 5) {
    mbar_wait(&bar_full, 0);
    tc_fence_after();
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(128, N, a_mn ? 1 : 0, b_mn ? 1 : 0);
      const uint64_t a_hi = umma_desc_hi_sw128(args.a_lbo, args.a_sbo);
      const uint64_t b_hi = umma_desc_hi_sw128(args.b_lbo, args.b_sbo);
      const long long t0 = clock64();
      for (int rep = 0; rep < args.reps; ++rep)
        for (int ks = 0; ks < K / 16; ++ks) {
          uint32_t a_off = a_mn ? ks * 2048u : (ks >> 2) * 16384u + (ks & 3) * 32u;
          uint32_t b_off = b_mn ? ks * 2048u : (ks >> 2) * (uint32_t(N) * 128u) + (ks & 3) * 32u;
          umma_bf16_ss(tmem_base, umma_desc(smem_u32(sA) + a_off, a_hi), umma_desc(smem_u32(sB) + b_off, b_hi), idesc,
                       (ks > 0 || rep > 0) ? 1u : 0u);
        }
      umma_commit(&bar_done);
      mbar_wait(&bar_done, 0);
      if (args.cycles) args.cycles[0] = clock64() - t0;
    }
    __syncwarp();
  }
  if (warp < 4) {
    mbar_wait(&bar_done, 0);
    tc_fence_after();
    const int row = warp * 32 + lane_id();
    for (int c0 = 0; c0 < N; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(tmem_base + (uint32_t(warp * 32) << 16) + c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) args.C[size_t(row) * N + c0 + j] = __uint_as_float(v[j]);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (do_store && !a_mn && threadIdx.x == 0) {
    // write the (swizzled) A tile back through a SWIZZLE_128B tensor map: must land un-swizzled in global
    for (int kc = 0; kc < K / 64; ++kc) tma_store_2d(&tmStore, sA + kc * 16384, kc * 64, 0);
    tma_store_commit();
    tma_store_wait_all0();
  }
  __syncthreads();
  if (warp == 5) tmem_dealloc<256>(tmem_base);
}

}  // namespace dsdf

extern "C" int dsdf_selftest_umma(int mode, const void* A, const void* B, float* C, void* A_store_out, int K, int N,
                                  unsigned a_lbo, unsigned a_sbo, unsigned b_lbo, unsigned b_sbo, void* stream) {
  using namespace dsdf;
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(K % 64 == 0 && K >= 64 && K <= 256, "selftest: K must be 64..256, multiple of 64");
  DSDF_REQUIRE(N % 64 == 0 && N >= 64 && N <= 256, "selftest: N must be 64..256, multiple of 64");
  const bool a_mn = mode & 1, b_mn = mode & 2;
  const bool timing = (mode >> 8) > 0;
  CUtensorMap tmA, tmB, tmS;
  int e;
  if (!a_mn) e = make_tmap_bf16_2d(&tmA, A, K, 128, uint64_t(K) * 2, 64, 128);
  else       e = make_tmap_bf16_2d(&tmA, A, 128, K, 128 * 2, 64, K);
  if (e) return e;
  if (!b_mn) e = make_tmap_bf16_2d(&tmB, B, K, N, uint64_t(K) * 2, 64, N);
  else       e = make_tmap_bf16_2d(&tmB, B, N, K, uint64_t(N) * 2, 64, K);
  if (e) return e;
  const void* store_ptr = (A_store_out && !timing) ? A_store_out : A;
  e = make_tmap_bf16_2d(&tmS, store_ptr, K, 128, uint64_t(K) * 2, 64, 128);
  if (e) return e;
  if (!timing && (mode & 8) && !A_store_out) return set_error("selftest: store mode needs an output buffer");
  SelftestArgs args{K, N, mode & 15, (mode >> 8) > 0 ? (mode >> 8) : 1, reinterpret_cast<long long*>((mode >> 8) > 0 ? A_store_out : nullptr),
                    a_lbo, a_sbo, b_lbo, b_sbo, reinterpret_cast<const __nv_bfloat16*>(A), C};
  if ((mode >> 8) > 0) { mode &= 7; args.mode = mode; }   // throughput runs: A_store_out receives the cycle count instead
  const size_t smem = 65536 + 131072 + 1024;
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  umma_selftest_kernel<<<1, ST_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(tmA, tmB, tmS, args);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// CTA-pair (cta_group::2) bring-up: C[256 x 256] = A[256 x K] * B[256 x K]^T with ONE cluster of two CTAs.
// Each CTA owns 128 rows of A and of C (its own TMEM) and HALF of B's N rows (rank r: B rows [128r, 128r+128)).
// Checks, against a PyTorch matmul: the pair TMEM allocation, the 2-SM TMA load signalling the leader's barrier, the
// M=256 instruction descriptor, the multicast commit, and the epilogue->MMA hand-over used by the fused chain kernels:
// mode bit2: A is written by the CTA's own threads (generic proxy, sw128_offset) and handed to the leader's MMA thread
// through a remote mbarrier arrive (peer CTA -> leader CTA).
namespace dsdf {

struct Selftest2Args {
  int K, mode;
  const __nv_bfloat16* a_raw;   // [256][K]
  float* C;                     // [256][256]
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(ST_THREADS, 1)
umma_pair_selftest_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, Selftest2Args args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int K = args.K;
  const bool manual_a = args.mode & 4;
  uint8_t* sA = smem;                         // [K/64 chunks][128 rows][64] = up to 64 KB
  uint8_t* sB = smem + 65536;                 // [K/64 chunks][128 N rows][64] = up to 64 KB (this CTA's half of B)
  __shared__ uint64_t bar_full, bar_done, bar_a_ready;
  __shared__ uint32_t tmem_base_slot;
  const int warp = threadIdx.x >> 5;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;

  if (threadIdx.x == 0) {
    mbar_init(&bar_full, 1);
    mbar_init(&bar_done, 1);
    mbar_init(&bar_a_ready, 2 * 4);            // 4 epilogue warps of each CTA
    fence_barrier_init();
  }
  if (warp == 5) tmem_alloc_pair<256>(&tmem_base_slot);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  const uint32_t full_leader = mapa_u32(smem_u32(&bar_full), 0);
  const uint32_t ready_leader = mapa_u32(smem_u32(&bar_a_ready), 0);
  const uint32_t bytes_per_cta = 2u * 128u * uint32_t(K) * 2u;   // A half + B half

  if (warp < 4) {
    if (manual_a) {
      for (int idx = threadIdx.x; idx < 128 * K; idx += 128) {
        int r = idx / K, c = idx % K;
        uint32_t off = (c >> 6) * (128u * 128u) + sw128_offset(r, c & 63);
        *reinterpret_cast<__nv_bfloat16*>(sA + off) = args.a_raw[size_t(rank * 128 + r) * K + c];
      }
      fence_proxy_async_all();
    }
    tc_fence_before();
    __syncwarp();
    if (lane_id() == 0) mbar_arrive_cluster(ready_leader);
  }
  if (warp == 4 && elect_one()) {
    if (leader) mbar_arrive_expect_tx(&bar_full, 2u * bytes_per_cta - (manual_a ? 2u * 128u * uint32_t(K) * 2u : 0u));
    for (int kc = 0; kc < K / 64; ++kc) {
      if (!manual_a) tma_load_2d_pair(sA + kc * 16384, &tmA, full_leader, kc * 64, int(rank) * 128);
      tma_load_2d_pair(sB + kc * 16384, &tmB, full_leader, kc * 64, int(rank) * 128);
    }
  }
  if (warp == 5 && leader) {
    mbar_wait(&bar_full, 0);
    mbar_wait_cluster(&bar_a_ready, 0);
    tc_fence_after();
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(256, 256, 0, 0);
      const uint64_t hi = umma_desc_hi_sw128(16, 1024);
      for (int ks = 0; ks < K / 16; ++ks) {
        const uint32_t off = (ks >> 2) * 16384u + (ks & 3) * 32u;
        umma_bf16_ss_pair(tmem_base, umma_desc(smem_u32(sA) + off, hi), umma_desc(smem_u32(sB) + off, hi), idesc, ks > 0 ? 1u : 0u);
      }
      umma_commit_pair(&bar_done, 0b11);
    }
    __syncwarp();
  }
  if (warp < 4) {
    mbar_wait(&bar_done, 0);
    tc_fence_after();
    const int row = warp * 32 + lane_id();
    for (int c0 = 0; c0 < 256; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(tmem_base + (uint32_t(warp * 32) << 16) + c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) args.C[size_t(rank * 128 + row) * 256 + c0 + j] = __uint_as_float(v[j]);
    }
    tc_fence_before();
  }
  __syncthreads();
  cluster_sync_all();
  if (warp == 5) tmem_dealloc_pair<256>(tmem_base);
}

}  // namespace dsdf

extern "C" int dsdf_selftest_umma_pair(int mode, const void* A, const void* B, float* C, int K, void* stream) {
  using namespace dsdf;
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(K % 64 == 0 && K >= 64 && K <= 256, "selftest: K must be 64..256, multiple of 64");
  CUtensorMap tmA, tmB;
  if (int e = make_tmap_bf16_2d(&tmA, A, K, 256, uint64_t(K) * 2, 64, 128)) return e;
  if (int e = make_tmap_bf16_2d(&tmB, B, K, 256, uint64_t(K) * 2, 64, 128)) return e;
  Selftest2Args args{K, mode, reinterpret_cast<const __nv_bfloat16*>(A), C};
  const size_t smem = 65536 + 65536 + 1024;
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(umma_pair_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  umma_pair_selftest_kernel<<<2, ST_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(tmA, tmB, args);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// Aux-step bring-up: the pair GEMM above plus aux_k (16 or 32) extra K columns whose operands use the SWIZZLE_NONE K-major
// canonical layout the forward chain's bias / xyz step relies on: A_aux [256][32] written by threads (k-core c of row r at
// c * 2048 + r * 16), B_aux [256][32] fetched by no-swizzle TMA boxes {8, 128} (one per k-core).  C = A B^T + A_aux B_aux^T.
// LBO / SBO of the aux descriptors are run-time arguments so that the host can confirm the convention.
namespace dsdf {

struct Selftest3Args {
  int K, aux_k;
  uint32_t lbo, sbo;
  const __nv_bfloat16* a_aux;   // [256][32]
  float* C;                     // [256][256]
};

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(ST_THREADS, 1)
umma_aux_selftest_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const __grid_constant__ CUtensorMap tmBaux, Selftest3Args args) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int K = args.K, n_cores = args.aux_k / 8;
  uint8_t* sA = smem;
  uint8_t* sB = smem + 65536;
  uint8_t* sAaux = smem + 131072;               // 4 k-cores x 2 KB
  uint8_t* sBaux = smem + 131072 + 8192;
  __shared__ uint64_t bar_full, bar_done, bar_a_ready;
  __shared__ uint32_t tmem_base_slot;
  const int warp = threadIdx.x >> 5;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  if (threadIdx.x == 0) {
    mbar_init(&bar_full, 1);
    mbar_init(&bar_done, 1);
    mbar_init(&bar_a_ready, 2 * 4);
    fence_barrier_init();
  }
  if (warp == 5) tmem_alloc_pair<256>(&tmem_base_slot);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  const uint32_t full_leader = mapa_u32(smem_u32(&bar_full), 0);
  const uint32_t ready_leader = mapa_u32(smem_u32(&bar_a_ready), 0);
  if (warp < 4) {
    const int r = threadIdx.x;                       // row of this CTA's A_aux
    for (int c = 0; c < n_cores; ++c) {
      const uint4 v = *reinterpret_cast<const uint4*>(args.a_aux + size_t(rank * 128 + r) * 32 + c * 8);
      *reinterpret_cast<uint4*>(sAaux + c * 2048 + r * 16) = v;
    }
    fence_proxy_async_all();
    tc_fence_before();
    __syncwarp();
    if (lane_id() == 0) mbar_arrive_cluster(ready_leader);
  }
  if (warp == 4 && elect_one()) {
    const uint32_t per_cta = 2u * 128u * uint32_t(K) * 2u + uint32_t(n_cores) * 2048u;
    if (leader) mbar_arrive_expect_tx(&bar_full, 2u * per_cta);
    for (int kc = 0; kc < K / 64; ++kc) {
      tma_load_2d_pair(sA + kc * 16384, &tmA, full_leader, kc * 64, int(rank) * 128);
      tma_load_2d_pair(sB + kc * 16384, &tmB, full_leader, kc * 64, int(rank) * 128);
    }
    for (int c = 0; c < n_cores; ++c) tma_load_2d_pair(sBaux + c * 2048, &tmBaux, full_leader, c * 8, int(rank) * 128);
  }
  if (warp == 5 && leader) {
    mbar_wait(&bar_full, 0);
    mbar_wait_cluster(&bar_a_ready, 0);
    tc_fence_after();
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(256, 256, 0, 0);
      const uint64_t hi = umma_desc_hi_sw128(16, 1024);
      const uint64_t hi_aux = umma_desc_hi_noswz(args.lbo, args.sbo);
      for (int s2 = 0; s2 < args.aux_k / 16; ++s2)
        umma_bf16_ss_pair(tmem_base, umma_desc(smem_u32(sAaux) + s2 * 4096, hi_aux), umma_desc(smem_u32(sBaux) + s2 * 4096, hi_aux), idesc,
                          s2 > 0 ? 1u : 0u);
      for (int ks = 0; ks < K / 16; ++ks) {
        const uint32_t off = (ks >> 2) * 16384u + (ks & 3) * 32u;
        umma_bf16_ss_pair(tmem_base, umma_desc(smem_u32(sA) + off, hi), umma_desc(smem_u32(sB) + off, hi), idesc, 1u);
      }
      umma_commit_pair(&bar_done, 0b11);
    }
    __syncwarp();
  }
  if (warp < 4) {
    mbar_wait(&bar_done, 0);
    tc_fence_after();
    const int row = warp * 32 + lane_id();
    for (int c0 = 0; c0 < 256; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(tmem_base + (uint32_t(warp * 32) << 16) + c0, v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 32; ++j) args.C[size_t(rank * 128 + row) * 256 + c0 + j] = __uint_as_float(v[j]);
    }
    tc_fence_before();
  }
  __syncthreads();
  cluster_sync_all();
  if (warp == 5) tmem_dealloc_pair<256>(tmem_base);
}

}  // namespace dsdf

extern "C" int dsdf_selftest_umma_aux(const void* A, const void* B, const void* A_aux, const void* B_aux, float* C, int K, int aux_k,
                                      unsigned lbo, unsigned sbo, void* stream) {
  using namespace dsdf;
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(K % 64 == 0 && K >= 64 && K <= 256, "selftest: K must be 64..256, multiple of 64");
  DSDF_REQUIRE(aux_k == 16 || aux_k == 32, "selftest: aux_k must be 16 or 32");
  CUtensorMap tmA, tmB, tmBaux;
  if (int e = make_tmap_bf16_2d(&tmA, A, K, 256, uint64_t(K) * 2, 64, 128)) return e;
  if (int e = make_tmap_bf16_2d(&tmB, B, K, 256, uint64_t(K) * 2, 64, 128)) return e;
  if (int e = make_tmap_bf16_2d_noswz(&tmBaux, B_aux, 32, 256, 64, 8, 128)) return e;
  Selftest3Args args{K, aux_k, lbo, sbo, reinterpret_cast<const __nv_bfloat16*>(A_aux), C};
  const size_t smem = 131072 + 16384 + 1024;
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(umma_aux_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  umma_aux_selftest_kernel<<<2, ST_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(tmA, tmB, tmBaux, args);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// Tensor-pipe rate probe: one CTA pair issues `iters` x 8 K-chunks of cta_group::2 MMAs (M = 256, N = n, K = 16 each,
// SS mode: A [128 x 512] and B from shared memory - contents are whatever the launch finds there, only the rate is
// measured) back to back and reports the SM clocks between the first issue and the completion of the last commit.
// n = 256: 4 MMAs per chunk; n = 128: 8 MMAs per chunk (the same MACs).  With lsu_load != 0 the four other warps of each
// CTA hammer the A tile with STS.128 meanwhile (an epilogue's share of the shared-memory port).
namespace dsdf {

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(ST_THREADS, 1)
mma_rate_kernel(int n, int iters, int lsu_load, unsigned long long* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar_done;
  __shared__ uint32_t tmem_base_slot;
  __shared__ volatile int stop_flag;
  const int warp = threadIdx.x >> 5;
  const bool leader = cluster_ctarank() == 0;
  if (threadIdx.x == 0) { mbar_init(&bar_done, 1); fence_barrier_init(); stop_flag = 0; }
  if (warp == 5) tmem_alloc_pair<512>(&tmem_base_slot);
  for (int i = threadIdx.x; i < (131072 + 65536) / 16; i += ST_THREADS) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_all();
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_slot;
  if (warp == 5 && leader) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc_bf16(256, uint32_t(n), 0, 0);
      const uint64_t hi = umma_desc_hi_sw128(16, 1024);
      const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem) + 131072;
      const long long t0 = clock64();
      for (int it = 0; it < iters; ++it)
        for (int kc = 0; kc < 8; ++kc)
          for (int g = 0; g < 256 / n; ++g)
            for (int ks = 0; ks < 4; ++ks)
              umma_bf16_ss_pair(tmem_base + g * n, umma_desc(a0 + kc * 16384u + ks * 32u, hi),
                                umma_desc(b0 + (kc & 3) * 16384u + g * 8192u + ks * 32u, hi), idesc, 1u);
      umma_commit_pair(&bar_done, 0b11);
      mbar_wait(&bar_done, 0);
      out[0] = (unsigned long long)(clock64() - t0);
      stop_flag = 1;
    }
    __syncwarp();
  } else if (warp < 4 && lsu_load) {
    // 128 threads: STS.128 into the A tile, row = thread, rotating 16-byte pieces (swizzle-free: contents do not matter)
    uint32_t a = smem_u32(smem) + threadIdx.x * 128u, k = 0;
    unsigned long long cnt = 0;
    while (!stop_flag || !leader) {
      sts_u32x4(a + ((k & 7u) << 4) + ((k >> 3) & 7u) * 16384u, k, k, k, k);
      ++k; ++cnt;
      if (!leader && cnt > (unsigned long long)iters * 8ull * 64ull) break;     // the peer CTA has no flag: bounded
    }
    if (leader && threadIdx.x == 0) out[1] = cnt;
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 5) tmem_dealloc_pair<512>(tmem_base);
}

}  // namespace dsdf

extern "C" int dsdf_selftest_mma_rate(int n, int iters, int lsu_load, unsigned long long* out_dev, void* stream) {
  using namespace dsdf;
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(n == 256 || n == 128, "mma rate probe: n must be 256 or 128");
  const size_t smem = 131072 + 65536 + 1024;
  DSDF_CHECK_CUDA(cudaFuncSetAttribute(mma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  mma_rate_kernel<<<2, ST_THREADS, smem, static_cast<cudaStream_t>(stream)>>>(n, iters, lsu_load, out_dev);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}
