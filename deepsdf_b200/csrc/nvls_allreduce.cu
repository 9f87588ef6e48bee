// All-reduce (sum) of the flat fp32 decoder gradient over NVSwitch multicast memory ("NVLS"): the one exchange step of the
// multi-GPU training path (replaces nn.DataParallel's gradient reduction, train_deep_sdf.py:353), as this library's own
// kernel instead of a NCCL call.  The gradient buffer of every rank lives in symmetric memory that is also mapped through ONE
// multicast address: a `multimem.ld_reduce` of that address returns the SUM of all ranks' copies, added inside the switch, and
// a `multimem.st` writes all ranks' copies at once.  Rank r reduces and re-broadcasts slice r of the buffer (a two-shot
// all-reduce whose two shots are one load and one store per 16 bytes), between two cross-GPU barriers on the symmetric
// signal pads:
//   barrier A   every rank's producer kernels (wgrad, assemble_grads) have finished - the kernel is stream-ordered behind them
//   reduce      v = multimem.ld_reduce.add [mc + i];  multimem.st [mc + i], v          for the 16-byte pieces of this rank's slice
//   barrier B   every rank's stores have landed everywhere before any consumer (clip, Adam) runs
// 8 GPUs, 7.4 MB: ~30 us against ~70 us for ncclAllReduce on the same box (tools/symm_probe.py) - what matters once a rank's share
// of the batch is small (strong scaling).  Allocation and the rendezvous that yields the multicast / signal-pad addresses are the
// caller's (torch.distributed._symmetric_memory on the Python side: plumbing); this file only computes.
#include "host_util.h"
#include "../../include/dsdf.h"

namespace dsdf {

namespace {

constexpr int NV_BLOCKS = 64;             // CTAs of the all-reduce kernel (each owns one barrier channel)
constexpr int NV_THREADS = 512;
constexpr int NV_UNROLL = 4;
constexpr unsigned long long NV_WAIT_NS = 120ull * 1000000000ull;    // bounded waits: a dead peer traps after two minutes instead
                                                                     // of hanging the GPU (peers may legitimately be seconds apart:
                                                                     // one rank writing a checkpoint, a first-call compile ...)
__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ uint32_t cas_release_sys(uint32_t* addr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.global.release.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t cas_acquire_sys(uint32_t* addr, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val) : "memory");
  return old;
}

// Cross-GPU barrier of block `channel` of every rank: thread p < world raises flag (channel, rank) on peer p's pad and waits
// for flag (channel, p) on its own pad, resetting it - the pads are all-zero between barriers.
__device__ __forceinline__ void rank_barrier(uint32_t* const* pads, int rank, int world, int channel) {
  __syncthreads();
  if (int(threadIdx.x) < world) {
    const int p = threadIdx.x;
    uint32_t* put = pads[p] + size_t(channel) * world + rank;
    uint32_t* get = pads[rank] + size_t(channel) * world + p;
    const unsigned long long deadline = now_ns() + NV_WAIT_NS;
    uint32_t spins = 0;
    while (cas_release_sys(put, 0u, 1u) != 0u) { if ((++spins & 0xFFFu) == 0u && now_ns() > deadline) __trap(); }
    while (cas_acquire_sys(get, 1u, 0u) != 1u) { if ((++spins & 0xFFFu) == 0u && now_ns() > deadline) __trap(); }
  }
  __syncthreads();
}

__global__ void __launch_bounds__(NV_THREADS) nvls_allreduce_kernel(float* __restrict__ mc, uint32_t* const* __restrict__ pads,
                                                                    int rank, int world, int64_t n_vec4) {
  rank_barrier(pads, rank, world, blockIdx.x);                       // A: all producers done
  const int64_t per = (n_vec4 + world - 1) / world;
  const int64_t beg = per * rank;
  const int64_t end = beg + per < n_vec4 ? beg + per : n_vec4;
  // four independent 16-byte pieces per thread and trip: the switch round trip (~2-3 us) is paid once per trip, not per piece
  const int64_t stride = int64_t(gridDim.x) * NV_THREADS;
  for (int64_t i0 = beg + int64_t(blockIdx.x) * NV_THREADS + threadIdx.x; i0 < end; i0 += NV_UNROLL * stride) {
    float v[NV_UNROLL][4];
#pragma unroll
    for (int u = 0; u < NV_UNROLL; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < end)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v[u][0]), "=f"(v[u][1]), "=f"(v[u][2]), "=f"(v[u][3]) : "l"(mc + 4 * i) : "memory");
    }
#pragma unroll
    for (int u = 0; u < NV_UNROLL; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < end)
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc + 4 * i), "f"(v[u][0]), "f"(v[u][1]),
                     "f"(v[u][2]), "f"(v[u][3]) : "memory");
    }
  }
  __threadfence_system();
  rank_barrier(pads, rank, world, gridDim.x + blockIdx.x);           // B: all stores landed
}

}  // namespace

}  // namespace dsdf

extern "C" int32_t dsdf_nvls_signal_words(int32_t world) { return 2 * dsdf::NV_BLOCKS * world; }

extern "C" int dsdf_nvls_allreduce_f32(float* multicast_ptr, const void* signal_pad_ptrs_dev, int32_t rank, int32_t world, int64_t n,
                                       void* stream) {
  using namespace dsdf;
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(multicast_ptr != nullptr && signal_pad_ptrs_dev != nullptr, "dsdf_nvls_allreduce_f32: null argument");
  DSDF_REQUIRE(world >= 2 && world <= 64 && rank >= 0 && rank < world, "dsdf_nvls_allreduce_f32: bad rank / world");
  DSDF_REQUIRE(n > 0 && (n & 3) == 0 && (reinterpret_cast<uintptr_t>(multicast_ptr) & 15) == 0,
               "dsdf_nvls_allreduce_f32: n must be a multiple of 4 floats and the buffer 16-byte aligned");
  DSDF_LAUNCHED;
  nvls_allreduce_kernel<<<NV_BLOCKS, NV_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
      multicast_ptr, static_cast<uint32_t* const*>(signal_pad_ptrs_dev), rank, world, n / 4);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}
