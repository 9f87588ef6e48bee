// Device-resident SDF sample store + on-device per-epoch subsampler: deep_sdf/data.py:75-110 (`unpack_sdf_samples`) and
// :142-194 (`SDFSamples.__getitem__`) without touching the host.  The reference reloads every scene's npz from disk for
// every batch (np.load + randperm + index_select on the CPU; >= 57 % of its A100 epoch); here all
// scenes' rows live in HBM ([rows,4] fp32 xyz+sdf, per scene its positive rows then its negative rows, NaN rows already
// dropped like `remove_nans`), and one kernel writes a scene-major batch [S * SamplesPerScene, 4] - the layout
// dsdf_fused_forward reads in place.
//
// Per scene (data.py:88-110): half = subsample/2; take `half` positive and `half` negative rows, or - when one sign has
// fewer than `half` rows - all of that sign and 2*half - n of the other; positives first, then negatives; rows drawn
// WITHOUT replacement (`torch.randperm(len)[:take]`).  The draw is a keyed bijection of [0, n): a 4-round Feistel network
// over the next even power of two with cycle walking, so row i of the batch is permutation(i) - no state, no sorting, any
// thread can compute its own source row.
#include "common.cuh"
#include "host_util.h"
#include "../../include/dsdf.h"

namespace dsdf {

__host__ __device__ __forceinline__ uint32_t mix32(uint32_t x) {           // murmur3 finaliser
  x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
  return x;
}

// bijection of [0, n), n >= 1, keyed by (k0, k1)
__host__ __device__ __forceinline__ uint32_t feistel_perm(uint32_t i, uint32_t n, uint32_t k0, uint32_t k1) {
  if (n <= 1) return 0;
  uint32_t bits = 32u - uint32_t(
#ifdef __CUDA_ARCH__
      __clz(int(n - 1))
#else
      __builtin_clz(n - 1)
#endif
  );
  if (bits & 1u) ++bits;                      // balanced halves
  if (bits < 2u) bits = 2u;
  const uint32_t hb = bits >> 1, hmask = (1u << hb) - 1u;
  uint32_t x = i;
  do {                                        // cycle walking: the domain is < 4 n, so a few rounds on average
    uint32_t l = x >> hb, r = x & hmask;
#pragma unroll
    for (uint32_t round = 0; round < 4; ++round) {
      const uint32_t f = mix32(r ^ (k0 + round * 0x9E3779B9u)) ^ mix32(k1 + round);
      const uint32_t nl = r, nr = (l ^ f) & hmask;
      l = nl; r = nr;
    }
    x = (l << hb) | r;
  } while (x >= n);
  return x;
}

// meta[s] = {pos_off, n_pos, neg_off, n_neg} (rows).  grid (ceil(sps/256), S)
__global__ void __launch_bounds__(256) sample_batch_kernel(const float4* __restrict__ store, const int64_t* __restrict__ meta,
                                                           const int64_t* __restrict__ scene_ids, int sps, uint64_t seed,
                                                           float4* __restrict__ out) {
  const int b = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= sps) return;
  const int64_t scene = scene_ids[b];
  const int64_t* m = meta + scene * 4;
  const int64_t n_pos = m[1], n_neg = m[3];
  const int64_t half = sps / 2;
  int64_t take_pos, take_neg;                                    // data.py:92-99
  if (n_pos < half) { take_pos = n_pos; take_neg = 2 * half - n_pos; }
  else if (n_neg < half) { take_neg = n_neg; take_pos = 2 * half - n_neg; }
  else { take_pos = take_neg = half; }
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i < take_pos + take_neg) {
    const bool neg = i >= take_pos;
    const uint32_t j = uint32_t(neg ? i - take_pos : i);
    const uint32_t n = uint32_t(neg ? n_neg : n_pos);
    const uint32_t k0 = mix32(uint32_t(seed) ^ mix32(uint32_t(scene) * 2u + (neg ? 1u : 0u)));
    const uint32_t k1 = mix32(uint32_t(seed >> 32) + 0x632BE5ABu * uint32_t(scene + 1));
    const uint32_t src = feistel_perm(j, n, k0, k1);
    v = __ldg(store + (neg ? m[2] : m[0]) + src);
  }
  out[int64_t(b) * sps + i] = v;
}

}  // namespace dsdf

using namespace dsdf;

extern "C" int dsdf_sample_batch(const float* store, const int64_t* meta, const int64_t* scene_ids, int32_t n_scenes,
                                 int32_t samples_per_scene, uint64_t seed, float* batch_out, void* stream) {
  if (int e = require_sm100()) return e;
  DSDF_REQUIRE(store && meta && scene_ids && batch_out && n_scenes > 0 && samples_per_scene > 0, "dsdf_sample_batch: bad arguments");
  DSDF_REQUIRE((reinterpret_cast<uintptr_t>(store) & 15) == 0 && (reinterpret_cast<uintptr_t>(batch_out) & 15) == 0,
               "sample store and batch must be 16-byte aligned ([rows,4] fp32)");
  DSDF_LAUNCHED;
  sample_batch_kernel<<<dim3((samples_per_scene + 255) / 256, n_scenes), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float4*>(store), meta, scene_ids, samples_per_scene, seed, reinterpret_cast<float4*>(batch_out));
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// host copy of the row bijection (tests and the host-side check of the sampler)
extern "C" uint32_t dsdf_sample_perm(uint32_t i, uint32_t n, uint64_t seed, int64_t scene, int32_t negative) {
  const uint32_t k0 = mix32(uint32_t(seed) ^ mix32(uint32_t(scene) * 2u + (negative ? 1u : 0u)));
  const uint32_t k1 = mix32(uint32_t(seed >> 32) + 0x632BE5ABu * uint32_t(scene + 1));
  return feistel_perm(i, n, k0, k1);
}
