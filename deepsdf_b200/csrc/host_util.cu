#include "host_util.h"

namespace dsdf {

thread_local char g_last_error[512] = {0};
long long g_kernel_launches = 0;

int require_sm100() {
  int dev = 0;
  DSDF_CHECK_CUDA(cudaGetDevice(&dev));
  int major = 0, minor = 0;
  DSDF_CHECK_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  DSDF_CHECK_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  if (major != 10) {
    snprintf(g_last_error, sizeof(g_last_error),
             "deepsdf_b200 kernels are built for sm_100a only; device %d is sm_%d%d (no fallback path)", dev, major, minor);
    return 3;
  }
  return 0;
}

int sm_count() {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
  return n;
}

PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<PFN_encodeTiled>(p);
  return fn;
}

static int make_tmap_bf16_2d_impl(CUtensorMap* out, const void* gptr, uint64_t inner, uint64_t outer, uint64_t pitch_bytes,
                                  uint32_t box_inner, uint32_t box_outer, CUtensorMapSwizzle swz) {
  PFN_encodeTiled enc = get_encode_tiled();
  DSDF_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled driver entry point not available");
  DSDF_REQUIRE((reinterpret_cast<uintptr_t>(gptr) & 15) == 0, "TMA global address must be 16-byte aligned");
  DSDF_REQUIRE((pitch_bytes & 15) == 0, "TMA row pitch must be a multiple of 16 bytes");
  DSDF_REQUIRE(box_inner * 2 <= 128 && box_outer <= 256, "TMA box exceeds SWIZZLE_128B limits");
  cuuint64_t gdim[2] = {inner, outer};
  cuuint64_t gstride[1] = {pitch_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(gptr), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_last_error, sizeof(g_last_error), "cuTensorMapEncodeTiled failed with CUresult %d", int(r));
    return 2;
  }
  return 0;
}

int make_tmap_bf16_2d(CUtensorMap* out, const void* gptr, uint64_t inner, uint64_t outer, uint64_t pitch_bytes,
                      uint32_t box_inner, uint32_t box_outer) {
  return make_tmap_bf16_2d_impl(out, gptr, inner, outer, pitch_bytes, box_inner, box_outer, CU_TENSOR_MAP_SWIZZLE_128B);
}

int make_tmap_bf16_2d_noswz(CUtensorMap* out, const void* gptr, uint64_t inner, uint64_t outer, uint64_t pitch_bytes,
                            uint32_t box_inner, uint32_t box_outer) {
  return make_tmap_bf16_2d_impl(out, gptr, inner, outer, pitch_bytes, box_inner, box_outer, CU_TENSOR_MAP_SWIZZLE_NONE);
}

}  // namespace dsdf
