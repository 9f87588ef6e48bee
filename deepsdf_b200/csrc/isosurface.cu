// Iso-surface extraction on the device-resident SDF volume (the tail of create_mesh, deep_sdf/mesh.py:88-155:
// convert_sdf_samples_to_ply calls skimage.measure.marching_cubes on a host copy of the [N,N,N] tensor).
// skimage (scikit-image, not pinned by the reference, absent here) implements Lewiner's marching cubes with its 256-case
// tables; this is MARCHING TETRAHEDRA instead - every cube is cut into the six Kuhn tetrahedra around its main diagonal
// (corners 0 -> 7), each tetrahedron has 16 sign cases and yields 0, 1 or 2 triangles.  The face diagonals of neighbouring
// cubes coincide, so the triangle soup is watertight; it is a different triangulation of the same piecewise-linear
// level set (vertices on grid edges AND on face / body diagonals), agreeing with marching cubes to O(voxel^2).
// HBM-bound integer work: two passes over the volume (count, emit) around an exclusive scan of the per-block counts.
//   volume[(ix*ny + iy)*nz + iz]   (create_mesh's ordering, mesh.py:45-47), inside <=> value < level
// Vertices of an edge are interpolated from its lower-index end point, so the two cubes (or tetrahedra) sharing an edge
// produce bit-identical coordinates.  Triangles are oriented with their normal towards increasing SDF (outwards).
// Explicit _rn arithmetic (no FMA contraction) keeps the CPU restatement oracle/isosurface_oracle.py bit-exact.
#include "common.cuh"
#include "host_util.h"

namespace dsdf {
namespace {

constexpr int ISO_THREADS = 256;
__constant__ int8_t c_tets[6][4] = {{0, 1, 3, 7}, {0, 1, 5, 7}, {0, 2, 3, 7}, {0, 2, 6, 7}, {0, 4, 5, 7}, {0, 4, 6, 7}};

struct IsoArgs {
  const float* vol;
  int32_t nx, ny, nz;
  float level;
  float origin[3], voxel[3];
  int64_t n_cubes;
};

__device__ __forceinline__ int tris_of_mask(uint32_t m) {
  const int n = __popc(m & 15u);
  return (n == 0 || n == 4) ? 0 : (n == 2 ? 2 : 1);
}

// corner values of cube c and its integer origin; returns false when c is out of range
__device__ __forceinline__ bool load_cube(const IsoArgs& a, int64_t c, float (&v)[8], int& ix, int& iy, int& iz) {
  if (c >= a.n_cubes) return false;
  const int mz = a.nz - 1, my = a.ny - 1;
  if (a.n_cubes <= 0x7FFFFFFF) {            // 32-bit index arithmetic (64-bit division costs more than the rest of the thread)
    const uint32_t c32 = uint32_t(c), r = c32 / uint32_t(mz);
    iz = int(c32 - r * uint32_t(mz));
    ix = int(r / uint32_t(my));
    iy = int(r - uint32_t(ix) * uint32_t(my));
  } else {
    iz = int(c % mz);
    const int64_t r = c / mz;
    iy = int(r % my);
    ix = int(r / my);
  }
#pragma unroll
  for (int b = 0; b < 8; ++b)
    v[b] = __ldg(a.vol + (int64_t(ix + (b & 1)) * a.ny + (iy + ((b >> 1) & 1))) * a.nz + (iz + (b >> 2)));
  return true;
}

__device__ __forceinline__ int count_cube(const float (&v)[8], float level) {
  uint32_t in = 0;
#pragma unroll
  for (int b = 0; b < 8; ++b) in |= (v[b] < level ? 1u : 0u) << b;
  if (in == 0u || in == 255u) return 0;
  int n = 0;
#pragma unroll
  for (int t = 0; t < 6; ++t) {
    uint32_t m = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) m |= ((in >> c_tets[t][q]) & 1u) << q;
    n += tris_of_mask(m);
  }
  return n;
}

__global__ void __launch_bounds__(ISO_THREADS) iso_count_kernel(const IsoArgs a, int32_t* __restrict__ block_counts) {
  const int64_t c = int64_t(blockIdx.x) * ISO_THREADS + threadIdx.x;
  float v[8];
  int ix, iy, iz;
  int n = load_cube(a, c, v, ix, iy, iz) ? count_cube(v, a.level) : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
  __shared__ int s[ISO_THREADS / 32];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = n;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
#pragma unroll
    for (int w = 0; w < ISO_THREADS / 32; ++w) t += s[w];
    block_counts[blockIdx.x] = t;
  }
}

struct P3 { float x, y, z; };

// vertex on the edge between cube corners ca and cb: interpolated from the end point with the lower grid index
__device__ __forceinline__ P3 edge_vertex(const IsoArgs& a, const float (&v)[8], int ix, int iy, int iz, int ca, int cb) {
  // corner bits are x | y<<1 | z<<2, the linear grid index has x slowest: compare (x, y, z) lexicographically
  const int ax = ca & 1, ay = (ca >> 1) & 1, az = ca >> 2, bx = cb & 1, by = (cb >> 1) & 1, bz = cb >> 2;
  const bool a_first = (ax != bx) ? (ax < bx) : ((ay != by) ? (ay < by) : (az < bz));
  const int lo = a_first ? ca : cb, hi = a_first ? cb : ca;
  const float vl = v[lo], vh = v[hi];
  const float t = __fdiv_rn(__fsub_rn(a.level, vl), __fsub_rn(vh, vl));
  const int lx = lo & 1, ly = (lo >> 1) & 1, lz = lo >> 2, hx = hi & 1, hy = (hi >> 1) & 1, hz = hi >> 2;
  P3 p;
  p.x = __fadd_rn(float(ix + lx), hx != lx ? t : 0.0f);
  p.y = __fadd_rn(float(iy + ly), hy != ly ? t : 0.0f);
  p.z = __fadd_rn(float(iz + lz), hz != lz ? t : 0.0f);
  return p;
}

__device__ __forceinline__ void put_triangle(const IsoArgs& a, float* __restrict__ out, int64_t tri, P3 p0, P3 p1, P3 p2, float rx,
                                             float ry, float rz) {
  // orient: normal (in grid-index space) along the inside -> outside direction r
  const float ux = __fsub_rn(p1.x, p0.x), uy = __fsub_rn(p1.y, p0.y), uz = __fsub_rn(p1.z, p0.z);
  const float wx = __fsub_rn(p2.x, p0.x), wy = __fsub_rn(p2.y, p0.y), wz = __fsub_rn(p2.z, p0.z);
  const float nx = __fsub_rn(__fmul_rn(uy, wz), __fmul_rn(uz, wy));
  const float ny = __fsub_rn(__fmul_rn(uz, wx), __fmul_rn(ux, wz));
  const float nz = __fsub_rn(__fmul_rn(ux, wy), __fmul_rn(uy, wx));
  const float d = __fadd_rn(__fadd_rn(__fmul_rn(nx, rx), __fmul_rn(ny, ry)), __fmul_rn(nz, rz));
  if (d < 0.0f) { const P3 t = p1; p1 = p2; p2 = t; }
  float* o = out + tri * 9;
  const P3 ps[3] = {p0, p1, p2};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    o[3 * k + 0] = __fadd_rn(a.origin[0], __fmul_rn(ps[k].x, a.voxel[0]));
    o[3 * k + 1] = __fadd_rn(a.origin[1], __fmul_rn(ps[k].y, a.voxel[1]));
    o[3 * k + 2] = __fadd_rn(a.origin[2], __fmul_rn(ps[k].z, a.voxel[2]));
  }
}

__global__ void __launch_bounds__(ISO_THREADS) iso_emit_kernel(const IsoArgs a, const int64_t* __restrict__ block_offsets,
                                                               float* __restrict__ verts) {
  const int64_t c = int64_t(blockIdx.x) * ISO_THREADS + threadIdx.x;
  float v[8];
  int ix = 0, iy = 0, iz = 0;
  const bool ok = load_cube(a, c, v, ix, iy, iz);
  const int n = ok ? count_cube(v, a.level) : 0;
  // block-exclusive scan of n
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = n;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
  __shared__ int s[ISO_THREADS / 32];
  if (lane == 31) s[warp] = incl;
  __syncthreads();
  int base = 0;
#pragma unroll
  for (int w = 0; w < ISO_THREADS / 32; ++w) if (w < warp) base += s[w];
  if (n == 0) return;
  int64_t tri = block_offsets[blockIdx.x] + base + incl - n;
  uint32_t in = 0;
#pragma unroll
  for (int b = 0; b < 8; ++b) in |= (v[b] < a.level ? 1u : 0u) << b;
  for (int t = 0; t < 6; ++t) {
    int cid[4];
    uint32_t m = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) { cid[q] = c_tets[t][q]; m |= ((in >> cid[q]) & 1u) << q; }
    const int n_in = __popc(m);
    if (n_in == 0 || n_in == 4) continue;
    auto dir = [&](int from, int to, float& rx, float& ry, float& rz) {       // corner `from` -> corner `to`
      rx = float((to & 1) - (from & 1)); ry = float(((to >> 1) & 1) - ((from >> 1) & 1)); rz = float((to >> 2) - (from >> 2));
    };
    float rx, ry, rz;
    if (n_in == 1 || n_in == 3) {
      const uint32_t lone_mask = (n_in == 1) ? m : (~m & 15u);
      const int ql = __ffs(int(lone_mask)) - 1;
      int o3[3], k = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) if (q != ql) o3[k++] = cid[q];
      const int al = cid[ql];
      if (n_in == 1) dir(al, o3[0], rx, ry, rz); else dir(o3[0], al, rx, ry, rz);
      put_triangle(a, verts, tri++, edge_vertex(a, v, ix, iy, iz, al, o3[0]), edge_vertex(a, v, ix, iy, iz, al, o3[1]),
                   edge_vertex(a, v, ix, iy, iz, al, o3[2]), rx, ry, rz);
    } else {
      int ins[2], outs[2], ki = 0, ko = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) { if ((m >> q) & 1u) ins[ki++] = cid[q]; else outs[ko++] = cid[q]; }
      dir(ins[0], outs[0], rx, ry, rz);
      const P3 q0 = edge_vertex(a, v, ix, iy, iz, ins[0], outs[0]), q1 = edge_vertex(a, v, ix, iy, iz, ins[0], outs[1]);
      const P3 q2 = edge_vertex(a, v, ix, iy, iz, ins[1], outs[1]), q3 = edge_vertex(a, v, ix, iy, iz, ins[1], outs[0]);
      put_triangle(a, verts, tri++, q0, q1, q2, rx, ry, rz);
      put_triangle(a, verts, tri++, q0, q2, q3, rx, ry, rz);
    }
  }
}

int fill_args(IsoArgs* a, const float* vol, int32_t nx, int32_t ny, int32_t nz, float level) {
  DSDF_REQUIRE(vol != nullptr && nx >= 2 && ny >= 2 && nz >= 2, "iso-surface: the volume needs at least 2 samples per axis");
  a->vol = vol; a->nx = nx; a->ny = ny; a->nz = nz; a->level = level;
  a->n_cubes = int64_t(nx - 1) * (ny - 1) * (nz - 1);
  DSDF_REQUIRE((a->n_cubes + ISO_THREADS - 1) / ISO_THREADS < (int64_t(1) << 31), "iso-surface: volume too large");
  return 0;
}

}  // namespace
}  // namespace dsdf

using namespace dsdf;

extern "C" int64_t dsdf_iso_num_blocks(int32_t nx, int32_t ny, int32_t nz) {
  if (nx < 2 || ny < 2 || nz < 2) return 0;
  return (int64_t(nx - 1) * (ny - 1) * (nz - 1) + ISO_THREADS - 1) / ISO_THREADS;
}

extern "C" int dsdf_iso_count(const float* volume, int32_t nx, int32_t ny, int32_t nz, float level, int32_t* block_counts,
                              void* stream) {
  if (int e = require_sm100()) return e;
  IsoArgs a;
  if (int e = fill_args(&a, volume, nx, ny, nz, level)) return e;
  a.origin[0] = a.origin[1] = a.origin[2] = 0.f; a.voxel[0] = a.voxel[1] = a.voxel[2] = 1.f;
  DSDF_LAUNCHED;
  iso_count_kernel<<<unsigned(dsdf_iso_num_blocks(nx, ny, nz)), ISO_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(a, block_counts);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dsdf_iso_emit(const float* volume, int32_t nx, int32_t ny, int32_t nz, float level, const float* origin3,
                             const float* voxel3, const int64_t* block_offsets, float* verts, void* stream) {
  if (int e = require_sm100()) return e;
  IsoArgs a;
  if (int e = fill_args(&a, volume, nx, ny, nz, level)) return e;
  DSDF_REQUIRE(origin3 && voxel3 && block_offsets && verts, "iso-surface: null argument");
  for (int k = 0; k < 3; ++k) { a.origin[k] = origin3[k]; a.voxel[k] = voxel3[k]; }
  DSDF_LAUNCHED;
  iso_emit_kernel<<<unsigned(dsdf_iso_num_blocks(nx, ny, nz)), ISO_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(a, block_offsets, verts);
  DSDF_CHECK_CUDA(cudaGetLastError());
  return 0;
}
