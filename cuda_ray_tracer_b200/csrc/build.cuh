// build.cuh — internal interface between build_sort.cu (bounds, Morton codes, onesweep radix sort)
// and build_tree.cu (hierarchy, refit, export, rt_build_lbvh).
#pragma once
#include <cstdlib>
#include <cstring>

#include "common.cuh"

// order-preserving float <-> uint encoding for atomicMin/atomicMax bounds reduction
__host__ __device__ inline uint32_t rt_float_order(float f) {
  uint32_t b;
#ifdef __CUDA_ARCH__
  b = __float_as_uint(f);
#else
  memcpy(&b, &f, 4);
#endif
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ inline float rt_float_unorder(uint32_t u) {
  uint32_t b = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  float f; memcpy(&f, &b, 4); return f;
#endif
}

// primitive AABB from the build view — get_primitive_aabb_device, src/lbvh_builder.cu:33-57 with the
// AABB constructors of include/interval.cuh:55-81 (sphere: ordered c-r..c+r; triangle: min/max of the
// three vertices, axes thinner than 0.01 padded by 0.01 on both sides)
struct RtBox { float xmin, ymin, zmin, xmax, ymax, zmax; };
__device__ inline RtBox rt_prim_box(float4 a, float4 b, float4 c) {
  RtBox o;
  if (c.w == RT_TAG_SPHERE) {
    float lo, hi;
    lo = a.x - a.w; hi = a.x + a.w; if (lo <= hi) { o.xmin = lo; o.xmax = hi; } else { o.xmin = hi; o.xmax = lo; }
    lo = a.y - a.w; hi = a.y + a.w; if (lo <= hi) { o.ymin = lo; o.ymax = hi; } else { o.ymin = hi; o.ymax = lo; }
    lo = a.z - a.w; hi = a.z + a.w; if (lo <= hi) { o.zmin = lo; o.zmax = hi; } else { o.zmin = hi; o.zmax = lo; }
    return o;
  }
  // p0 = a.xyz, p1 = (a.w, b.x, b.y), p2 = (b.z, b.w, c.x)
  o.xmin = fminf(fminf(a.x, a.w), b.z); o.xmax = fmaxf(fmaxf(a.x, a.w), b.z);
  o.ymin = fminf(fminf(a.y, b.x), b.w); o.ymax = fmaxf(fmaxf(a.y, b.x), b.w);
  o.zmin = fminf(fminf(a.z, b.y), c.x); o.zmax = fmaxf(fmaxf(a.z, b.y), c.x);
  if (o.xmax - o.xmin < 0.01f) { o.xmin = o.xmin - 0.01f; o.xmax = o.xmax + 0.01f; }
  if (o.ymax - o.ymin < 0.01f) { o.ymin = o.ymin - 0.01f; o.ymax = o.ymax + 0.01f; }
  if (o.zmax - o.zmin < 0.01f) { o.zmin = o.zmin - 0.01f; o.zmax = o.zmax + 0.01f; }
  return o;
}

// Leaf slack: how far outside its leaf AABB a hit point the reference's primitive test ACCEPTS can lie.
// The triangle test tolerates barycentrics down to -1e-3 (src/struct.cu:150-152, EPSILON), so accepted
// points fill a slightly larger triangle than the one the AABB was built from (include/interval.cuh:61-81);
// a sphere hit near a silhouette carries the cancellation error of sqrt(r^2 - d^2) (struct.cu:88-98).  The
// traversal uses the slack (max over the subtree, stored in the node half's aux word) to cull by distance
// only when no acceptable hit inside the subtree can be closer — see traverse.cuh, exact near-first lane.
// Evaluated in double from the float e1/e2 actually used by the test, so slivers with inaccurate
// e1/e2 get the (large) slack they need; anything non-finite or ill-conditioned gets +INF (never culled).
// sphere: 2e-3 r for the cancellation in sqrt(r^2 - d^2) away from the silhouette, plus 2e-6 per unit of |c| + r for a
// small sphere far from the origin, where the rounding of c - o and t*d + o - c no longer hides inside 2e-3 r
// (traverse.cuh, RT_ORIGIN_SLACK carries the ray-origin share of the same bound)
__device__ inline float rt_sphere_slack(float4 a) {
  return 2e-3f * fabsf(a.w) + 2e-6f * (fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fabsf(a.z)) + fabsf(a.w));
}
__device__ inline float rt_leaf_slack(float4 a, float4 b, float4 c, float4 t0, float4 t1, float4 t2, const RtBox& box) {
  if (c.w == RT_TAG_SPHERE) return rt_sphere_slack(a);
  const double p0[3] = {a.x, a.y, a.z}, p1[3] = {a.w, b.x, b.y}, p2[3] = {b.z, b.w, c.x};
  const double nor[3] = {t0.w, t1.x, t1.y}, e1[3] = {t1.z, t1.w, t2.x}, e2[3] = {t2.y, t2.z, t2.w};
  if (nor[0] == 0.0 && nor[1] == 0.0 && nor[2] == 0.0) return 0.f;  // den == 0 for every ray: never hit
  double d1[3], d2[3], g11 = 0, g12 = 0, g21 = 0, g22 = 0, ne1 = 0, ne2 = 0, nd = 0;
  for (int k = 0; k < 3; k++) {
    d1[k] = p1[k] - p0[k]; d2[k] = p2[k] - p0[k];
    g11 += e1[k] * d1[k]; g12 += e1[k] * d2[k]; g21 += e2[k] * d1[k]; g22 += e2[k] * d2[k];
    ne1 += fabs(e1[k]); ne2 += fabs(e2[k]); nd += fabs(d1[k]) + fabs(d2[k]);
  }
  const double det = g11 * g22 - g12 * g21;
  if (!(fabs(det) > 1e-3) || !(fabs(det) < 1e3) || !isfinite(ne1 + ne2 + nd)) return INFINITY;
  // tolerance on the barycentrics: EPSILON, plus the rounding of the float dot products that compute them
  const double eps = 0.001 * 1.01 + 1e-5 + 4e-6 * (ne1 + ne2) * nd;
  const double lo[3] = {box.xmin, box.ymin, box.zmin}, hi[3] = {box.xmax, box.ymax, box.zmax};
  const double cb[3][2] = {{-eps, -eps}, {1.0 + 2.0 * eps, -eps}, {-eps, 1.0 + 2.0 * eps}};
  double slack = 0.0;
  for (int q = 0; q < 3; q++) {
    const double al = (g22 * cb[q][0] - g12 * cb[q][1]) / det, be = (-g21 * cb[q][0] + g11 * cb[q][1]) / det;
    for (int k = 0; k < 3; k++) {
      const double x = p0[k] + al * d1[k] + be * d2[k];
      slack = fmax(slack, fmax(lo[k] - x, x - hi[k]));
    }
  }
  if (!isfinite(slack)) return INFINITY;
  return (float)(slack * 1.02) * 1.0001f + 1e-30f;
}

// build_sort.cu
int rt_build_codes_and_sort(rt_scene* s, int morton_bits, unsigned flags, cudaStream_t st, uint32_t* launches);

// ---- programmatic dependent launch along the build's chain of kernels ----------------------------------------------
// The build is nine short kernels in a row, each waiting for the previous one (init, bounds, Morton, 4-8 sort passes,
// bottom-up); none fills the GPU for long, so the ~3 us it takes to drain one grid and start the next is paid nine
// times.  Launched with cudaLaunchAttributeProgrammaticStreamSerialization, a kernel's blocks may become resident while
// the previous grid's last blocks still run (the previous kernel says so first thing, rt_pdl_trigger); they do whatever
// does not touch global memory (zeroing shared counters) and then block in rt_pdl_wait until the previous grid has
// COMPLETED and its writes are visible.  Every kernel of the chain waits unconditionally at its top, so completion is
// transitive.  Both instructions are no-ops in a kernel launched the ordinary way.  RT_NO_PDL=1: ordinary launches (A/B).
__device__ __forceinline__ void rt_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void rt_pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
// Measured (profiles/build_pdl_r02.txt): 100K 0.114 -> 0.103 ms, 1M 0.293 -> 0.292 (30-bit) / 0.432 -> 0.413 (63-bit), but 10M
// 30-bit 2.125 -> 2.278 ms: long passes gain nothing from a 3 us hand-over and lose to the next pass's blocks sitting on the
// SMs early.  So the chain is used up to 4M primitives.
#define RT_PDL_MAX_N (1 << 22)
inline bool rt_pdl_for(int n) { return n <= RT_PDL_MAX_N && getenv("RT_NO_PDL") == nullptr; }  // env read per build: scripts/build_ab.py flips it in-process
template <typename... Params, typename... Args>
inline cudaError_t rt_launch_chained(bool dependent, void (*kernel)(Params...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = dependent ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, Params(args)...);
}
