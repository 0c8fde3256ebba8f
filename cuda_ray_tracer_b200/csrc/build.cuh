// build.cuh — internal interface between build_sort.cu (bounds, Morton codes, onesweep radix sort)
// and build_tree.cu (hierarchy, refit, export, rt_build_lbvh).
#pragma once
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

// build_sort.cu
int rt_build_codes_and_sort(rt_scene* s, int morton_bits, unsigned flags, cudaStream_t st, uint32_t* launches);
