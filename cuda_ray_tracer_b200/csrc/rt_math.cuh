// rt_math.cuh — device vector arithmetic with the reference's operation order.
//
// Every function restates an operator of include/vec3.cuh:21-107 (12-byte vec3, fequal, normalize)
// so that a ray, a hit distance or a normal computed here has the bits the reference computes: the
// library is built with -fmad=false (no a*b+c contraction), '/' and sqrtf are IEEE, and
// sums are written in the reference's association ((x+y)+z).
#pragma once
#include <cuda_runtime.h>
#include <math.h>

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 mk3(float a, float b, float c) { V3 v; v.x = a; v.y = b; v.z = c; return v; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return mk3(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator/(V3 a, float s) { return mk3(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ V3 mul3(V3 a, V3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }  // RGB * RGB, struct.cuh:30-33
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }  // vec3.cuh:62-64,95-98
__device__ __forceinline__ float length3(V3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }  // vec3.cuh:66-68

// vec3.cuh:7-18
__device__ __forceinline__ bool fequal(float a, float b) {
  const float diff = fabsf(a - b);
  const float largest = fmaxf(fabsf(a), fabsf(b));
  if (largest < 1e-6f) return diff < 1e-6f;
  return diff / largest < 1e-6f;
}
// vec3.cuh:70-82
__device__ __forceinline__ V3 normalize3(V3 a) {
  const float mag = length3(a);
  if (fequal(mag, 0.0f)) return mk3(0.0f, 0.0f, 0.0f);
  const float inv_mag = 1.0f / mag;
  return mk3(a.x * inv_mag, a.y * inv_mag, a.z * inv_mag);
}
// RGB::operator== against RGB(0,0,0), struct.cuh:20-23 (the "is this material shiny / transparent" test)
__device__ __forceinline__ bool rgb_is_zero(V3 c) { return fequal(c.x, 0.0f) && fequal(c.y, 0.0f) && fequal(c.z, 0.0f); }
