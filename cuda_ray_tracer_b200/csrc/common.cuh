// common.cuh — internal types of libb200rt.so (device data layout, error handling).
//
// Data layout in HBM (all arrays 256-byte aligned inside one arena):
//   prim_bv  [n][3] float4  "build view", file order.   triangle: p0.xyz p1.x | p1.yz p2.xy | p2.z slack 0 tag
//                                                      sphere:   c.xyz r      | 0 0 0 0     | 0 slack 0 tag
//   prim_tv  [n][3] float4  "trace view", file order.   triangle: p0.xyz nor.x | nor.yz e1.xy | e1.z e2.xyz
//                                                      sphere:   c.xyz r      | 0 ...
//   prim_mat [n][3] float4  Materials (11 floats + pad), file order
//   sorted_tv[n][6] float4  prim_tv gathered into Morton order (leaf k reads record k), 96-byte records:
//                            the three prim_tv words + (file-order index, 0, 0, 0) | the leaf's own AABB in node-half
//                            layout (xmin ymin xmax ymax | zmin zmax 0 0): read only when a hit is accepted, for the
//                            exact lane's certificate (traverse.cuh)
//   nodes    [n-1][4] float4 64-byte internal node = two 32-byte child halves
//                            half: xmin ymin xmax ymax | zmin zmax ref aux   (ref<0: leaf, see below); the pairs
//                            (xmin,ymin) (xmax,ymax) (zmin,zmax) feed the packed f32x2 slab test (traverse.cuh)
// Everything the tracer touches is indexed by file-order primitive index or by sorted position;
// the reference's (type,id) PrimitiveReference pairs survive only for rt_export_reference_layout.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/rt_api.h"

#ifndef RT_INLINE_SUN
#define RT_INLINE_SUN 0   // experiment: inline sun shadow rays (traverse.cuh)
#endif
#define RT_TV_STRIDE 6  // float4 words per sorted leaf record (96 bytes = three 32-byte sectors)
// File-order primitive records: 48 bytes (three float4) per primitive in each of the two views.
// (Measured and rejected, profiles/build_ab_r02.txt: ONE interleaved 128-byte record per primitive, so that the builder's
// gather through the sorted permutation touches one location instead of two.  The gather got cheaper, but the two
// streaming passes over the build view (bounds, Morton codes) then read 128-byte strides for 48 useful bytes: 0.336 ms
// per build at 1M triangles against 0.302.)
#define RT_PRIM_STRIDE 3
// format of the 4-wide nodes (RT_BUILD_WIDE4): 0 = four fp32 child halves, 128 bytes; 1 = bfloat16 plane offsets from the
// node's min corner, 96 bytes (build_tree.cu collapse4_kernel, traverse.cuh)
#ifndef RT_WIDE_BF16
#define RT_WIDE_BF16 0
#endif
#define RT_TAG_TRI 1.0f
#define RT_TAG_SPHERE 2.0f

// child reference inside a node half: >=0 internal node slot; <0 leaf:
//   bit31 = 1, bit30 = sphere flag, bits 0..29 = sorted position k
#define RT_LEAF_BIT 0x80000000u
#define RT_SPHERE_BIT 0x40000000u
#define RT_LEAF_MASK 0x3FFFFFFFu

void rt_set_error(const std::string& msg);
#define RT_CUDA(call)                                                                                \
  do {                                                                                               \
    cudaError_t e_ = (call);                                                                         \
    if (e_ != cudaSuccess) {                                                                         \
      rt_set_error(std::string(#call) + " failed at " + __FILE__ + ":" + std::to_string(__LINE__) + \
                   ": " + cudaGetErrorString(e_));                                                   \
      return RT_ERR_CUDA;                                                                            \
    }                                                                                                \
  } while (0)

struct LightSun { float dir[3]; float color[3]; };
struct LightBulb { float point[3]; float color[3]; };
struct PlaneRec { float nor[3]; float point[3]; float mat[11]; float pad; };  // 72 bytes

// scalar render parameters, passed to kernels by value (mirrors RawConfig's scalars, config.hpp:75-98)
struct RtCamera {
  int width, height, bounces, gi;
  float dof_focus, dof_lens, expose;
  int fisheye, panorama;
  float forward[3], right[3], up[3], eye[3];
};

struct RtDeviceScene {  // device pointers, passed to kernels by value
  int n;                // primitives
  const int* root;      // device word holding the root node slot (or leaf ref if n == 1): read by the kernels, never by the host
  const float4* nodes;
  const float4* wnodes;  // 4-wide nodes (RT_BUILD_WIDE4) or nullptr: 128 bytes = the four grandchildren's halves, same slot numbering
  const float4* sorted_tv;
  const uint32_t* perm;  // sorted position -> file-order index
  const float4* prim_mat;
  const LightSun* suns; int num_suns;
  const LightBulb* bulbs; int num_bulbs;
  const PlaneRec* planes; int num_planes;
};

struct rt_scene {
  int device = 0;
  rt_scene_desc cfg;  // scalar fields only (pointers nulled)
  int n = 0, ns = 0, nt = 0;
  bool any_shiny = false, any_trans = false, any_rough = false;
  // arena
  char* arena = nullptr; size_t arena_bytes = 0;
  float4 *prim_bv = nullptr, *prim_tv = nullptr, *prim_mat = nullptr, *sorted_tv = nullptr, *nodes = nullptr, *wnodes = nullptr;
  int2* refs = nullptr;        // file-order (type,id)
  uint32_t* tri_prim = nullptr;  // triangle id -> file-order prim index (for rt_scene_update_triangles)
  LightSun* suns = nullptr; LightBulb* bulbs = nullptr; PlaneRec* planes = nullptr;
  // build products / scratch
  uint64_t *keys_a = nullptr, *keys_b = nullptr;   // 4- or 8-byte keys live in these buffers
  uint32_t *perm_a = nullptr, *perm_b = nullptr;
  uint32_t* hist = nullptr;        // [8 passes][256]
  uint32_t* lookback = nullptr;    // onesweep tile states
  uint32_t* tickets = nullptr;     // per-pass dynamic tile counters (+ misc counters)
  int* node_flags = nullptr;       // bottom-up arrival flags / other-bound exchange
  int* parent = nullptr;           // top-down variant: parent info per node / leaf
  int* node_karras = nullptr;      // [n-1] slot -> reference (Karras) internal index
  float* bounds = nullptr;         // 6 floats (min3,max3) as ordered ints during reduction
  uint32_t* sorted_keys32 = nullptr; uint64_t* sorted_keys64 = nullptr; uint32_t* sorted_perm = nullptr;
  bool built = false; int built_bits = 0; bool built_topdown = false, built_wide = false;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;   // around the last build (rt_scene_build_ms)
  // RT_RENDER_OVERLAP_BUILD: the point in the caller's stream before this frame's update/build work
  cudaEvent_t ev_fork = nullptr; cudaStream_t fork_stream = nullptr; bool fork_armed = false;
  float* upd_stage[3] = {nullptr, nullptr, nullptr};  // device staging of host vertex arrays (rt_scene_update_triangles)
  // render state (allocated lazily, reused)
  void* rstate = nullptr;
};

int rt_render_free_state(rt_scene* s);
int rt_arm_fork(rt_scene* s, cudaStream_t st);
