/* include/rt_api.h — C ABI of the B200-native LBVH + ray-tracing core (libb200rt.so).
 *
 * Drop-in boundary for the two hot paths of GJ0407790/cuda_ray_tracer.  The reference has no
 * FFI; its seam is four C++ host functions exchanging one POD struct of device pointers
 * (SURVEY §8b).  Each entry point below names the reference interface it replaces (file:line in
 * the reference tree).  Plain pointers and sizes only; every call returns 0 or a negative
 * rt_status and never exits the process (the reference's CUDA_CHECK → exit(EXIT_FAILURE),
 * main.cu:14-23, becomes a status + rt_last_error()).  There is no CPU fallback: every compute
 * entry point fails with RT_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef RT_API_H
#define RT_API_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  RT_OK = 0, RT_ERR_INVALID = -1, RT_ERR_CUDA = -2, RT_ERR_NOT_BUILT = -3, RT_ERR_IO = -4, RT_ERR_PARSE = -5,
  RT_ERR_NOMEM = -6
} rt_status;

/* ---- scene description: the host-side content of StlConfig / RawConfig ----------------------
 * replaces: initRawConfigFromStl + copyConfigDataToDevice (include/config_utils.cuh:11-16,
 * src/config_utils.cu:18-199) and the RawConfig POD (include/config.hpp:75-126).
 * Array layouts are the reference's: 12-byte vec3 (include/vec3.cuh:21-25), 44-byte Materials
 * (include/object.cuh:17-38: color3, shininess3, trans3, ior, roughness), 8-byte
 * PrimitiveReference {int32 type (0 sphere, 1 triangle), uint32 id_in_type_array}
 * (include/object.cuh:72-88), Sun/Bulb 24 bytes (object.cuh:232-266), Plane 84 bytes
 * (object.cuh:124-149: a,b,c,d, nor3, point3, Materials).  Pointers may be host or device
 * memory (they are read with cudaMemcpyDefault), e.g. NCCL-broadcast device buffers. */
typedef struct rt_scene_desc {
  int32_t width, height;            /* png W H                      config.hpp:76-77  */
  float   color[3];                 /* current colour               config.hpp:78     */
  int32_t bounces, aa;              /*                              config.hpp:79-80  */
  float   dof_focus, dof_lens;      /*                              config.hpp:81-82  */
  float   forward[3], right[3], up[3], eye[3], target_up[3]; /*     config.hpp:84-88  */
  float   expose;                   /* +INF = off                   config.hpp:90     */
  int32_t fisheye, panorama;        /*                              config.hpp:91-92  */
  float   ior, rough;               /*                              config.hpp:94-95  */
  int32_t gi;                       /*                              config.hpp:96     */
  float   trans[3], shine[3];       /*                              config.hpp:97-98  */
  float   scene_min[3], scene_max[3]; /* RawConfig::scene_min/max_corner, config.hpp:112-113;
                                         the reference leaves them at +INF/-INF */
  int32_t num_spheres;   const float* sphere_c; const float* sphere_r; const float* sphere_mat;
  int32_t num_triangles; const float* tri_p0; const float* tri_p1; const float* tri_p2;
                         const float* tri_nor; const float* tri_e1; const float* tri_e2; const float* tri_mat;
  int32_t num_prims;     const int32_t* refs;   /* file order, config.hpp:65 */
  int32_t num_suns;      const float* suns;
  int32_t num_bulbs;     const float* bulbs;
  int32_t num_planes;    const float* planes;
} rt_scene_desc;

typedef struct rt_scene rt_scene;   /* opaque; owns one device arena on one GPU */

/* replaces copyConfigDataToDevice / freeRawConfigDeviceMemory (config_utils.cuh:16,20) */
int rt_scene_create(const rt_scene_desc* desc, int device, rt_scene** out);
int rt_scene_destroy(rt_scene* scene);
/* new vertex positions for every triangle (deforming scenes, "LBVH rebuilt every frame"):
 * recomputes nor/e1/e2 on the device with the Triangle constructor's arithmetic
 * (include/object.cuh:177-191).  Pointers host or device, [num_triangles][3] floats each. */
int rt_scene_update_triangles(rt_scene* scene, const float* p0, const float* p1, const float* p2, void* stream);

/* ---- LBVH build ------------------------------------------------------------------------------
 * replaces build_lbvh_karas(RawConfig&, int morton_bits) (include/lbvh_builder.cuh:14,
 * src/lbvh_builder.cu:401-521) and build_morton_codes_and_sort_primitives
 * (include/lbvh_utils.cuh:22, src/lbvh_utils.cu:77-129).  morton_bits: 30 (the reference's
 * hard-coded 10 bits/axis, lbvh_utils.cu:84) or 63 (21 bits/axis, not in the reference).
 * Asynchronous on `stream` (a cudaStream_t, NULL = default stream) unless ms_out != NULL, in which
 * case the call synchronises and returns the device time of the build in milliseconds. */
enum {
  RT_BOUNDS_FROM_DESC   = 0u,  /* use desc.scene_min/max verbatim: +INF/-INF reproduces the shipped
                                  reference, where every Morton code is 0            */
  RT_BOUNDS_COMPUTED    = 1u,  /* reduce the primitive AABBs on the device (what parse.cpp:147-155,
                                  193-200 accumulates and drops)                                  */
  RT_BUILD_TOPDOWN      = 2u,  /* per-node Karras range/split search (lbvh_builder.cu:103-322) +
                                  separate refit instead of the fused bottom-up builder           */
  RT_BUILD_LOCAL_CLIMB  = 4u,  /* fused bottom-up builder whose levels inside a 512-leaf chunk meet in shared
                                  memory (then the global flags); same tree,
                                  measured no faster (the primitive gather bounds the kernel), kept for A/B */
  RT_BUILD_GLOBAL_CLIMB = 16u, /* fused bottom-up builder with EVERY level through the global flags (round 1's kernel).  The
                                  default since round 2 lets the 32 leaves of a warp meet by shuffle first; same tree.
                                  Kept for A/B (scripts/build_ab.py) and so the tests hold both against the oracle */
  RT_BUILD_WIDE4        = 8u   /* EXPERIMENTAL: after the LBVH, derive a 4-wide tree over the same leaves (a node's
                                  children = its grandchildren, 128-byte nodes).  The near-first lanes of rt_render /
                                  rt_trace then step through it (half the dependent steps per ray); the exported
                                  LBVH, the reference-order lane and the exact lane's fallback keep the binary tree.
                                  Same hit ids: the exact lane's answer is defined on leaves      */
};
int rt_build_lbvh(rt_scene* scene, int morton_bits, unsigned flags, void* stream, float* ms_out);
/* Device time of the last rt_build_lbvh (the number lbvh_builder.cu:489 prints), read AFTER the fact so
 * that a frame loop can keep the build asynchronous (ms_out = NULL): waits for that build only. */
int rt_scene_build_ms(rt_scene* scene, float* ms_out);

/* ---- render ----------------------------------------------------------------------------------
 * replaces render(pixel_t* d_image, int w, int h, int aa, RawConfig* d_cfg) (include/draw.cuh:10,
 * src/draw.cu:215-239) and everything below it (shootPrimaryRay, hitNearest, traverse_lbvh,
 * diffuse/reflection/refraction/GI, checkPlane).  d_rgba: device RGBA8 [h][w] (pixel_t,
 * include/libpng.h:23-27).  Optional outputs: d_hit_id int32 [h][w] = closest hit of sample 0's
 * primary ray (file-order primitive index, -1 miss, -2-p plane p); d_written u8 [h][w] = 1 where
 * a pixel was produced (F3: the reference's aa>=2 launch shape covers only part of the image
 * when aa<32; uncovered pixels are left 0 here, uninitialised in the reference). */
enum {
  RT_SPP_REF          = 0u,   /* aa<=1: render_kernel; aa>=2: 32 samples/pixel over the reference's
                                 launch shape (draw.cu:135-213,226-238)                 */
  RT_SPP_TRUE         = 1u,   /* aa>=2: exactly aa samples for every pixel, seed (1234+pixel, s)   */
  RT_TRAVERSE_NEAR_FIRST = 2u, /* plain nearer-child-first traversal with distance culling.  Same hit as
                                 the reference whenever every hit point lies inside its own AABBs; the
                                 triangle test's -1e-3 barycentric tolerance (struct.cu:150-152) lets a
                                 hit sit just outside, and then the reference's answer depends on its
                                 visiting order.  The DEFAULT lane is also near-first but provably returns
                                 the reference's answer (slack-inflated culling + per-ray certificate +
                                 reference-order fallback); this flag only removes that
                                 machinery, to measure what exactness costs                          */
  RT_ROUGH_ZYX        = 4u,   /* draw order of the three unsequenced standerdD() arguments
                                 (draw.cu:335-337): z,y,x instead of x,y,z                          */
  RT_COUNTERS         = 8u,   /* fill node/leaf counters (slower kernels) and traverse_ms          */
  RT_TIME_TRAVERSE    = 16u,  /* fill traverse_ms (CUDA events around the traverse launches) only  */
  RT_TRAVERSE_REF_ORDER = 32u, /* every ray in the reference's left-first visiting order
                                 (bvh_traversal.cu:149-179): the exact lane's fallback, for all rays */
  RT_RENDER_OVERLAP_BUILD = 64u, /* frame loops (update -> build -> render on ONE stream): run the part of the
                                 frame that does not read the BVH (image clear, primary ray generation of the
                                 first pass) on an internal stream that forks from `stream` at the frame's
                                 first rt_scene_update_triangles / rt_build_lbvh call, i.e. concurrently with
                                 the build.  The caller promises that nothing it enqueued on `stream` after
                                 that call touches the render's output buffers.  Same pixels either way.
                                 rt_scene_begin_frame moves the fork point earlier (e.g. before the caller's
                                 own vertex upload)                                                        */
  RT_RENDER_SERIAL_PASSES = 128u, /* one pass at a time on `stream` alone.  Default: two passes in flight on two
                                 streams with a state pool each, so that the drain of one pass's kernels (its
                                 longest rays) overlaps the other pass's work; traverse_ms then adds up launch
                                 durations that overlap each other.  Same pixels either way.               */
  RT_TRAVERSE_PACKET  = 256u  /* near-first lanes, opt-in: one warp walks the tree ONCE for 32 consecutive rays of the
                                 queue (one position, one stack, every node fetched once per warp) instead of 32
                                 independent lanes.  Same hits (the exact lane's answer does not depend on the
                                 visiting order).  Pays when consecutive rays visit the same nodes
                                 (large triangles); on the 1M-triangle soup only 17 of 32 lanes want the node their
                                 warp fetched and it is 20 % slower than independent lanes, which already get the
                                 cache benefit of the pixel-major queue order (profiles/packet_kernel_r02.json)    */
};
int rt_scene_begin_frame(rt_scene* scene, void* stream);
typedef struct rt_tile_set {    /* image partition for multi-GPU: tile t -> rank t % num_ranks      */
  int32_t tile_w, tile_h, num_ranks, rank;
} rt_tile_set;
typedef struct rt_counters {
  uint64_t rays;            /* hitNearest calls with bounce != 0 (every BVH traversal)              */
  uint64_t samples;         /* primary samples                                                      */
  uint64_t nodes_visited;   /* 64-byte node fetches (RT_COUNTERS)                                   */
  uint64_t leaf_tests;      /* primitive tests (RT_COUNTERS)                                        */
  uint64_t shaded_hits;     /* hits whose material was fetched                                      */
  uint32_t wavefront_iters, passes;
  float    render_ms;       /* device time generate..resolve                                        */
  uint32_t kernel_launches; /* kernels launched by this call                                        */
  float    traverse_ms;     /* device time inside traverse kernels (RT_COUNTERS / RT_TIME_TRAVERSE) */
  uint32_t traverse_launches;
  uint64_t fallback_rays;   /* exact lane: rays re-traced in reference order (certificate failed)   */
  float    d2h_ms;          /* rt_render_host: wall time of the device->host copy of the image (main.cu:73-79) */
  uint32_t reserved0;
  uint64_t packet_steps;    /* node steps of the packet kernel, one per WARP step (RT_COUNTERS): nodes_visited /
                               (32 * packet_steps) = share of the lanes that wanted the node a warp fetched  */
} rt_counters;
int rt_render(rt_scene* scene, int w, int h, int aa, unsigned flags, const rt_tile_set* tiles,
              uint8_t* d_rgba, int32_t* d_hit_id, uint8_t* d_written, rt_counters* counters, void* stream);
/* same, host buffers: device work + device->host copy of the image (main.cu:67-77) */
int rt_render_host(rt_scene* scene, int w, int h, int aa, unsigned flags, const rt_tile_set* tiles,
                   uint8_t* h_rgba, int32_t* h_hit_id, uint8_t* h_written, rt_counters* counters);
/* allocates the device image rt_render_host renders into (cudaMalloc of d_image, main.cu:56-58) ahead of the first frame */
int rt_render_reserve(rt_scene* scene, int w, int h);
/* closest hits for explicit rays (device arrays [n][3] origins and normalised directions):
 * hitNearest for a bounce!=0 ray (draw.cu:292-318).  Outputs device int32[n] / float[n]. */
int rt_trace(rt_scene* scene, int n, const float* d_org, const float* d_dir, unsigned flags,
             int32_t* d_hit_id, float* d_hit_t, rt_counters* counters, void* stream);

/* rt_trace without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs,
 * outputs, flags (lane selection) and results as rt_trace, asynchronous on `stream`; its scratch (packed rays, hit
 * records, fallback list, tickets) lives in one file-scope device array of fixed size instead of buffers allocated on
 * demand, so the call performs no cudaMalloc / cudaFree / host allocation of CUDA memory.  Limit: n <=
 * rt_trace_static_max_rays() (131072 rays, 5.5 MiB held by every GPU that loads the library until exit); larger n
 * returns RT_ERR_INVALID.  All calls on a device share the array: calls that overlap on the GPU race — serialise them. */
int rt_trace_static(rt_scene* scene, int n, const float* d_org, const float* d_dir, unsigned flags,
                    int32_t* d_hit_id, float* d_hit_t, void* stream);
int rt_trace_static_max_rays(void);

/* ---- multi-GPU -------------------------------------------------------------------
 * No counterpart in the reference: it is a single-GPU program (main.cu:25-94, job.slurm:7-11).  The
 * path shards by image tiles: the scene is replicated, every GPU builds the same LBVH itself and
 * renders the tiles t with t % ranks == rank (rt_render's rt_tile_set), the owned tiles are gathered
 * on the root.  rt_comm wraps one NCCL communicator (libnccl.so.2 is loaded when the first
 * communicator is created, not before): either ONE process drives all GPUs (rt_comm_init, the C++
 * host `raytracer --gpus N`), or there is one process per GPU (torchrun: rt_comm_unique_id on one
 * rank, its 128 bytes handed to every rank by the host, rt_comm_init_rank everywhere).  Calls that
 * take arrays (`scenes`, `d_rgba`, `streams`) index them by LOCAL device, 0..rt_comm_num_local-1. */
typedef struct rt_comm rt_comm;
int rt_comm_init(int ndev, const int* devices /* NULL = 0..ndev-1 */, rt_comm** out);
int rt_comm_unique_id(void* id128);
int rt_comm_init_rank(int nranks, int rank, const void* id128, int device, rt_comm** out);
int rt_comm_destroy(rt_comm* comm);
int rt_comm_size(const rt_comm* comm);                 /* GPUs in the communicator            */
int rt_comm_num_local(const rt_comm* comm);            /* GPUs driven by this process         */
int rt_comm_local_rank(const rt_comm* comm, int local);
int rt_comm_local_device(const rt_comm* comm, int local);
/* tile shape for `ranks` GPUs and an image `width` pixels wide: 256 pixels, tiles per row sharing as few
 * factors with `ranks` as possible (else a rank owns whole columns of tiles) */
void rt_tile_shape(int ranks, int width, int* tile_w, int* tile_h);
/* copyConfigDataToDevice (src/config_utils.cu:72-199) on every GPU: the root's arrays (host or device
 * pointers, as for rt_scene_create; root_desc is ignored in processes that do not own `root`) are
 * broadcast over NCCL and each local GPU creates its arena from them.  scenes_out[num_local]. */
int rt_scene_broadcast(rt_comm* comm, const rt_scene_desc* root_desc, int root, rt_scene** scenes_out);
/* rt_scene_update_triangles on every GPU: 36 bytes per triangle from the root over NCCL.  Asynchronous
 * on streams[local] (NULL: default streams). */
int rt_comm_update_triangles(rt_comm* comm, rt_scene* const* scenes, const float* p0, const float* p1,
                             const float* p2, int root, void* const* streams);
int rt_comm_build_lbvh(rt_comm* comm, rt_scene* const* scenes, int morton_bits, unsigned flags, float* ms_out);
/* d_rgba[local]: full-size device image holding the tiles that GPU rendered; afterwards the root's image
 * holds all of them.  Every other GPU sends its own tiles only (1/ranks of the image), packed. */
int rt_framebuffer_gather(rt_comm* comm, uint8_t* const* d_rgba, int w, int h, int tile_w, int tile_h,
                          int root, void* const* streams);
/* render() on every local GPU (one host thread each) + gather on `root` + optional copy of the root's
 * image to h_rgba (main.cu:67-77).  tile_w/tile_h <= 0: rt_tile_shape. */
/* device memory for hosts that do not link the CUDA runtime themselves (the per-GPU images of `raytracer --gpus N`) */
void* rt_device_alloc(int device, size_t bytes);
void  rt_device_free(int device, void* p);
int rt_render_multi(rt_comm* comm, rt_scene* const* scenes, int w, int h, int aa, unsigned flags,
                    int tile_w, int tile_h, int root, uint8_t* const* d_rgba, uint8_t* h_rgba,
                    rt_counters* counters);

/* ---- parity dumps in the reference's exact layouts ---------------------------------------------
 * codes: uint32[n] sorted d_morton_codes (config.hpp:109) (NULL ok); codes64: uint64[n] for
 * morton_bits 63 (NULL ok); refs8: sorted PrimitiveReference[n]; nodes44: LBVHNode[2n-1]
 * (include/lbvh.cuh:6-28; internal i in [0,n-2], leaf k at n-1+k; fields the reference leaves
 * uninitialised are zero; visited_atomic_counter of internal nodes is 2); perm: uint32[n] sorted
 * position -> file-order index.  All HOST pointers. */
int rt_export_reference_layout(rt_scene* scene, uint32_t* codes, uint64_t* codes64, void* refs8,
                               void* nodes44, uint32_t* perm);
int rt_scene_num_prims(const rt_scene* scene);

/* ---- host side of the drop-in surface --------------------------------------------------------- */
typedef struct rt_host_scene rt_host_scene;
/* replaces parseInput(char* argv[], StlConfig&) (include/parse.hpp:10, src/parse.cpp:16-222) */
int  rt_parse_scene_file(const char* path, rt_host_scene** out);
int  rt_host_scene_desc(const rt_host_scene* hs, rt_scene_desc* out);   /* borrows hs's arrays */
const char* rt_host_scene_filename(const rt_host_scene* hs);             /* png W H <name>      */
void rt_host_scene_bounds(const rt_host_scene* hs, float* mn3, float* mx3); /* parse.cpp:28 running bounds */
void rt_host_scene_free(rt_host_scene* hs);
/* Triangle constructor (include/object.cuh:177-191) on the host for n triangles */
void rt_triangle_setup(int n, const float* p0, const float* p1, const float* p2, float* nor, float* e1, float* e2);
/* replaces save_image(image_t*, const char*) (include/libpng.h:63, src/libpng.cpp:73-107): RGBA8,
 * non-interlaced PNG written with zlib */
int  rt_png_write(const char* path, int w, int h, const uint8_t* rgba);
int  rt_png_read(const char* path, int* w, int* h, uint8_t** rgba_out);  /* load_image, libpng.h:53 */
void rt_free(void* p);

/* XORWOW state after curand_init(seed, subsequence, 0) (draw.cu:105,162): v0..v4, d.  Host. */
void rt_xorwow_init(uint64_t seed, uint64_t subsequence, uint32_t* state6);

const char* rt_last_error(void);
const char* rt_version(void);

#ifdef __cplusplus
}
#endif
#endif
