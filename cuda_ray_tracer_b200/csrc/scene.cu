// scene.cu — rt_scene_create / destroy / update: upload in the reference's SoA layout, repack on
// the device into the arena layout described in common.cuh.
// Replaces src/config_utils.cu:18-232 (≤17 cudaMalloc + pageable cudaMemcpy pairs and a leak of the
// SoA component arrays, :205-209) by ONE arena allocation, asynchronous bulk copies staged inside it, one sync.
#include <cstring>
#include <mutex>
#include <vector>

#include "build.cuh"

static thread_local std::string g_err;
void rt_set_error(const std::string& msg) { g_err = msg; }
extern "C" const char* rt_last_error(void) { return g_err.c_str(); }
extern "C" const char* rt_version(void) { return RT_INLINE_SUN ? "b200rt 0.2 (sm_100a) inline-sun" : "b200rt 0.2 (sm_100a)"; }

namespace {

struct PackArgs {
  const float *sc, *sr, *sm, *t0, *t1, *t2, *tn, *te1, *te2, *tm;
  const int2* refs;
  int n;
};

// one thread per file-order primitive: gather the reference SoA record, emit the three views
__global__ void pack_prims_kernel(PackArgs a, float4* bv, float4* tv, float4* mat, uint32_t* tri_prim, int* flags) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  int2 r = a.refs[i];
  const float* m;
  if (r.x == 0) {  // sphere
    const float* c = a.sc + 3 * (size_t)r.y;
    float rad = a.sr[r.y];
    bv[RT_PRIM_STRIDE * (size_t)i + 0] = make_float4(c[0], c[1], c[2], rad);
    bv[RT_PRIM_STRIDE * (size_t)i + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
    bv[RT_PRIM_STRIDE * (size_t)i + 2] = make_float4(0.f, rt_sphere_slack(make_float4(c[0], c[1], c[2], rad)), 0.f, RT_TAG_SPHERE);  // .y = hit slack (build.cuh)
    tv[RT_PRIM_STRIDE * (size_t)i + 0] = make_float4(c[0], c[1], c[2], rad);
    tv[RT_PRIM_STRIDE * (size_t)i + 1] = make_float4(0.f, 0.f, 0.f, 0.f);
    tv[RT_PRIM_STRIDE * (size_t)i + 2] = make_float4(0.f, 0.f, 0.f, 0.f);
    m = a.sm + 11 * (size_t)r.y;
  } else {
    size_t t = (size_t)r.y;
    const float *p0 = a.t0 + 3 * t, *p1 = a.t1 + 3 * t, *p2 = a.t2 + 3 * t, *nn = a.tn + 3 * t, *e1 = a.te1 + 3 * t,
                *e2 = a.te2 + 3 * t;
    bv[RT_PRIM_STRIDE * (size_t)i + 0] = make_float4(p0[0], p0[1], p0[2], p1[0]);
    bv[RT_PRIM_STRIDE * (size_t)i + 1] = make_float4(p1[1], p1[2], p2[0], p2[1]);
    const float4 ba = make_float4(p0[0], p0[1], p0[2], p1[0]), bb = make_float4(p1[1], p1[2], p2[0], p2[1]);
    float4 bc = make_float4(p2[2], 0.f, 0.f, RT_TAG_TRI);
    const float4 ta = make_float4(p0[0], p0[1], p0[2], nn[0]), tb = make_float4(nn[1], nn[2], e1[0], e1[1]),
                 tc = make_float4(e1[2], e2[0], e2[1], e2[2]);
    bc.y = rt_leaf_slack(ba, bb, bc, ta, tb, tc, rt_prim_box(ba, bb, bc));  // hit slack, a property of the triangle setup
    bv[RT_PRIM_STRIDE * (size_t)i + 2] = bc;
    tv[RT_PRIM_STRIDE * (size_t)i + 0] = ta; tv[RT_PRIM_STRIDE * (size_t)i + 1] = tb; tv[RT_PRIM_STRIDE * (size_t)i + 2] = tc;
    tri_prim[t] = (uint32_t)i;
    m = a.tm + 11 * t;
  }
  mat[3 * (size_t)i + 0] = make_float4(m[0], m[1], m[2], m[3]);
  mat[3 * (size_t)i + 1] = make_float4(m[4], m[5], m[6], m[7]);
  mat[3 * (size_t)i + 2] = make_float4(m[8], m[9], m[10], 0.f);
  // scene-wide material flags drive the size of the wavefront continuation stack (render.cu).
  // Same approximate-zero test as RGB::operator== (include/struct.cuh:20-23 via fequal, vec3.cuh:7-18).
  bool shiny = !(fabsf(m[3]) < 1e-6f && fabsf(m[4]) < 1e-6f && fabsf(m[5]) < 1e-6f);
  bool trans = !(fabsf(m[6]) < 1e-6f && fabsf(m[7]) < 1e-6f && fabsf(m[8]) < 1e-6f);
  bool rough = m[10] > 0.0f;
  int f = (shiny ? 1 : 0) | (trans ? 2 : 0) | (rough ? 4 : 0);
  if (f) atomicOr(flags, f);
}

__global__ void pack_planes_kernel(const float* planes21, int np, PlaneRec* out, int* flags) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= np) return;
  const float* p = planes21 + 21 * (size_t)i;
  PlaneRec r;
  for (int k = 0; k < 3; k++) { r.nor[k] = p[4 + k]; r.point[k] = p[7 + k]; }
  for (int k = 0; k < 11; k++) r.mat[k] = p[10 + k];
  r.pad = 0.f;
  out[i] = r;
  const float* m = r.mat;
  bool shiny = !(fabsf(m[3]) < 1e-6f && fabsf(m[4]) < 1e-6f && fabsf(m[5]) < 1e-6f);
  bool trans = !(fabsf(m[6]) < 1e-6f && fabsf(m[7]) < 1e-6f && fabsf(m[8]) < 1e-6f);
  int f = (shiny ? 1 : 0) | (trans ? 2 : 0) | (m[10] > 0.0f ? 4 : 0);
  if (f) atomicOr(flags, f);
}

// Triangle constructor on the device (include/object.cuh:177-191).  Compiled with -fmad=false and
// IEEE div/sqrt, so results equal the g++-compiled host constructor bit for bit.
__device__ inline float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ inline float3 sub3(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ inline float3 cross3(float3 a, float3 b) { return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ inline float dot3(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ inline bool fequal_dev(float a, float b) {
  float diff = fabsf(a - b), largest = fmaxf(fabsf(a), fabsf(b));
  if (largest < 1e-6f) return diff < 1e-6f;
  return diff / largest < 1e-6f;
}
__device__ inline float3 normalize3(float3 a) {
  float mag = sqrtf(a.x * a.x + a.y * a.y + a.z * a.z);
  if (fequal_dev(mag, 0.0f)) return f3(0.f, 0.f, 0.f);
  float inv = 1.0f / mag;
  return f3(a.x * inv, a.y * inv, a.z * inv);
}
__global__ void update_tris_kernel(int nt, const float* p0a, const float* p1a, const float* p2a, const uint32_t* tri_prim,
                                   float4* bv, float4* tv) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nt) return;
  float3 p0 = f3(p0a[3 * t], p0a[3 * t + 1], p0a[3 * t + 2]), p1 = f3(p1a[3 * t], p1a[3 * t + 1], p1a[3 * t + 2]),
         p2 = f3(p2a[3 * t], p2a[3 * t + 1], p2a[3 * t + 2]);
  float3 nor = normalize3(cross3(sub3(p1, p0), sub3(p2, p0)));
  float3 a1 = cross3(sub3(p2, p0), nor), a2 = cross3(sub3(p1, p0), nor);
  float s1 = 1.0f / dot3(a1, sub3(p1, p0)), s2 = 1.0f / dot3(a2, sub3(p2, p0));
  float3 e1 = f3(a1.x * s1, a1.y * s1, a1.z * s1), e2 = f3(a2.x * s2, a2.y * s2, a2.z * s2);
  size_t i = tri_prim[t];
  const float4 ba = make_float4(p0.x, p0.y, p0.z, p1.x), bb = make_float4(p1.y, p1.z, p2.x, p2.y);
  float4 bc = make_float4(p2.z, 0.f, 0.f, RT_TAG_TRI);
  const float4 ta = make_float4(p0.x, p0.y, p0.z, nor.x), tb = make_float4(nor.y, nor.z, e1.x, e1.y), tc = make_float4(e1.z, e2.x, e2.y, e2.z);
  bc.y = rt_leaf_slack(ba, bb, bc, ta, tb, tc, rt_prim_box(ba, bb, bc));  // hit slack travels with the triangle setup
  bv[RT_PRIM_STRIDE * i + 0] = ba; bv[RT_PRIM_STRIDE * i + 1] = bb; bv[RT_PRIM_STRIDE * i + 2] = bc;
  tv[RT_PRIM_STRIDE * i + 0] = ta; tv[RT_PRIM_STRIDE * i + 1] = tb; tv[RT_PRIM_STRIDE * i + 2] = tc;
}

size_t align_up(size_t x) { return (x + 255) & ~(size_t)255; }

// Upload staging.  The reference marshals with up to 17 cudaMalloc + pageable cudaMemcpy pairs
// (src/config_utils.cu:72-199).  Here the caller's SoA arrays are copied with cudaMemcpyAsync into ONE staging
// region — the part of the arena that only the build writes later (sorted leaf records + nodes, 160 B per
// primitive >= the 116 B per triangle / 60 B per sphere of input) — so scene creation makes exactly one device
// allocation, never frees anything, and synchronises once.  Arrays that already live in device memory are read in
// place; pinned host arrays (cudaHostAlloc / cudaHostRegister) go over at full PCIe rate without a bounce copy.
struct Stager {
  char* base = nullptr; size_t cap = 0, off = 0;
  void* overflow = nullptr;  // only when the inputs do not fit the arena scratch (e.g. thousands of planes)
  cudaError_t place(const void* src, size_t bytes, const void** dev) {
    *dev = nullptr;
    if (bytes == 0) return cudaSuccess;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, src) == cudaSuccess && (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged)) { *dev = src; return cudaSuccess; }
    cudaGetLastError();
    const size_t need = (bytes + 255) & ~(size_t)255;
    if (off + need > cap) return cudaErrorMemoryAllocation;
    *dev = base + off;
    off += need;
    return cudaMemcpyAsync(const_cast<void*>(*dev), src, bytes, cudaMemcpyHostToDevice, 0);
  }
  ~Stager() { if (overflow) cudaFree(overflow); }
};

}  // namespace

extern "C" int rt_scene_num_prims(const rt_scene* s) { return s ? s->n : 0; }

extern "C" int rt_scene_create(const rt_scene_desc* d, int device, rt_scene** out) {
  if (!d || !out) { rt_set_error("rt_scene_create: null argument"); return RT_ERR_INVALID; }
  *out = nullptr;
  if (d->num_prims != d->num_spheres + d->num_triangles || d->num_prims < 0) {
    rt_set_error("rt_scene_create: num_prims != num_spheres + num_triangles"); return RT_ERR_INVALID;
  }
  if ((unsigned)d->num_prims > RT_LEAF_MASK) { rt_set_error("rt_scene_create: more than 2^30-1 primitives"); return RT_ERR_INVALID; }
  int ndev = 0;
  RT_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) { rt_set_error("rt_scene_create: no such CUDA device"); return RT_ERR_CUDA; }
  RT_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  RT_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {  // kernels are built for sm_100a only; there is no fallback path
    rt_set_error(std::string("rt_scene_create: device is ") + prop.name + " (sm_" + std::to_string(prop.major) + std::to_string(prop.minor) +
                 "), this library is built for sm_100a only");
    return RT_ERR_CUDA;
  }
  rt_scene* s = new rt_scene();
  s->device = device; s->cfg = *d;
  s->n = d->num_prims; s->ns = d->num_spheres; s->nt = d->num_triangles;
  // null the borrowed pointers
  s->cfg.sphere_c = s->cfg.sphere_r = s->cfg.sphere_mat = nullptr;
  s->cfg.tri_p0 = s->cfg.tri_p1 = s->cfg.tri_p2 = s->cfg.tri_nor = s->cfg.tri_e1 = s->cfg.tri_e2 = s->cfg.tri_mat = nullptr;
  s->cfg.refs = nullptr; s->cfg.suns = s->cfg.bulbs = s->cfg.planes = nullptr;
  const size_t n = (size_t)s->n, n1 = n > 0 ? n : 1;
  const size_t tiles_max = n1 / 2048 + 2;  // onesweep tiles are >= 2048 keys
  size_t off = 0;
  auto carve = [&](size_t bytes) { size_t o = off; off += align_up(bytes ? bytes : 16); return o; };
  size_t o_bv = carve(n1 * 16 * RT_PRIM_STRIDE), o_tv = carve(n1 * 16 * RT_PRIM_STRIDE), o_mat = carve(n1 * 48), o_stv = carve(n1 * 16 * RT_TV_STRIDE), o_nodes = carve(n1 * 64);
  size_t o_wn = carve(n1 * 128);
  size_t o_refs = carve(n1 * 8), o_tp = carve((size_t)(s->nt > 0 ? s->nt : 1) * 4);
  size_t o_sun = carve((size_t)d->num_suns * sizeof(LightSun)), o_bulb = carve((size_t)d->num_bulbs * sizeof(LightBulb)),
         o_pl = carve((size_t)d->num_planes * sizeof(PlaneRec));
  size_t o_ka = carve(n1 * 8), o_kb = carve(n1 * 8), o_pa = carve(n1 * 4), o_pb = carve(n1 * 4);
  size_t o_hist = carve(8 * 256 * 4), o_lb = carve(8 * tiles_max * 256 * 4), o_tk = carve(64 * 4);
  size_t o_flags = carve(n1 * 4), o_par = carve(2 * n1 * 4), o_kar = carve(n1 * 4), o_bnd = carve(64);
  s->arena_bytes = off;
  cudaError_t e = cudaMalloc((void**)&s->arena, s->arena_bytes);
  if (e != cudaSuccess) { rt_set_error(std::string("rt_scene_create: arena cudaMalloc: ") + cudaGetErrorString(e)); delete s; return RT_ERR_NOMEM; }
  char* A = s->arena;
  s->prim_bv = (float4*)(A + o_bv); s->prim_tv = (float4*)(A + o_tv); s->prim_mat = (float4*)(A + o_mat);
  s->sorted_tv = (float4*)(A + o_stv); s->nodes = (float4*)(A + o_nodes); s->wnodes = (float4*)(A + o_wn); s->refs = (int2*)(A + o_refs);
  s->tri_prim = (uint32_t*)(A + o_tp); s->suns = (LightSun*)(A + o_sun); s->bulbs = (LightBulb*)(A + o_bulb);
  s->planes = (PlaneRec*)(A + o_pl); s->keys_a = (uint64_t*)(A + o_ka); s->keys_b = (uint64_t*)(A + o_kb);
  s->perm_a = (uint32_t*)(A + o_pa); s->perm_b = (uint32_t*)(A + o_pb); s->hist = (uint32_t*)(A + o_hist);
  s->lookback = (uint32_t*)(A + o_lb); s->tickets = (uint32_t*)(A + o_tk); s->node_flags = (int*)(A + o_flags);
  s->parent = (int*)(A + o_par); s->node_karras = (int*)(A + o_kar); s->bounds = (float*)(A + o_bnd);
  auto fail = [&](cudaError_t err, const char* what) {
    rt_set_error(std::string("rt_scene_create: ") + what + ": " + cudaGetErrorString(err));
    cudaFree(s->arena); delete s; return RT_ERR_CUDA;
  };
  {
    const size_t ns = (size_t)s->ns, nt = (size_t)s->nt, np = (size_t)d->num_planes;
    Stager S;
    S.base = A + o_stv; S.cap = (o_nodes + n1 * 64) - o_stv;
    const size_t need = ns * 60 + nt * 116 + np * 84 + 12 * 256;
    if (need > S.cap) {  // does not happen for primitive data; a scene with very many planes
      if ((e = cudaMalloc(&S.overflow, need))) return fail(e, "staging");
      S.base = (char*)S.overflow; S.cap = need;
    }
    const void *sc, *sr, *sm, *t0, *t1, *t2, *tn, *te1, *te2, *tm, *planes;
    if ((e = S.place(d->sphere_c, ns * 12, &sc)) || (e = S.place(d->sphere_r, ns * 4, &sr)) || (e = S.place(d->sphere_mat, ns * 44, &sm)) ||
        (e = S.place(d->tri_p0, nt * 12, &t0)) || (e = S.place(d->tri_p1, nt * 12, &t1)) || (e = S.place(d->tri_p2, nt * 12, &t2)) ||
        (e = S.place(d->tri_nor, nt * 12, &tn)) || (e = S.place(d->tri_e1, nt * 12, &te1)) || (e = S.place(d->tri_e2, nt * 12, &te2)) ||
        (e = S.place(d->tri_mat, nt * 44, &tm)) || (e = S.place(d->planes, np * 84, &planes)))
      return fail(e, "upload");
    int* flagbuf = (int*)(A + o_bnd) + 32;  // spare bytes of the (256-byte aligned) bounds block
    if ((e = cudaMemsetAsync(flagbuf, 0, 4, 0))) return fail(e, "flags");
    if (n && (e = cudaMemcpyAsync(s->refs, d->refs, n * 8, cudaMemcpyDefault, 0))) return fail(e, "refs");
    if (d->num_suns && (e = cudaMemcpyAsync(s->suns, d->suns, (size_t)d->num_suns * 24, cudaMemcpyDefault, 0))) return fail(e, "suns");
    if (d->num_bulbs && (e = cudaMemcpyAsync(s->bulbs, d->bulbs, (size_t)d->num_bulbs * 24, cudaMemcpyDefault, 0))) return fail(e, "bulbs");
    if (n) {
      PackArgs a{(const float*)sc, (const float*)sr, (const float*)sm, (const float*)t0, (const float*)t1,
                 (const float*)t2, (const float*)tn, (const float*)te1, (const float*)te2, (const float*)tm, s->refs, s->n};
      pack_prims_kernel<<<(unsigned)((n + 255) / 256), 256>>>(a, s->prim_bv, s->prim_tv, s->prim_mat, s->tri_prim, flagbuf);
    }
    if (d->num_planes) pack_planes_kernel<<<(d->num_planes + 63) / 64, 64>>>((const float*)planes, d->num_planes, s->planes, flagbuf);
    int f = 0;
    if ((e = cudaGetLastError()) || (e = cudaMemcpyAsync(&f, flagbuf, 4, cudaMemcpyDeviceToHost, 0)) || (e = cudaDeviceSynchronize())) return fail(e, "pack kernels");
    s->any_shiny = f & 1; s->any_trans = f & 2; s->any_rough = f & 4;
  }
  if ((e = cudaEventCreate(&s->ev0)) || (e = cudaEventCreate(&s->ev1)) || (e = cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming)))
    return fail(e, "events");
  *out = s;
  return RT_OK;
}

extern "C" int rt_scene_destroy(rt_scene* s) {
  if (!s) return RT_OK;
  cudaSetDevice(s->device);
  rt_render_free_state(s);
  if (s->ev0) cudaEventDestroy(s->ev0);
  if (s->ev1) cudaEventDestroy(s->ev1);
  if (s->ev_fork) cudaEventDestroy(s->ev_fork);
  for (int k = 0; k < 3; k++) cudaFree(s->upd_stage[k]);
  cudaFree(s->arena);
  delete s;
  return RT_OK;
}

// RT_RENDER_OVERLAP_BUILD: remember the point in `st` before this frame's first update/build call.  rt_render may start the
// BVH-independent head of the frame (image clear, primary ray generation) from there on a side stream.
int rt_arm_fork(rt_scene* s, cudaStream_t st) {
  if (s->fork_armed && s->fork_stream == st) return RT_OK;
  RT_CUDA(cudaEventRecord(s->ev_fork, st));
  s->fork_stream = st; s->fork_armed = true;
  return RT_OK;
}

extern "C" int rt_scene_begin_frame(rt_scene* s, void* stream) {
  if (!s) { rt_set_error("rt_scene_begin_frame: null scene"); return RT_ERR_INVALID; }
  RT_CUDA(cudaSetDevice(s->device));
  s->fork_armed = false;
  return rt_arm_fork(s, (cudaStream_t)stream);
}

extern "C" int rt_scene_update_triangles(rt_scene* s, const float* p0, const float* p1, const float* p2, void* stream) {
  if (!s || !p0 || !p1 || !p2) { rt_set_error("rt_scene_update_triangles: null argument"); return RT_ERR_INVALID; }
  if (s->nt == 0) return RT_OK;
  RT_CUDA(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  rt_arm_fork(s, st);
  const float* src[3] = {p0, p1, p2};
  const float* dev[3];
  bool staged = false;
  for (int k = 0; k < 3; k++) {
    cudaPointerAttributes at;
    cudaError_t e = cudaPointerGetAttributes(&at, src[k]);
    if (e == cudaSuccess && at.type == cudaMemoryTypeDevice) { dev[k] = src[k]; continue; }
    cudaGetLastError();
    if (!s->upd_stage[k]) RT_CUDA(cudaMalloc((void**)&s->upd_stage[k], (size_t)s->nt * 12));  // kept for the next frame
    RT_CUDA(cudaMemcpyAsync(s->upd_stage[k], src[k], (size_t)s->nt * 12, cudaMemcpyDefault, st));
    dev[k] = s->upd_stage[k];
    staged = true;
  }
  update_tris_kernel<<<(s->nt + 255) / 256, 256, 0, st>>>(s->nt, dev[0], dev[1], dev[2], s->tri_prim, s->prim_bv, s->prim_tv);
  RT_CUDA(cudaGetLastError());
  s->built = false;
  (void)staged;
  return RT_OK;
}
