// oracle/ref_host_driver.cu — drives the REFERENCE's own traversal/shading code on the host.
// TEST INFRASTRUCTURE ONLY.  Compiled by oracle/Makefile together with the reference sources
// (in place under /root/reference, -include ref_prelude.h) into oracle/_ref/libref_host.so.
// Nothing here restates reference arithmetic: parseInput, hitNearest, shootPrimaryRay, randD,
// RGBtosRGB, curand_* are the reference's / cuRAND's own symbols.  Only the per-pixel loops of
// render_kernel / render_kernel_warp_aa (src/draw.cu:94-213, __global__ and therefore not
// host-callable) are re-driven here, and the K1..K6 build result is supplied by the caller.
#include <atomic>
#include <cstring>
#include <thread>
#include <vector>

#include "config.hpp"
#include "draw.cuh"
#include "helper.cuh"
#include "bvh_traversal.cuh"
#include "parse.hpp"
#include "lbvh.cuh"

struct ref_scene {
  StlConfig stl;
  RawConfig rc;
  std::vector<point3> sc; std::vector<float> sr; std::vector<Materials> sm;
  std::vector<point3> tp0, tp1, tp2, te1, te2; std::vector<vec3> tnor; std::vector<Materials> tm;
  SphereDataSoA ssoa; TriangleDataSoA tsoa;
  std::vector<PrimitiveReference> refs;
  std::vector<LBVHNode> nodes;
  AABB bounds;
};

static void fill_raw(ref_scene* r) {
  // host mirror of initRawConfigFromStl + copyConfigDataToDevice (src/config_utils.cu:18-199),
  // with the "device" pointers pointing at host vectors
  StlConfig& s = r->stl; RawConfig& c = r->rc;
  memset(&c, 0, sizeof(c));
  c.width = s.width; c.height = s.height; c.color = s.color; c.bounces = s.bounces; c.aa = s.aa;
  c.dof_focus = s.dof_focus; c.dof_lens = s.dof_lens; c.forward = s.forward; c.right = s.right; c.up = s.up;
  c.eye = s.eye; c.target_up = s.target_up; c.expose = s.expose; c.fisheye = s.fisheye; c.panorama = s.panorama;
  c.ior = s.ior; c.rough = s.rough; c.gi = s.gi; c.trans = s.trans; c.shine = s.shine;
  c.num_spheres = (int)s.host_spheres_data.size(); c.num_triangles = (int)s.host_triangles_data.size();
  c.num_total_primitives = (int)s.host_primitive_references.size();
  for (auto& p : s.host_spheres_data) { r->sc.push_back(p.c); r->sr.push_back(p.r); r->sm.push_back(p.mat); }
  for (auto& t : s.host_triangles_data) {
    r->tp0.push_back(t.p0); r->tp1.push_back(t.p1); r->tp2.push_back(t.p2); r->tnor.push_back(t.nor);
    r->te1.push_back(t.e1); r->te2.push_back(t.e2); r->tm.push_back(t.mat);
  }
  r->ssoa.c = r->sc.data(); r->ssoa.r = r->sr.data(); r->ssoa.mat = r->sm.data();
  r->tsoa.p0 = r->tp0.data(); r->tsoa.p1 = r->tp1.data(); r->tsoa.p2 = r->tp2.data(); r->tsoa.nor = r->tnor.data();
  r->tsoa.e1 = r->te1.data(); r->tsoa.e2 = r->te2.data(); r->tsoa.mat = r->tm.data();
  c.d_spheres_soa = &r->ssoa; c.d_triangles_soa = &r->tsoa;
  r->refs = s.host_primitive_references;
  c.d_primitive_references = r->refs.data();
  c.num_sun = (int)s.sun_data.size(); c.d_all_suns = s.sun_data.data();
  c.num_bulbs = (int)s.bulb_data.size(); c.d_all_bulbs = s.bulb_data.data();
  c.num_planes = (int)s.plane_data.size(); c.d_all_planes = s.plane_data.data();
}

extern "C" {

ref_scene* ref_load(const char* path) {
  ref_scene* r = new ref_scene();
  char* argv[2] = {(char*)"raytracer", (char*)path};
  parseInput(argv, r->stl);  // exits the process on a bad file, like the reference
  fill_raw(r);
  return r;
}
void ref_free(ref_scene* r) { delete r; }

// Programmatic scenes (synthetic soups for bench.py --impl reference): the same host objects parse.cpp
// builds for `xyz`/`tri`/`sphere`/`sun` lines (src/parse.cpp:130-216), made with the reference's own
// constructors (Triangle(Vertex,Vertex,Vertex,RGB): include/object.cuh:165-191), without the text detour.
ref_scene* ref_new(int width, int height, int bounces, int aa, const float* eye3, const float* forward3) {
  ref_scene* r = new ref_scene();
  StlConfig& s = r->stl;
  s.width = width; s.height = height; s.bounces = bounces; s.aa = aa;
  s.eye = point3(eye3[0], eye3[1], eye3[2]);
  s.forward = vec3(forward3[0], forward3[1], forward3[2]);   // parse.cpp:61-65
  s.right = cross(s.forward, s.up).normalize();
  s.up = cross(s.right, s.forward).normalize();
  return r;
}
void ref_add_triangles(ref_scene* r, int n, const float* p0, const float* p1, const float* p2, const float* mat11) {
  StlConfig& s = r->stl;
  for (int i = 0; i < n; i++) {
    const float* m = mat11 + 11 * (size_t)i;
    Triangle t(Vertex(p0[3 * i], p0[3 * i + 1], p0[3 * i + 2]), Vertex(p1[3 * i], p1[3 * i + 1], p1[3 * i + 2]),
               Vertex(p2[3 * i], p2[3 * i + 1], p2[3 * i + 2]), RGB(m[0], m[1], m[2]));
    t.mat.shininess = RGB(m[3], m[4], m[5]); t.mat.trans = RGB(m[6], m[7], m[8]); t.mat.ior = m[9]; t.mat.roughness = m[10];
    s.host_triangles_data.push_back(t);
    s.host_primitive_references.push_back(PrimitiveReference(PrimitiveType::TRIANGLE, (unsigned)s.host_triangles_data.size() - 1));
  }
}
void ref_add_sun(ref_scene* r, const float* dir3, const float* color3) {
  r->stl.sun_data.push_back(Sun(dir3[0], dir3[1], dir3[2], RGB(color3[0], color3[1], color3[2])));
}
void ref_finish(ref_scene* r) { fill_raw(r); }

void ref_counts(ref_scene* r, int* out6) {
  out6[0] = r->rc.num_spheres; out6[1] = r->rc.num_triangles; out6[2] = r->rc.num_total_primitives;
  out6[3] = r->rc.num_sun; out6[4] = r->rc.num_bulbs; out6[5] = r->rc.num_planes;
}
void ref_config(ref_scene* r, int* ints8, float* f32) {
  RawConfig& c = r->rc;
  ints8[0] = c.width; ints8[1] = c.height; ints8[2] = c.bounces; ints8[3] = c.aa; ints8[4] = c.gi;
  ints8[5] = c.fisheye; ints8[6] = c.panorama; ints8[7] = 0;
  int k = 0;
  f32[k++] = c.color.r; f32[k++] = c.color.g; f32[k++] = c.color.b; f32[k++] = c.dof_focus; f32[k++] = c.dof_lens;
  vec3* vs[5] = {&c.forward, &c.right, &c.up, &c.eye, &c.target_up};
  for (auto v : vs) { f32[k++] = v->x; f32[k++] = v->y; f32[k++] = v->z; }
  f32[k++] = c.expose; f32[k++] = c.ior; f32[k++] = c.rough;
  f32[k++] = c.trans.r; f32[k++] = c.trans.g; f32[k++] = c.trans.b;
  f32[k++] = c.shine.r; f32[k++] = c.shine.g; f32[k++] = c.shine.b;
  while (k < 32) f32[k++] = 0.f;
}
// parsed arrays in reference layout, for parser parity
const float* ref_tri(ref_scene* r, int which) {
  switch (which) { case 0: return &r->tp0[0].x; case 1: return &r->tp1[0].x; case 2: return &r->tp2[0].x;
    case 3: return &r->tnor[0].x; case 4: return &r->te1[0].x; default: return &r->te2[0].x; }
}
const float* ref_tri_mat(ref_scene* r) { return (const float*)r->tm.data(); }
const float* ref_sphere_c(ref_scene* r) { return &r->sc[0].x; }
const float* ref_sphere_r(ref_scene* r) { return r->sr.data(); }
const float* ref_sphere_mat(ref_scene* r) { return (const float*)r->sm.data(); }
const int* ref_refs(ref_scene* r) { return (const int*)r->stl.host_primitive_references.data(); }
const float* ref_planes(ref_scene* r) { return (const float*)r->stl.plane_data.data(); }
const float* ref_suns(ref_scene* r) { return (const float*)r->stl.sun_data.data(); }
const float* ref_bulbs(ref_scene* r) { return (const float*)r->stl.bulb_data.data(); }

// K1..K6 results (built elsewhere) in the reference's own LBVHNode layout (include/lbvh.cuh:6-28)
void ref_set_bvh(ref_scene* r, int n, const int* sorted_refs2, const float* bbox6, const unsigned* left, const unsigned* right) {
  r->refs.resize((size_t)n);
  for (int k = 0; k < n; k++) r->refs[k] = PrimitiveReference((PrimitiveType)sorted_refs2[2 * k], (unsigned)sorted_refs2[2 * k + 1]);
  r->rc.d_primitive_references = r->refs.data();
  r->nodes.assign((size_t)(2 * n - 1), LBVHNode());
  for (int i = 0; i < 2 * n - 1; i++) {
    LBVHNode& nd = r->nodes[i];
    const float* b = bbox6 + 6 * (size_t)i;
    nd.bbox = AABB(Interval(b[0], b[1]), Interval(b[2], b[3]), Interval(b[4], b[5]));
    if (i < n - 1) { nd.left_child_offset = left[i]; nd.right_child_offset = right[i]; nd.num_primitives_in_leaf = 0; nd.primitive_offset = 0; }
    else { nd.left_child_offset = 0; nd.right_child_offset = 0; nd.num_primitives_in_leaf = 1; nd.primitive_offset = (unsigned)(i - (n - 1)); }
  }
  r->rc.d_lbvh_nodes = r->nodes.data(); r->rc.num_lbvh_nodes = 2 * n - 1;
}

static void pixel_ref(RawConfig* cfg, int W, int H, int aa, int pix, pixel_t* out) {
  const int w = pix % W, h = pix / W;
  if (aa <= 1) {  // render_kernel, draw.cu:94-133
    curandState state; curand_init(1234, pix, 0, &state);
    RGBA rgba;
    if (aa == 0) rgba = shootPrimaryRay((float)w, (float)h, &state, cfg);
    else {
      RGBA acc;
      for (int i = 0; i < aa; i++) {
        float nw = w + randD(-0.5f, 0.5f, &state);
        float nh = h + randD(-0.5f, 0.5f, &state);
        acc = acc + shootPrimaryRay(nw, nh, &state, cfg);
      }
      rgba = acc.mean(aa);
    }
    out->r = RGBtosRGB(rgba.r) * 255; out->g = RGBtosRGB(rgba.g) * 255; out->b = RGBtosRGB(rgba.b) * 255; out->a = rgba.a * 255;
    return;
  }
  // render_kernel_warp_aa, draw.cu:135-213: 32 lanes, butterfly sum
  float v[4][32];
  for (int lane = 0; lane < 32; lane++) {
    curandState state; curand_init(1234 + pix, lane, 0, &state);
    float jx = randD(-0.5f, 0.5f, &state), jy = randD(-0.5f, 0.5f, &state);
    RGBA s = shootPrimaryRay((float)w + jx, (float)h + jy, &state, cfg);
    v[0][lane] = s.r; v[1][lane] = s.g; v[2][lane] = s.b; v[3][lane] = s.a;
  }
  for (int m = 16; m > 0; m /= 2)
    for (int ch = 0; ch < 4; ch++) {
      float nv[32];
      for (int i = 0; i < 32; i++) nv[i] = v[ch][i] + v[ch][i ^ m];
      memcpy(v[ch], nv, sizeof(nv));
    }
  const float inv = 1.0f / 32.0f;
  auto u8 = [](float f) { return (unsigned char)(fminf(fmaxf(f, 0.0f), 1.0f) * 255.0f + 0.5f); };
  out->r = u8(RGBtosRGB(v[0][0] * inv)); out->g = u8(RGBtosRGB(v[1][0] * inv));
  out->b = u8(RGBtosRGB(v[2][0] * inv)); out->a = u8(v[3][0] * inv);
}

// rows [y0,y1) of the image with the reference's render() semantics (F3 launch shape included)
void ref_render(ref_scene* r, int W, int H, int aa, int y0, int y1, int threads, unsigned char* rgba, unsigned char* written) {
  RawConfig cfg = r->rc; cfg.width = W; cfg.height = H; cfg.aa = aa;
  const long long P = (long long)W * H;
  long long covered = P;
  if (aa >= 2) { long long grid = (P * aa + 127) / 128; covered = grid * 4 < P ? grid * 4 : P; }
  // work items of 16 consecutive pixels (not rows: a band of 4 rows would keep 4 threads busy and leave the rest idle)
  const long long p_begin = (long long)y0 * W, p_end = (long long)y1 * W;
  std::atomic<long long> next(p_begin);
  auto work = [&]() {
    RawConfig local = cfg;
    for (;;) {
      const long long b = next.fetch_add(16);
      if (b >= p_end) break;
      for (long long pp = b; pp < (b + 16 < p_end ? b + 16 : p_end); pp++) {
        int pix = (int)pp;
        if (pix >= covered) { if (written) written[pix] = 0; continue; }
        pixel_ref(&local, W, H, aa, pix, (pixel_t*)(rgba + 4 * (size_t)pix));
        if (written) written[pix] = 1;
      }
    }
  };
  if (threads <= 1) work();
  else { std::vector<std::thread> th; for (int t = 0; t < threads; t++) th.emplace_back(work); for (auto& t : th) t.join(); }
}

// Closest-hit primitive ids out of the UNMODIFIED hitNearest: ObjectInfo carries no id
// (include/object.cuh:44-56), so every primitive's colour is replaced by a tag that encodes it
// (traversal never reads materials), and the tag of the returned hit is decoded.
void ref_trace_ids(ref_scene* r, int n, const float* org3, const float* dir3, int* ids, float* ts) {
  std::vector<Materials> sm = r->sm, tm = r->tm; std::vector<Plane> pl = r->stl.plane_data;
  auto tag = [](Materials& m, int kind, unsigned id) { m.color = RGB((float)kind, (float)(id & 0xFFFF), (float)(id >> 16)); };
  for (size_t i = 0; i < sm.size(); i++) tag(sm[i], 1, (unsigned)i);
  for (size_t i = 0; i < tm.size(); i++) tag(tm[i], 2, (unsigned)i);
  for (size_t i = 0; i < pl.size(); i++) tag(pl[i].mat, 3, (unsigned)i);
  SphereDataSoA ss = r->ssoa; ss.mat = sm.data();
  TriangleDataSoA ts_ = r->tsoa; ts_.mat = tm.data();
  RawConfig cfg = r->rc; cfg.d_spheres_soa = &ss; cfg.d_triangles_soa = &ts_; cfg.d_all_planes = pl.data();
  // file-order index of (type,id): position in the unsorted reference list
  std::vector<int> sphere_pos(sm.size()), tri_pos(tm.size());
  const auto& fr = r->stl.host_primitive_references;
  for (size_t i = 0; i < fr.size(); i++) {
    if (fr[i].type == PrimitiveType::SPHERE) sphere_pos[fr[i].id_in_type_array] = (int)i; else tri_pos[fr[i].id_in_type_array] = (int)i;
  }
  for (int i = 0; i < n; i++) {
    Ray ray; ray.eye = point3(org3[3 * i], org3[3 * i + 1], org3[3 * i + 2]);
    ray.dir = vec3(dir3[3 * i], dir3[3 * i + 1], dir3[3 * i + 2]); ray.bounce = 1;
    ObjectInfo o = hitNearest(ray, &cfg);
    int id = -1;
    if (o.isHit) {
      int kind = (int)o.mat.color.r; unsigned pid = (unsigned)o.mat.color.g | ((unsigned)o.mat.color.b << 16);
      id = kind == 1 ? sphere_pos[pid] : kind == 2 ? tri_pos[pid] : -2 - (int)pid;
    }
    ids[i] = id;
    if (ts) ts[i] = o.isHit ? o.distance : INFINITY;
  }
}

// primary-ray ids for a W x H image at exact pixel coordinates (aa == 0 semantics, draw.cu:111-114)
void ref_primary_rays(ref_scene* r, int W, int H, float* org3, float* dir3) {
  RawConfig cfg = r->rc; cfg.width = W; cfg.height = H;
  for (int pix = 0; pix < W * H; pix++) {
    curandState state; curand_init(1234, pix, 0, &state);
    Ray ray((float)(pix % W), (float)(pix / W), &state, &cfg);
    org3[3 * pix] = ray.eye.x; org3[3 * pix + 1] = ray.eye.y; org3[3 * pix + 2] = ray.eye.z;
    dir3[3 * pix] = ray.dir.x; dir3[3 * pix + 1] = ray.dir.y; dir3[3 * pix + 2] = ray.dir.z;
  }
}

// cuRAND's own host path: state words after curand_init, then n raw draws, uniforms and normals
void ref_curand(unsigned long long seed, unsigned long long subseq, unsigned* state6, int n, unsigned* raw, float* uni, float* nrm) {
  curandState st; curand_init(seed, subseq, 0, &st);
  for (int i = 0; i < 5; i++) state6[i] = st.v[i];
  state6[5] = st.d;
  curandState a = st, b = st, c = st;
  for (int i = 0; i < n; i++) { if (raw) raw[i] = curand(&a); if (uni) uni[i] = curand_uniform(&b); if (nrm) nrm[i] = curand_normal(&c); }
}

}  // extern "C"
