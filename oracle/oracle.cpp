// oracle/oracle.cpp — CPU oracle.  TEST INFRASTRUCTURE ONLY (see oracle.h).
//
// Sequential restatement of the reference hot path.  Compile with -ffp-contract=off so that no
// a*b+c is fused: the bit-exact lane ("lane A") is IEEE fp32 without contraction on
// every side.  Parity status: pinned against (1) the reference's N<=16 node dump known answers
// for tri.txt, (2) the reference's own traversal/shading sources compiled for the
// host (oracle/_ref/libref_host.so, built by oracle/Makefile from /root/reference in place) and
// (3) cuRAND's own host path for XORWOW — see tests/test_oracle_vs_ref.py and tests/golden/.
//
// Citations are file:line in /root/reference.
#include "oracle.h"

#include <algorithm>
#include <atomic>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

namespace {

// ---------------------------------------------------------------------------------------------
// vec3 / fequal / normalize — include/vec3.cuh:7-107
// ---------------------------------------------------------------------------------------------
struct V3 { float x = 0.f, y = 0.f, z = 0.f; };
inline V3 mk(float a, float b, float c) { V3 v; v.x = a; v.y = b; v.z = c; return v; }
inline V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
inline V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return mk(a.x * s, a.y * s, a.z * s); }
inline V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float length(V3 a) { return std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }

// vec3.cuh:7-18
inline bool fequal(float a, float b, float epsilon = 1e-6f) {
  float diff = fabsf(a - b);
  float largest = fmaxf(fabsf(a), fabsf(b));
  if (largest < 1e-6f) return diff < epsilon;
  return diff / largest < epsilon;
}
// vec3.cuh:72-82
inline V3 normalize(V3 a) {
  float mag = length(a);
  if (fequal(mag, 0.0f)) return mk(0.f, 0.f, 0.f);
  float inv_mag = 1.0f / mag;
  return mk(a.x * inv_mag, a.y * inv_mag, a.z * inv_mag);
}

// struct.cuh:12-62
struct RGB3 { float r = 0.f, g = 0.f, b = 0.f; };
inline RGB3 rgb(float r, float g, float b) { RGB3 c; c.r = r; c.g = g; c.b = b; return c; }
inline bool rgb_eq(RGB3 a, RGB3 b) { return fequal(a.r, b.r) && fequal(a.g, b.g) && fequal(a.b, b.b); }
inline RGB3 operator-(RGB3 a, RGB3 b) { return rgb(a.r - b.r, a.g - b.g, a.b - b.b); }
inline RGB3 operator*(RGB3 a, RGB3 b) { return rgb(a.r * b.r, a.g * b.g, a.b * b.b); }
struct RGBA4 { float r = 0.f, g = 0.f, b = 0.f, a = 0.f; };
inline RGBA4 rgba(float r, float g, float b, float a) { RGBA4 c; c.r = r; c.g = g; c.b = b; c.a = a; return c; }
inline RGBA4 operator+(RGBA4 x, RGBA4 y) { return rgba(x.r + y.r, x.g + y.g, x.b + y.b, x.a + y.a); }
inline RGBA4 operator*(RGB3 c, RGBA4 o) { return rgba(c.r * o.r, c.g * o.g, c.b * o.b, o.a); }

// object.cuh:17-38 — 44 bytes: color, shininess, trans, ior, roughness
struct Mat {
  RGB3 color, shin, trans;
  float ior = 1.458f, rough = 0.f;
};
static_assert(sizeof(Mat) == 44, "Materials layout");

struct Sphere { V3 c; float r; Mat m; };
struct Tri { V3 p0, p1, p2, nor, e1, e2; Mat m; };
struct Plane { float a, b, c, d; V3 nor, point; Mat m; };
static_assert(sizeof(Plane) == 84, "Plane layout");
struct Sun { V3 dir; RGB3 color; };
struct Bulb { V3 point; RGB3 color; };
struct Ref { int32_t type; uint32_t id; };  // object.cuh:72-88  (0 sphere, 1 triangle)

// interval.cuh:46-104
struct Box { float xmin = INFINITY, xmax = -INFINITY, ymin = INFINITY, ymax = -INFINITY, zmin = INFINITY, zmax = -INFINITY; };
inline Box box2(V3 a, V3 b) {  // interval.cuh:55-59
  Box o;
  if (a.x <= b.x) { o.xmin = a.x; o.xmax = b.x; } else { o.xmin = b.x; o.xmax = a.x; }
  if (a.y <= b.y) { o.ymin = a.y; o.ymax = b.y; } else { o.ymin = b.y; o.ymax = a.y; }
  if (a.z <= b.z) { o.zmin = a.z; o.zmax = b.z; } else { o.zmin = b.z; o.zmax = a.z; }
  return o;
}
inline Box box3(V3 a, V3 b, V3 c) {  // interval.cuh:61-81
  Box o;
  o.xmin = std::fmin(std::fmin(a.x, b.x), c.x); o.xmax = std::fmax(std::fmax(a.x, b.x), c.x);
  o.ymin = std::fmin(std::fmin(a.y, b.y), c.y); o.ymax = std::fmax(std::fmax(a.y, b.y), c.y);
  o.zmin = std::fmin(std::fmin(a.z, b.z), c.z); o.zmax = std::fmax(std::fmax(a.z, b.z), c.z);
  if (o.xmax - o.xmin < 0.01f) { o.xmin = o.xmin - 0.01f; o.xmax = o.xmax + 0.01f; }
  if (o.ymax - o.ymin < 0.01f) { o.ymin = o.ymin - 0.01f; o.ymax = o.ymax + 0.01f; }
  if (o.zmax - o.zmin < 0.01f) { o.zmin = o.zmin - 0.01f; o.zmax = o.zmax + 0.01f; }
  return o;
}
inline Box box_union(const Box& a, const Box& b) {  // interval.cuh:83-88
  Box o;
  o.xmin = std::fmin(a.xmin, b.xmin); o.xmax = std::fmax(a.xmax, b.xmax);
  o.ymin = std::fmin(a.ymin, b.ymin); o.ymax = std::fmax(a.ymax, b.ymax);
  o.zmin = std::fmin(a.zmin, b.zmin); o.zmax = std::fmax(a.zmax, b.zmax);
  return o;
}

// object.cuh:177-191 (Triangle constructor; compiled by g++ in the reference, Makefile:38-39)
inline void tri_setup(Tri& t) {
  t.nor = normalize(cross(t.p1 - t.p0, t.p2 - t.p0));
  V3 a1 = cross(t.p2 - t.p0, t.nor);
  V3 a2 = cross(t.p1 - t.p0, t.nor);
  t.e1 = (1 / (dot(a1, t.p1 - t.p0))) * a1;
  t.e2 = (1 / (dot(a2, t.p2 - t.p0))) * a2;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// scene — include/config.hpp:26-72 (StlConfig)
// ---------------------------------------------------------------------------------------------
struct o_scene {
  int width = 0, height = 0;
  std::string filename = "file.txt";
  RGB3 color = rgb(1.f, 1.f, 1.f);
  int bounces = 4, aa = 0;
  float dof_focus = 0.f, dof_lens = 0.f;
  V3 forward = mk(0.f, 0.f, -1.f), right = mk(1.f, 0.f, 0.f), up = mk(0.f, 1.f, 0.f), eye = mk(0.f, 0.f, 0.f),
     target_up = mk(0.f, 1.f, 0.f);
  float expose = INFINITY;
  bool fisheye = false, panorama = false;
  float ior = 1.458f, rough = 0.f;
  int gi = 0;
  RGB3 trans, shine;

  std::vector<Sphere> spheres;
  std::vector<Tri> tris;
  std::vector<Ref> refs;  // file order
  std::vector<Sun> suns;
  std::vector<Bulb> bulbs;
  std::vector<Plane> planes;
  std::vector<V3> verts;
  Box running;  // parse.cpp:28 — computed, then dropped by the reference (F1)

  // SoA mirrors (reference layout) for handing the same inputs to the CUDA path
  std::vector<float> s_c, s_r, s_m, t_soa[6], t_m, sun_f, bulb_f, plane_f;
  std::vector<int32_t> refs_i;
  bool soa_valid = false;

  // build outputs
  int built_bits = 0;
  std::vector<uint32_t> codes30;
  std::vector<uint64_t> codes63;
  std::vector<uint32_t> perm;
  std::vector<Ref> srefs;
  std::vector<int32_t> srefs_i;
  std::vector<Box> nbox;
  std::vector<float> nbox_f;
  std::vector<uint32_t> nleft, nright;
  std::vector<int32_t> nparent;
};

namespace {

void add_bounds(o_scene& s, const Box& b) {  // parse.cpp:150-155,195-200
  s.running.xmin = fminf(s.running.xmin, b.xmin); s.running.xmax = fmaxf(s.running.xmax, b.xmax);
  s.running.ymin = fminf(s.running.ymin, b.ymin); s.running.ymax = fmaxf(s.running.ymax, b.ymax);
  s.running.zmin = fminf(s.running.zmin, b.zmin); s.running.zmax = fmaxf(s.running.zmax, b.zmax);
}

Mat current_mat(const o_scene& s) {
  Mat m; m.color = s.color; m.shin = s.shine; m.trans = s.trans; m.ior = s.ior; m.rough = s.rough; return m;
}

void push_sphere(o_scene& s, V3 c, float r, const Mat& m) {  // parse.cpp:130-156
  Sphere sp; sp.c = c; sp.r = r; sp.m = m;
  s.spheres.push_back(sp);
  Ref rf; rf.type = 0; rf.id = (uint32_t)s.spheres.size() - 1; s.refs.push_back(rf);
  V3 rv = mk(r, r, r);
  add_bounds(s, box2(c - rv, c + rv));
  s.soa_valid = false;
}
void push_tri(o_scene& s, V3 a, V3 b, V3 c, const Mat& m) {  // parse.cpp:174-201
  Tri t; t.p0 = a; t.p1 = b; t.p2 = c; t.m = m; tri_setup(t);
  s.tris.push_back(t);
  Ref rf; rf.type = 1; rf.id = (uint32_t)s.tris.size() - 1; s.refs.push_back(rf);
  add_bounds(s, box3(a, b, c));
  s.soa_valid = false;
}
void push_plane(o_scene& s, float a, float b, float c, float d, const Mat& m) {  // object.cuh:136-141
  Plane p; p.a = a; p.b = b; p.c = c; p.d = d;
  p.nor = normalize(mk(a, b, c));
  p.point = (-d * mk(a, b, c)) / (float)(std::pow((double)a, 2) + std::pow((double)b, 2) + std::pow((double)c, 2));
  p.m = m;
  s.planes.push_back(p);
  s.soa_valid = false;
}

// src/parse.cpp:41-222
bool parse_line(o_scene& s, const std::vector<std::string>& w) {
  if (w.empty()) return true;
  auto F = [&](int i) { return std::stof(w[i]); };
  auto I = [&](int i) { return std::stoi(w[i]); };
  const std::string& k = w[0];
  size_t n = w.size();
  if (k == "png" && n == 4) { s.width = I(1); s.height = I(2); s.filename = w[3]; }
  else if (k == "bounces" && n == 2) s.bounces = I(1);
  else if (k == "forward" && n == 4) {
    s.forward = mk(F(1), F(2), F(3));
    s.right = normalize(cross(s.forward, s.up));
    s.up = normalize(cross(s.right, s.forward));
  } else if (k == "up" && n == 4) {
    s.target_up = mk(F(1), F(2), F(3));
    s.right = normalize(cross(s.forward, s.target_up));
    s.up = normalize(cross(s.right, s.forward));
  } else if (k == "eye" && n == 4) s.eye = mk(F(1), F(2), F(3));
  else if (k == "expose" && n == 2) s.expose = F(1);
  else if (k == "dof" && n == 3) { s.dof_focus = F(1); s.dof_lens = F(2); }
  else if (k == "aa" && n == 2) s.aa = I(1);
  else if (k == "panorama" && n == 1) s.panorama = true;
  else if (k == "fisheye" && n == 1) s.fisheye = true;
  else if (k == "gi" && n == 2) s.gi = I(1);
  else if (k == "color" && n == 4) s.color = rgb(F(1), F(2), F(3));
  else if (k == "roughness" && n == 2) s.rough = F(1);
  else if (k == "shininess" && n == 2) s.shine = rgb(F(1), F(1), F(1));
  else if (k == "shininess" && n == 4) s.shine = rgb(F(1), F(2), F(3));
  else if (k == "transparency" && n == 2) s.trans = rgb(F(1), F(1), F(1));
  else if (k == "transparency" && n == 4) s.trans = rgb(F(1), F(2), F(3));
  else if (k == "ior" && n == 2) s.ior = F(1);
  else if (k == "sphere" && n == 5) push_sphere(s, mk(F(1), F(2), F(3)), F(4), current_mat(s));
  else if (k == "plane" && n == 5) push_plane(s, F(1), F(2), F(3), F(4), current_mat(s));
  else if (k == "xyz" && n == 4) s.verts.push_back(mk(F(1), F(2), F(3)));
  else if (k == "tri" && n == 4) {
    // parse.cpp:175-180: indices are held in *float* variables (exact below 2^24 vertices)
    float i, j, kk;
    int size = (int)s.verts.size();
    i = (I(1) > 0) ? I(1) - 1 : size + I(1);
    j = (I(2) > 0) ? I(2) - 1 : size + I(2);
    kk = (I(3) > 0) ? I(3) - 1 : size + I(3);
    push_tri(s, s.verts[(size_t)i], s.verts[(size_t)j], s.verts[(size_t)kk], current_mat(s));
  } else if (k == "sun" && n == 4) { Sun u; u.dir = mk(F(1), F(2), F(3)); u.color = s.color; s.suns.push_back(u); }
  else if (k == "bulb" && n == 4) { Bulb u; u.point = mk(F(1), F(2), F(3)); u.color = s.color; s.bulbs.push_back(u); }
  else return false;
  return true;
}

void put3(std::vector<float>& v, V3 a) { v.push_back(a.x); v.push_back(a.y); v.push_back(a.z); }
void putm(std::vector<float>& v, const Mat& m) {
  const float* f = reinterpret_cast<const float*>(&m);
  v.insert(v.end(), f, f + 11);
}
void ensure_soa(o_scene* sc) {  // de-interleave, src/config_utils.cu:80-133
  if (sc->soa_valid) return;
  o_scene& s = *sc;
  s.s_c.clear(); s.s_r.clear(); s.s_m.clear(); s.t_m.clear(); s.sun_f.clear(); s.bulb_f.clear(); s.plane_f.clear();
  for (auto& v : s.t_soa) v.clear();
  s.refs_i.clear();
  for (auto& p : s.spheres) { put3(s.s_c, p.c); s.s_r.push_back(p.r); putm(s.s_m, p.m); }
  for (auto& t : s.tris) {
    put3(s.t_soa[0], t.p0); put3(s.t_soa[1], t.p1); put3(s.t_soa[2], t.p2);
    put3(s.t_soa[3], t.nor); put3(s.t_soa[4], t.e1); put3(s.t_soa[5], t.e2);
    putm(s.t_m, t.m);
  }
  for (auto& r : s.refs) { s.refs_i.push_back(r.type); s.refs_i.push_back((int32_t)r.id); }
  for (auto& u : s.suns) { put3(s.sun_f, u.dir); s.sun_f.push_back(u.color.r); s.sun_f.push_back(u.color.g); s.sun_f.push_back(u.color.b); }
  for (auto& u : s.bulbs) { put3(s.bulb_f, u.point); s.bulb_f.push_back(u.color.r); s.bulb_f.push_back(u.color.g); s.bulb_f.push_back(u.color.b); }
  for (auto& p : s.planes) {
    s.plane_f.push_back(p.a); s.plane_f.push_back(p.b); s.plane_f.push_back(p.c); s.plane_f.push_back(p.d);
    put3(s.plane_f, p.nor); put3(s.plane_f, p.point); putm(s.plane_f, p.m);
  }
  s.soa_valid = true;
}

// ---------------------------------------------------------------------------------------------
// K1 Morton codes — src/lbvh_utils.cu:10-75
// ---------------------------------------------------------------------------------------------
inline uint32_t expand_bits10(uint32_t v) {  // lbvh_utils.cu:10-17
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
inline uint64_t expand_bits21(uint64_t v) {  // 21-bit generalisation (H9; not in the reference)
  v &= 0x1FFFFFull;
  v = (v | (v << 32)) & 0x1F00000000FFFFull;
  v = (v | (v << 16)) & 0x1F0000FF0000FFull;
  v = (v | (v << 8)) & 0x100F00F00F00F00Full;
  v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
  v = (v | (v << 2)) & 0x1249249249249249ull;
  return v;
}
inline uint32_t quantize(float coord, float mn, float range, int bits) {  // lbvh_utils.cu:24-30
  if (range <= 1e-6f) return 0;
  float normalized = (coord - mn) / range;
  normalized = fmaxf(0.0f, fminf(1.0f, normalized));
  return static_cast<uint32_t>(normalized * ((1 << bits) - 1));
}
V3 centroid(const o_scene& s, const Ref& r) {  // lbvh_utils.cu:48-59
  if (r.type == 0) return s.spheres[r.id].c;
  const Tri& t = s.tris[r.id];
  return (t.p0 + t.p1 + t.p2) / 3.0f;
}
void get_bounds(const o_scene& s, int mode, float* mn, float* mx) {
  if (mode == O_BOUNDS_COMPUTED) {
    mn[0] = s.running.xmin; mn[1] = s.running.ymin; mn[2] = s.running.zmin;
    mx[0] = s.running.xmax; mx[1] = s.running.ymax; mx[2] = s.running.zmax;
  } else {  // config.hpp:66 default AABB, never assigned (parse.cpp:28-39): min=+INF max=-INF
    mn[0] = mn[1] = mn[2] = INFINITY; mx[0] = mx[1] = mx[2] = -INFINITY;
  }
}
uint64_t morton_of(const o_scene& s, const Ref& r, const float* mn, const float* mx, int bits_total) {
  V3 c = centroid(s, r);
  float rx = mx[0] - mn[0], ry = mx[1] - mn[1], rz = mx[2] - mn[2];  // lbvh_utils.cu:66-68
  if (bits_total == 30) {
    uint32_t qx = quantize(c.x, mn[0], rx, 10), qy = quantize(c.y, mn[1], ry, 10), qz = quantize(c.z, mn[2], rz, 10);
    return expand_bits10(qx) | (expand_bits10(qy) << 1) | (expand_bits10(qz) << 2);  // lbvh_utils.cu:19-22
  }
  uint64_t qx = quantize(c.x, mn[0], rx, 21), qy = quantize(c.y, mn[1], ry, 21), qz = quantize(c.z, mn[2], rz, 21);
  return expand_bits21(qx) | (expand_bits21(qy) << 1) | (expand_bits21(qz) << 2);
}

// ---------------------------------------------------------------------------------------------
// K4 Karras hierarchy — src/lbvh_builder.cu:76-322 (literal)
// ---------------------------------------------------------------------------------------------
struct Keys {
  const uint32_t* k30 = nullptr; const uint64_t* k63 = nullptr; unsigned n = 0;
};
inline int clz32(uint32_t x) { return x ? __builtin_clz(x) : 32; }
inline int clz64(uint64_t x) { return x ? __builtin_clzll(x) : 64; }
inline int delta(int a, int b, const Keys& K) {  // lbvh_builder.cu:76-101
  bool ia = (a < 0 || a >= (int)K.n), ib = (b < 0 || b >= (int)K.n);
  if (ia || ib) return -1;
  if (K.k30) {
    uint32_t ka = K.k30[a], kb = K.k30[b];
    if (ka == kb) return 32 + clz32((uint32_t)a ^ (uint32_t)b);
    return clz32(ka ^ kb);
  }
  uint64_t ka = K.k63[a], kb = K.k63[b];
  if (ka == kb) return 64 + clz32((uint32_t)a ^ (uint32_t)b);
  return clz64(ka ^ kb);
}
inline void determine_range(const Keys& K, int i, int& lo, int& hi) {  // lbvh_builder.cu:103-182
  const int dl = delta(i, i - 1, K), dr = delta(i, i + 1, K);
  if (K.n <= 1) { lo = hi = i; return; }
  int d, dmin;
  if (dr > dl) { d = 1; dmin = dl; } else { d = -1; dmin = dr; }
  unsigned l_max = 1;
  int nb = i + (int)l_max * d;
  int cur = delta(i, nb, K);
  while (cur > dmin) {
    l_max <<= 1;
    nb = i + (int)l_max * d;
    if (nb < 0 || nb >= (int)K.n) break;
    cur = delta(i, nb, K);
  }
  unsigned l = 0;
  for (unsigned t = l_max >> 1; t > 0; t >>= 1) {
    int nbb = i + (int)(l + t) * d;
    if (nbb >= 0 && nbb < (int)K.n) {
      if (delta(i, nbb, K) > dmin) l += t;
    }
  }
  const int j = i + (int)l * d;
  if (i < j) { lo = i; hi = j; } else { lo = j; hi = i; }
}
inline int find_split(const Keys& K, int first, int last) {  // lbvh_builder.cu:186-221
  if (first == last) return first;
  const int common = delta(first, last, K);
  int split = first, step = last - first;
  do {
    step = (step + 1) >> 1;
    const int ns = split + step;
    if (ns < last) {
      if (delta(first, ns, K) > common) split = ns;
    }
  } while (step > 1);
  return split;
}

// ---------------------------------------------------------------------------------------------
// XORWOW — cuRAND curand_kernel.h (_curand_init_inplace, _skipahead_sequence_inplace, curand),
// curand_uniform.h:69-72, curand_normal.h:70-87,313-326.  Call sites draw.cu:105,162; helper.cu:82-101.
// ---------------------------------------------------------------------------------------------
struct Rng {
  uint32_t v[5]; uint32_t d; int bm_flag; float bm_extra;
};
inline uint32_t xw_next(Rng& s) {
  uint32_t t = (s.v[0] ^ (s.v[0] >> 2));
  s.v[0] = s.v[1]; s.v[1] = s.v[2]; s.v[2] = s.v[3]; s.v[3] = s.v[4];
  s.v[4] = (s.v[4] ^ (s.v[4] << 4)) ^ (t ^ (t << 1));
  s.d += 362437;
  return s.v[4] + s.d;
}
// 160x160 GF(2) matrix as 160 images of basis vectors (same convention as cuRAND's precalc matrices)
struct Gf2 { uint32_t img[160][5]; };
inline void gf2_apply(const Gf2& M, const uint32_t* in, uint32_t* out) {
  uint32_t r[5] = {0, 0, 0, 0, 0};
  for (int i = 0; i < 5; i++)
    for (int j = 0; j < 32; j++)
      if (in[i] & (1u << j))
        for (int k = 0; k < 5; k++) r[k] ^= M.img[i * 32 + j][k];
  for (int k = 0; k < 5; k++) out[k] = r[k];
}
struct SkipTables {
  // pw[j] = T^(2^67 * 2^j): one subsequence step is 2^67 draws (curand_kernel.h skipahead_sequence)
  std::vector<Gf2> pw;
  SkipTables() {
    Gf2 T;
    for (int b = 0; b < 160; b++) {
      Rng s; memset(&s, 0, sizeof(s));
      s.v[b / 32] = 1u << (b % 32);
      xw_next(s);
      for (int k = 0; k < 5; k++) T.img[b][k] = s.v[k];
    }
    auto square = [](const Gf2& A) { Gf2 R; for (int b = 0; b < 160; b++) gf2_apply(A, A.img[b], R.img[b]); return R; };
    Gf2 P = T;
    for (int i = 0; i < 67; i++) P = square(P);
    pw.push_back(P);
    for (int j = 1; j < 40; j++) pw.push_back(square(pw.back()));
  }
};
const SkipTables& skip_tables() { static SkipTables t; return t; }

inline void xw_init(Rng& s, unsigned long long seed, unsigned long long subseq) {
  uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
  uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
  uint32_t t0 = 1099087573u * s0;
  uint32_t t1 = 2591861531u * s1;
  s.d = 6615241u + t1 + t0;
  s.v[0] = 123456789u + t0;
  s.v[1] = 362436069u ^ t0;
  s.v[2] = 521288629u + t1;
  s.v[3] = 88675123u ^ t1;
  s.v[4] = 5783321u + t0;
  const SkipTables& T = skip_tables();
  for (int j = 0; subseq && j < 40; j++, subseq >>= 1)
    if (subseq & 1) gf2_apply(T.pw[j], s.v, s.v);
  s.bm_flag = 0; s.bm_extra = 0.f;
}
inline float xw_uniform(Rng& s) {  // curand_uniform.h:69-72
  uint32_t x = xw_next(s);
  return x * 2.3283064e-10f + (2.3283064e-10f / 2.0f);
}
inline float xw_normal(Rng& s) {  // curand_normal.h:313-326 + 70-87 (host branch: sinf/cosf)
  if (s.bm_flag != 1) {
    uint32_t x = xw_next(s), y = xw_next(s);
    float u = x * 2.3283064e-10f + (2.3283064e-10f / 2);
    float v = y * (2.3283064e-10f * 6.2831855f) + ((2.3283064e-10f * 6.2831855f) / 2);
    float sq = sqrtf(-2.0f * logf(u));
    float rx = sinf(v), ry = cosf(v);
    rx *= sq; ry *= sq;
    s.bm_extra = ry; s.bm_flag = 1;
    return rx;
  }
  s.bm_flag = 0;
  return s.bm_extra;
}
inline float randD(float a, float b, Rng& s) { float u = xw_uniform(s); return a + (b - a) * u; }  // helper.cu:82-85
inline float standerdD(float sd, Rng& s) { return xw_normal(s) * sd; }                             // helper.cu:87-89
// draw.cu:335-337,395-397: vec3(standerdD(..), standerdD(..), standerdD(..)) — the three draws are
// unsequenced constructor arguments.  g_args_rtl=0: x,y,z order (left to right); 1: z,y,x (what g++
// does for the host build of the reference).  Which one the GPU build takes is measured on the box
// against oracle/_ref/ref_gpu_parity (tests/test_gpu_vs_ref_gpu.py) and recorded in DESIGN.md.
int g_args_rtl = 0;
inline V3 rough_offset(float rough, Rng& st) {
  float a, b, cc;
  if (g_args_rtl) { cc = standerdD(rough, st); b = standerdD(rough, st); a = standerdD(rough, st); }
  else { a = standerdD(rough, st); b = standerdD(rough, st); cc = standerdD(rough, st); }
  return mk(a, b, cc);
}
inline V3 spherePoint(Rng& s) {                                                                  // helper.cu:91-101
  float z = 2.0f * randD(0.0f, 1.0f, s) - 1.0f;
  float theta = 2.0f * 3.14159265f * randD(0.0f, 1.0f, s);
  float r = std::sqrt(1.0f - z * z);
  float x = r * std::cos(theta);
  float y = r * std::sin(theta);
  return mk(x, y, z);
}

// ---------------------------------------------------------------------------------------------
// traversal + intersection — src/bvh_traversal.cu, src/struct.cu
// ---------------------------------------------------------------------------------------------
struct Ray { V3 eye, dir; int bounce = 0; };
inline Ray mkray(V3 eye, V3 dir, int bounce) { Ray r; r.eye = eye; r.dir = normalize(dir); r.bounce = bounce; return r; }  // object.cuh:69
struct Hit {  // object.cuh:44-56 ObjectInfo (+ id, which the reference does not carry)
  bool isHit = false; float distance = -1.0f; V3 point, normal; Mat mat; int32_t id = -1;
  Hit() { mat.color = rgb(0, 0, 0); }  // Materials() default: RGB() = 0
};

struct Ctx {
  const o_scene* s; o_counters cnt;
  Ctx(const o_scene* sc) : s(sc) { memset(&cnt, 0, sizeof(cnt)); }
};

inline bool hit_aabb(const Box& b, V3 o, V3 inv, float tmin, float tmax) {  // bvh_traversal.cu:11-44
  float tx1 = (b.xmin - o.x) * inv.x, tx2 = (b.xmax - o.x) * inv.x;
  float tnx = fminf(tx1, tx2), tfx = fmaxf(tx1, tx2);
  float ty1 = (b.ymin - o.y) * inv.y, ty2 = (b.ymax - o.y) * inv.y;
  float tny = fminf(ty1, ty2), tfy = fmaxf(ty1, ty2);
  float tz1 = (b.zmin - o.z) * inv.z, tz2 = (b.zmax - o.z) * inv.z;
  float tnz = fminf(tz1, tz2), tfz = fmaxf(tz1, tz2);
  float t_enter = fmaxf(fmaxf(tnx, tny), tnz);
  float t_exit = fminf(fminf(tfx, tfy), tfz);
  return t_enter < t_exit && t_enter < tmax && t_exit > tmin;
}
inline Hit hit_sphere(const Ray& ray, const Sphere& sp) {  // struct.cu:64-109
  V3 cr0 = sp.c - ray.eye;
  bool inside = (dot(cr0, cr0) < sp.r * sp.r);
  float tc = dot(cr0, ray.dir);
  if (!inside && tc < 0.0f) return Hit();
  V3 dv = ray.eye + (tc * ray.dir) - sp.c;
  float d2 = dot(dv, dv);
  if (!inside && (sp.r * sp.r) < d2) return Hit();
  float toff = std::sqrt((sp.r * sp.r) - d2);
  float t = inside ? tc + toff : tc - toff;
  V3 p = t * ray.dir + ray.eye;
  V3 nor = inside ? (sp.c - p) : (p - sp.c);
  nor = normalize(nor);
  Hit h; h.isHit = true; h.distance = t; h.point = p; h.normal = nor; h.mat = sp.m;
  return h;
}
inline Hit hit_tri(const Ray& ray, const Tri& t) {  // struct.cu:111-163
  float den = dot(ray.dir, t.nor);
  if (fabsf(den) < 1e-9f) return Hit();
  float ti = dot((t.p0 - ray.eye), t.nor) / den;
  if (ti <= 0.001f) return Hit();
  V3 p = ti * ray.dir + ray.eye;
  float b1 = dot(t.e1, p - t.p0);
  float b2 = dot(t.e2, p - t.p0);
  float b0 = 1.0f - b1 - b2;
  bool inside = (b0 >= -0.001f) && (b1 >= -0.001f) && (b2 >= -0.001f);
  if (!inside) return Hit();
  Hit h; h.isHit = true; h.distance = ti; h.point = p; h.normal = (den < 0.0f) ? t.nor : -t.nor; h.mat = t.m;
  return h;
}

Hit traverse(Ctx& c, const Ray& ray, float tmax0) {  // bvh_traversal.cu:92-183
  const o_scene& s = *c.s;
  Hit best; best.distance = tmax0;
  float tmax = tmax0;
  const int n = (int)s.refs.size();
  if (s.nbox.empty()) return best;
  V3 inv = mk(1.0f / ray.dir.x, 1.0f / ray.dir.y, 1.0f / ray.dir.z);
  const float tmin = 0.0001f;
  unsigned stack[64]; int sp = 0;
  unsigned cur = 0;
  while (true) {
    c.cnt.nodes_visited++;
    if ((int)cur >= n - 1) {  // leaf (nodes n-1 .. 2n-2; lbvh_builder.cu:59-71)
      unsigned k = cur - (unsigned)(n - 1);
      const Ref& r = s.srefs[k];
      c.cnt.leaf_tests++;
      Hit h = (r.type == 0) ? hit_sphere(ray, s.spheres[r.id]) : hit_tri(ray, s.tris[r.id]);
      if (h.isHit && h.distance > 1e-6f && h.distance < tmax) {  // bvh_traversal.cu:80-86
        tmax = h.distance; best = h; best.id = (int32_t)s.perm[k];
      }
      if (sp == 0) break;
      cur = stack[--sp];
      continue;
    }
    unsigned L = s.nleft[cur], R = s.nright[cur];
    bool hl = hit_aabb(s.nbox[L], ray.eye, inv, tmin, tmax);
    bool hr = hit_aabb(s.nbox[R], ray.eye, inv, tmin, tmax);
    if (hl && hr) {
      cur = L;
      if (sp < 64) stack[sp++] = R;
      if (sp > c.cnt.max_stack) c.cnt.max_stack = sp;
    } else if (hl) cur = L;
    else if (hr) cur = R;
    else { if (sp == 0) break; cur = stack[--sp]; }
  }
  return best;
}

Hit check_plane(Ctx& c, const Ray& ray) {  // draw.cu:581-615
  const o_scene& s = *c.s;
  float t_sol = INFINITY; V3 p_sol, nor; Mat mats; mats.color = rgb(0, 0, 0); int which = -1;
  for (size_t i = 0; i < s.planes.size(); i++) {
    const Plane& pl = s.planes[i];
    float t = dot((pl.point - ray.eye), pl.nor) / (dot(ray.dir, pl.nor));
    if (t <= 1e-6f) continue;
    V3 ip = t * ray.dir + ray.eye;
    if (t < t_sol && t > 0.001f) {
      t_sol = t; p_sol = ip;
      nor = (dot(pl.nor, ray.dir) < 0.0f) ? pl.nor : -pl.nor;
      mats = pl.m; which = (int)i;
    }
  }
  if (t_sol >= (float)(INT_MAX - 10)) return Hit();
  Hit h; h.isHit = true; h.distance = t_sol; h.point = p_sol; h.normal = nor; h.mat = mats; h.id = -2 - which;
  return h;
}

Hit hit_nearest(Ctx& c, const Ray& ray) {  // draw.cu:292-318
  if (ray.bounce == 0) return Hit();
  c.cnt.rays++;
  Hit b = traverse(c, ray, INFINITY);
  Hit p = check_plane(c, ray);
  if (b.isHit && p.isHit) return (b.distance < p.distance) ? b : p;
  if (b.isHit) return b;
  if (p.isHit) return p;
  return Hit();
}

// ---------------------------------------------------------------------------------------------
// shading — src/draw.cu:260-659, src/helper.cu:12-45
// ---------------------------------------------------------------------------------------------
inline float RGBtosRGB(float l) {  // helper.cu:12-27
  float sol;
  if (l < 0.0031308f) sol = 12.92f * l;
  else sol = 1.055f * std::pow(l, 1 / 2.4f) - 0.055f;
  sol = std::fmin(1.0f, std::fmax(0.0f, sol));
  return sol;
}
inline float setExpose(float cc, const o_scene& s) {  // helper.cu:40-45 (double subtraction)
  return s.expose == float(INFINITY) ? cc : 1.0 - std::exp(-s.expose * cc);
}
inline RGBA4 colorSun(float lambert, RGB3 oc, RGB3 lc, const o_scene& s) {  // draw.cu:627-636
  float r = oc.r * (lc.r * lambert), g = oc.g * (lc.g * lambert), b = oc.b * (lc.b * lambert);
  return rgba(setExpose(r, s), setExpose(g, s), setExpose(b, s), 0.0f);
}
inline RGBA4 colorBulb(float lambert, RGB3 oc, RGB3 lc, float t, const o_scene& s) {  // draw.cu:649-659
  float i = 1.0f / std::pow((double)t, 2);
  float r = oc.r * (lc.r * lambert), g = oc.g * (lc.g * lambert), b = oc.b * (lc.b * lambert);
  return rgba(setExpose(r, s) * i, setExpose(g, s) * i, setExpose(b, s) * i, 0.0f);
}

RGBA4 reflectionLight(Ctx& c, const Ray& ray, const Hit& obj, Rng& st);
RGBA4 refractionLight(Ctx& c, const Ray& ray, const Hit& obj, Rng& st);

RGBA4 diffuseLight(Ctx& c, const Hit& obj, Rng& st) {  // draw.cu:329-377
  const o_scene& s = *c.s;
  c.cnt.shaded_hits++;
  RGBA4 color;
  V3 normal = obj.normal;
  if (obj.mat.rough > 0.0f) {
    normal = normal + rough_offset(obj.mat.rough, st);
  }
  normal = normalize(normal);
  for (size_t i = 0; i < s.suns.size(); i++) {
    const Sun& light = s.suns[i];
    Ray sh = mkray(obj.point + obj.normal * 0.001f, light.dir, 1);
    Hit info = hit_nearest(c, sh);
    if (info.isHit) continue;
    float lambert = std::fmax(dot(normal, normalize(light.dir)), 0.0f);
    color = color + colorSun(lambert, obj.mat.color, light.color, s);
  }
  for (size_t i = 0; i < s.bulbs.size(); i++) {
    const Bulb& light = s.bulbs[i];
    V3 bd = light.point - obj.point;
    Ray sh = mkray(obj.point + obj.normal * 0.001f, bd, 1);
    Hit info = hit_nearest(c, sh);
    if (info.isHit && info.distance < length(bd)) continue;
    float lambert = std::fmax(dot(normal, normalize(bd)), 0.0f);
    color = color + colorBulb(lambert, obj.mat.color, light.color, length(bd), s);
  }
  return color;
}

RGBA4 reflectionLight(Ctx& c, const Ray& ray, const Hit& obj, Rng& st) {  // draw.cu:389-437
  if (rgb_eq(obj.mat.shin, rgb(0, 0, 0)) || ray.bounce <= 0) return RGBA4();
  V3 normal = obj.normal;
  if (obj.mat.rough > 0.0f) {
    normal = normal + rough_offset(obj.mat.rough, st);
  }
  normal = normalize(normal);
  V3 rdir = ray.dir - 2.0f * (dot(normal, ray.dir)) * normal;
  Ray second = mkray(obj.point + obj.normal * 0.001f, rdir, ray.bounce - 1);
  Hit so = hit_nearest(c, second);
  RGB3 shine, trans;
  if (ray.bounce == 1) { shine = rgb(0, 0, 0); trans = rgb(0, 0, 0); }
  else { shine = so.mat.shin; trans = so.mat.trans; }
  RGBA4 color;
  if (so.isHit) {
    RGBA4 diffuse = diffuseLight(c, so, st);
    RGBA4 reflect = reflectionLight(c, second, so, st);
    RGBA4 refract = refractionLight(c, ray, obj, st);  // outer args, draw.cu:424
    RGB3 one = rgb(1.f, 1.f, 1.f);
    color = shine * reflect + (one - shine) * trans * refract + (one - shine) * (one - trans) * diffuse;
  } else color = rgba(0.f, 0.f, 0.f, 1.f);
  return color;
}

RGBA4 refractionLight(Ctx& c, const Ray& ray, const Hit& obj, Rng& st) {  // draw.cu:456-527
  if (rgb_eq(obj.mat.trans, rgb(0, 0, 0)) || ray.bounce <= 0) return RGBA4();
  V3 rdir; Ray inside_ray, final_ray;
  float ior = 1.0f / obj.mat.ior;
  V3 dir = ray.dir;
  V3 normal = normalize(obj.normal);
  V3 ip = obj.point;
  int bounce = ray.bounce;
  float k = 1.0f - std::pow((double)ior, 2) * (1.0f - std::pow((double)dot(normal, dir), 2));
  if (k < 0) {
    rdir = dir - 2.0f * (dot(normal, dir)) * normal;
    final_ray = mkray(ip + normal * 0.001f, rdir, --bounce);
  } else {
    rdir = ior * dir - (ior * (dot(normal, dir)) + std::sqrt(k)) * normal;
    inside_ray = mkray(ip - normal * 0.0001f, rdir, bounce);
    Hit other = hit_nearest(c, inside_ray);
    normal = normalize(other.normal);
    ior = other.mat.ior;
    dir = inside_ray.dir;
    ip = other.point;
    k = 1.0f - ior * ior * (1.0f - std::pow((double)dot(normal, dir), 2));
    rdir = ior * dir - (ior * (dot(normal, dir)) + std::sqrt(k)) * normal;
    final_ray = mkray(ip - normal * 0.0001f, rdir, --bounce);
  }
  Hit fo = hit_nearest(c, final_ray);
  RGB3 shine, trans;
  if (bounce == 0) { shine = rgb(0, 0, 0); trans = rgb(0, 0, 0); }
  else { shine = fo.mat.shin; trans = fo.mat.trans; }
  RGBA4 color;
  if (fo.isHit) {
    RGBA4 diffuse = diffuseLight(c, fo, st);
    RGBA4 reflect = reflectionLight(c, final_ray, fo, st);
    RGBA4 refract = refractionLight(c, final_ray, fo, st);
    RGB3 one = rgb(1.f, 1.f, 1.f);
    color = shine * reflect + (one - shine) * trans * refract + (one - shine) * (one - trans) * diffuse;
  } else color = rgba(0.f, 0.f, 0.f, 1.f);
  return color;
}

RGBA4 globalIllumination(Ctx& c, const Hit& obj, int gi_bounce, Rng& st) {  // draw.cu:540-568
  const o_scene& s = *c.s;
  if (s.gi == 0 || gi_bounce == 0) return RGBA4();
  V3 normal = obj.normal, ip = obj.point;
  V3 gdir = normalize(normal + spherePoint(st));
  Ray gray = mkray(ip + normal * 0.001f, gdir, gi_bounce - 1);
  Hit go = hit_nearest(c, gray);
  RGBA4 color;
  if (go.isHit) {
    RGBA4 diffuse = diffuseLight(c, go, st);
    RGBA4 reflect = reflectionLight(c, gray, go, st);
    RGBA4 refract = refractionLight(c, gray, go, st);
    RGBA4 gic = go.mat.color * globalIllumination(c, go, gi_bounce - 1, st);
    RGB3 one = rgb(1.f, 1.f, 1.f);
    color = go.mat.shin * reflect + (one - go.mat.shin) * go.mat.trans * refract +
            (one - go.mat.shin) * (one - go.mat.trans) * (diffuse + gic);
    color.a = 1.0f;
  }
  return color;
}

// struct.cu:16-62
Ray primary_ray(const o_scene& s, int W, int H, float x, float y, Rng& st) {
  Ray r;
  float max_dim = std::fmax((float)W, (float)H);
  float sx = (2.0f * x - W) / max_dim;
  float sy = (H - 2.0f * y) / max_dim;
  r.eye = s.eye;
  V3 dir;
  if (s.fisheye) {
    dir = (sx * s.right + sy * s.up) + (float)std::sqrt(1.0f - std::pow((double)sx, 2) - std::pow((double)sy, 2)) * s.forward;
  } else if (s.panorama) {
    sx = (float)x / (float)W;
    sy = (float)y / (float)H;
    float theta = (sx - 0.5f) * 2.0f * 3.14159265358979323846f;
    float phi = (sy - 0.5f) * 3.14159265358979323846f;
    dir = std::cos(phi) * (std::cos(theta) * s.forward + std::sin(theta) * s.right) - std::sin(phi) * s.up;
    dir = normalize(dir);
  } else if (s.dof_focus != 0.0f) {
    float theta = randD(0.0f, 2.0f * 3.14159265358979323846f, st);
    float rr = randD(0.0f, s.dof_lens, st);
    float lx = rr * std::cos(theta);
    float ly = rr * std::sin(theta);
    r.eye = r.eye + lx * s.up + ly * s.right;
    V3 old_dir = s.forward + sx * s.right + sy * s.up;
    dir = (s.eye + normalize(old_dir) * s.dof_focus - r.eye) / s.dof_focus;
  } else {
    dir = s.forward + sx * s.right + sy * s.up;
  }
  r.bounce = s.bounces;
  r.dir = normalize(dir);
  return r;
}

RGBA4 shootPrimaryRay(Ctx& c, int W, int H, float x, float y, Rng& st, int32_t* id_out) {  // draw.cu:260-285
  const o_scene& s = *c.s;
  Ray ray = primary_ray(s, W, H, x, y, st);
  Hit obj = hit_nearest(c, ray);
  if (id_out) *id_out = obj.isHit ? obj.id : -1;
  RGBA4 color;
  if (obj.isHit) {
    RGBA4 diffuse = diffuseLight(c, obj, st);
    RGBA4 reflect = reflectionLight(c, ray, obj, st);
    RGBA4 refract = refractionLight(c, ray, obj, st);
    RGBA4 gic = obj.mat.color * globalIllumination(c, obj, s.gi, st);
    RGB3 one = rgb(1.f, 1.f, 1.f);
    color = obj.mat.shin * reflect + (one - obj.mat.shin) * obj.mat.trans * refract +
            (one - obj.mat.shin) * (one - obj.mat.trans) * (diffuse + gic);
    color.a = 1.0f;
  }
  return color;
}

inline uint8_t to_u8_round(float f) { return (uint8_t)(fminf(fmaxf(f, 0.0f), 1.0f) * 255.0f + 0.5f); }  // draw.cu:9-11

// one pixel with the semantics of render_kernel (draw.cu:94-133) or render_kernel_warp_aa (:135-213)
void render_pixel(Ctx& c, int W, int H, int aa, int spp_mode, int pix, uint8_t* out4, int32_t* id_out, float* lin4) {
  const int w = pix % W, h = pix / W;
  RGBA4 res;
  bool round_store;
  if (aa <= 1 && !(spp_mode == O_SPP_TRUE && aa == 1 && false)) {
    Rng st; xw_init(st, 1234ull, (unsigned long long)pix);  // draw.cu:105
    if (aa == 0) res = shootPrimaryRay(c, W, H, (float)w, (float)h, st, id_out);  // :111-114
    else {
      RGBA4 acc;
      for (int i = 0; i < aa; i++) {
        float nw = w + randD(-0.5f, 0.5f, st);
        float nh = h + randD(-0.5f, 0.5f, st);
        acc = acc + shootPrimaryRay(c, W, H, nw, nh, st, i == 0 ? id_out : nullptr);
      }
      float inv = 1.0f / aa;  // struct.cuh:55-59 mean()
      res = rgba(acc.r * inv, acc.g * inv, acc.b * inv, acc.a * inv);
    }
    round_store = false;
  } else {
    // warp kernel: 32 samples (ref) or aa samples (true), seed (1234+pixel, lane) draw.cu:162
    const int ns = (spp_mode == O_SPP_REF) ? 32 : aa;
    int np2 = 1; while (np2 < ns) np2 <<= 1;
    std::vector<RGBA4> v((size_t)np2);
    for (int lane = 0; lane < ns; lane++) {
      Rng st; xw_init(st, 1234ull + (unsigned long long)pix, (unsigned long long)lane);
      float jx = randD(-0.5f, 0.5f, st);
      float jy = randD(-0.5f, 0.5f, st);
      float sw = (float)w + jx, sh = (float)h + jy;
      v[lane] = shootPrimaryRay(c, W, H, sw, sh, st, lane == 0 ? id_out : nullptr);
    }
    // butterfly sum, draw.cu:181-189: v[i] += v[i ^ m] for m = n/2 .. 1; lane 0 holds the result
    for (int m = np2 / 2; m > 0; m /= 2) {
      std::vector<RGBA4> nv(v);
      for (int i = 0; i < np2; i++) nv[i] = v[i] + v[i ^ m];
      v.swap(nv);
    }
    const float inv = 1.0f / (float)ns;  // draw.cu:195 (1/32)
    res = rgba(v[0].r * inv, v[0].g * inv, v[0].b * inv, v[0].a * inv);
    round_store = true;
  }
  if (lin4) { lin4[0] = res.r; lin4[1] = res.g; lin4[2] = res.b; lin4[3] = res.a; }
  if (round_store) {  // draw.cu:202-211
    out4[0] = to_u8_round(RGBtosRGB(res.r)); out4[1] = to_u8_round(RGBtosRGB(res.g));
    out4[2] = to_u8_round(RGBtosRGB(res.b)); out4[3] = to_u8_round(res.a);
  } else {  // draw.cu:129-132 (implicit float -> uint8 conversion truncates)
    out4[0] = (uint8_t)(RGBtosRGB(res.r) * 255); out4[1] = (uint8_t)(RGBtosRGB(res.g) * 255);
    out4[2] = (uint8_t)(RGBtosRGB(res.b) * 255); out4[3] = (uint8_t)(res.a * 255);
  }
}

void add_counters(o_counters* dst, const o_counters& c) {
  dst->rays += c.rays; dst->nodes_visited += c.nodes_visited; dst->leaf_tests += c.leaf_tests;
  dst->shaded_hits += c.shaded_hits; if (c.max_stack > dst->max_stack) dst->max_stack = c.max_stack;
}

}  // namespace

// =============================================================================================
// C interface
// =============================================================================================
extern "C" {

void o_set_args_rtl(int v) { g_args_rtl = v; }
o_scene* o_scene_new(void) { return new o_scene(); }
void o_scene_free(o_scene* s) { delete s; }

o_scene* o_scene_parse(const char* path) {  // parse.cpp:16-39
  std::ifstream in(path);
  if (!in) return nullptr;
  o_scene* s = new o_scene();
  std::string line;
  while (std::getline(in, line)) {
    std::stringstream ss(line);
    std::vector<std::string> words; std::string w;
    while (ss >> w) words.push_back(w);
    bool ok = false;
    try { ok = parse_line(*s, words); } catch (...) { ok = false; }
    if (!ok) { delete s; return nullptr; }
  }
  return s;
}

// ints: width,height,bounces,aa,gi,fisheye,panorama,0
// floats: color3, dof_focus, dof_lens, forward3, right3, up3, eye3, target_up3, expose, ior, rough, trans3, shine3
void o_scene_get_config(const o_scene* s, int* ints, float* f) {
  ints[0] = s->width; ints[1] = s->height; ints[2] = s->bounces; ints[3] = s->aa; ints[4] = s->gi;
  ints[5] = s->fisheye; ints[6] = s->panorama; ints[7] = 0;
  int k = 0;
  f[k++] = s->color.r; f[k++] = s->color.g; f[k++] = s->color.b; f[k++] = s->dof_focus; f[k++] = s->dof_lens;
  const V3* vs[5] = {&s->forward, &s->right, &s->up, &s->eye, &s->target_up};
  for (auto v : vs) { f[k++] = v->x; f[k++] = v->y; f[k++] = v->z; }
  f[k++] = s->expose; f[k++] = s->ior; f[k++] = s->rough;
  f[k++] = s->trans.r; f[k++] = s->trans.g; f[k++] = s->trans.b;
  f[k++] = s->shine.r; f[k++] = s->shine.g; f[k++] = s->shine.b;
  while (k < 32) f[k++] = 0.f;
}
void o_scene_set_config(o_scene* s, const int* ints, const float* f) {
  s->width = ints[0]; s->height = ints[1]; s->bounces = ints[2]; s->aa = ints[3]; s->gi = ints[4];
  s->fisheye = ints[5] != 0; s->panorama = ints[6] != 0;
  int k = 0;
  s->color = rgb(f[0], f[1], f[2]); k = 3; s->dof_focus = f[k++]; s->dof_lens = f[k++];
  V3* vs[5] = {&s->forward, &s->right, &s->up, &s->eye, &s->target_up};
  for (auto v : vs) { v->x = f[k++]; v->y = f[k++]; v->z = f[k++]; }
  s->expose = f[k++]; s->ior = f[k++]; s->rough = f[k++];
  s->trans = rgb(f[k], f[k + 1], f[k + 2]); k += 3;
  s->shine = rgb(f[k], f[k + 1], f[k + 2]);
}
const char* o_scene_filename(const o_scene* s) { return s->filename.c_str(); }

int o_scene_counts(const o_scene* s, int* ns, int* nt, int* np, int* nsun, int* nb, int* npl) {
  if (ns) *ns = (int)s->spheres.size(); if (nt) *nt = (int)s->tris.size(); if (np) *np = (int)s->refs.size();
  if (nsun) *nsun = (int)s->suns.size(); if (nb) *nb = (int)s->bulbs.size(); if (npl) *npl = (int)s->planes.size();
  return (int)s->refs.size();
}
#define SOA(s) ensure_soa(const_cast<o_scene*>(s))
const float* o_scene_sphere_c(const o_scene* s) { SOA(s); return s->s_c.data(); }
const float* o_scene_sphere_r(const o_scene* s) { SOA(s); return s->s_r.data(); }
const float* o_scene_sphere_mat(const o_scene* s) { SOA(s); return s->s_m.data(); }
const float* o_scene_tri(const o_scene* s, int which) { SOA(s); return s->t_soa[which].data(); }
const float* o_scene_tri_mat(const o_scene* s) { SOA(s); return s->t_m.data(); }
const int32_t* o_scene_refs(const o_scene* s) { SOA(s); return s->refs_i.data(); }
const float* o_scene_suns(const o_scene* s) { SOA(s); return s->sun_f.data(); }
const float* o_scene_bulbs(const o_scene* s) { SOA(s); return s->bulb_f.data(); }
const float* o_scene_planes(const o_scene* s) { SOA(s); return s->plane_f.data(); }
void o_scene_bounds(const o_scene* s, int mode, float* mn, float* mx) { get_bounds(*s, mode, mn, mx); }

static Mat mat_from(const float* m) { Mat r; memcpy(static_cast<void*>(&r), m, 44); return r; }
void o_scene_add_triangles(o_scene* s, int n, const float* p0, const float* p1, const float* p2, const float* m) {
  for (int i = 0; i < n; i++)
    push_tri(*s, mk(p0[3 * i], p0[3 * i + 1], p0[3 * i + 2]), mk(p1[3 * i], p1[3 * i + 1], p1[3 * i + 2]),
             mk(p2[3 * i], p2[3 * i + 1], p2[3 * i + 2]), mat_from(m + 11 * (size_t)i));
}
void o_scene_add_spheres(o_scene* s, int n, const float* c, const float* r, const float* m) {
  for (int i = 0; i < n; i++) push_sphere(*s, mk(c[3 * i], c[3 * i + 1], c[3 * i + 2]), r[i], mat_from(m + 11 * (size_t)i));
}
void o_scene_add_sun(o_scene* s, const float* d, const float* c) {
  Sun u; u.dir = mk(d[0], d[1], d[2]); u.color = rgb(c[0], c[1], c[2]); s->suns.push_back(u); s->soa_valid = false;
}
void o_scene_add_bulb(o_scene* s, const float* p, const float* c) {
  Bulb u; u.point = mk(p[0], p[1], p[2]); u.color = rgb(c[0], c[1], c[2]); s->bulbs.push_back(u); s->soa_valid = false;
}
void o_scene_add_plane(o_scene* s, const float* abcd, const float* m) { push_plane(*s, abcd[0], abcd[1], abcd[2], abcd[3], mat_from(m)); }

void o_triangle_setup(int n, const float* p0, const float* p1, const float* p2, float* nor, float* e1, float* e2) {
  for (int i = 0; i < n; i++) {
    Tri t;
    t.p0 = mk(p0[3 * i], p0[3 * i + 1], p0[3 * i + 2]); t.p1 = mk(p1[3 * i], p1[3 * i + 1], p1[3 * i + 2]);
    t.p2 = mk(p2[3 * i], p2[3 * i + 1], p2[3 * i + 2]);
    tri_setup(t);
    nor[3 * i] = t.nor.x; nor[3 * i + 1] = t.nor.y; nor[3 * i + 2] = t.nor.z;
    e1[3 * i] = t.e1.x; e1[3 * i + 1] = t.e1.y; e1[3 * i + 2] = t.e1.z;
    e2[3 * i] = t.e2.x; e2[3 * i + 1] = t.e2.y; e2[3 * i + 2] = t.e2.z;
  }
}

void o_morton_unsorted(const o_scene* s, int bounds_mode, int bits, uint64_t* out) {
  float mn[3], mx[3]; get_bounds(*s, bounds_mode, mn, mx);
  for (size_t i = 0; i < s->refs.size(); i++) out[i] = morton_of(*s, s->refs[i], mn, mx, bits);
}

// build_lbvh_karas, src/lbvh_builder.cu:401-521 (orchestration), K1..K6
int o_build(o_scene* sc, int bounds_mode, int bits) {
  o_scene& s = *sc;
  const int n = (int)s.refs.size();
  if (n == 0) return -1;
  if (bits != 30 && bits != 63) return -2;
  std::vector<uint64_t> codes((size_t)n);
  o_morton_unsorted(sc, bounds_mode, bits, codes.data());
  // K2: thrust::sort_by_key = stable LSD radix sort (lbvh_utils.cu:110-115)
  s.perm.resize((size_t)n);
  for (int i = 0; i < n; i++) s.perm[i] = (uint32_t)i;
  std::stable_sort(s.perm.begin(), s.perm.end(), [&](uint32_t a, uint32_t b) { return codes[a] < codes[b]; });
  s.codes30.clear(); s.codes63.clear(); s.srefs.resize((size_t)n); s.srefs_i.resize((size_t)n * 2);
  for (int k = 0; k < n; k++) {
    if (bits == 30) s.codes30.push_back((uint32_t)codes[s.perm[k]]); else s.codes63.push_back(codes[s.perm[k]]);
    s.srefs[k] = s.refs[s.perm[k]];
    s.srefs_i[2 * k] = s.srefs[k].type; s.srefs_i[2 * k + 1] = (int32_t)s.srefs[k].id;
  }
  s.built_bits = bits;
  const int nn = 2 * n - 1;
  s.nbox.assign((size_t)nn, Box()); s.nleft.assign((size_t)std::max(n - 1, 0), 0xFFFFFFFFu);
  s.nright.assign((size_t)std::max(n - 1, 0), 0xFFFFFFFFu); s.nparent.assign((size_t)nn, -1);
  Keys K; K.n = (unsigned)n;
  if (bits == 30) K.k30 = s.codes30.data(); else K.k63 = s.codes63.data();
  // K4 generate_internal_nodes_karas_kernel, lbvh_builder.cu:224-322
  for (int i = 0; i < n - 1; i++) {
    int lo, hi; determine_range(K, i, lo, hi);
    int split = find_split(K, lo, hi);
    unsigned base = (unsigned)(n - 1);
    int dsplit = delta(split, split + 1, K);
    unsigned L, R;
    if (split == lo) L = base + (unsigned)split;
    else L = (delta(lo, split, K) > dsplit) ? (unsigned)split : base + (unsigned)lo;
    if (split + 1 == hi) R = base + (unsigned)hi;
    else R = (delta(split + 1, hi, K) > dsplit) ? (unsigned)(split + 1) : base + (unsigned)hi;
    s.nleft[i] = L; s.nright[i] = R;
    s.nparent[L] = i; s.nparent[R] = i;
    if (i == 0) s.nparent[0] = -1;
  }
  // K6 set_aabb_kernel_adapted, lbvh_builder.cu:324-387: leaves, then unions bottom-up.
  // fmin/fmax unions are order independent, so a counter-driven sequential sweep reproduces the result.
  for (int k = 0; k < n; k++) {
    const Ref& r = s.srefs[k];
    if (r.type == 0) { const Sphere& sp = s.spheres[r.id]; V3 rv = mk(sp.r, sp.r, sp.r); s.nbox[n - 1 + k] = box2(sp.c - rv, sp.c + rv); }
    else { const Tri& t = s.tris[r.id]; s.nbox[n - 1 + k] = box3(t.p0, t.p1, t.p2); }
  }
  std::vector<uint8_t> visited((size_t)std::max(n - 1, 0), 0);
  for (int k = 0; k < n; k++) {
    int cur = n - 1 + k;
    int p = s.nparent[cur];
    while (p != -1 && p < n - 1) {
      if (visited[p]++ == 0) break;
      s.nbox[p] = box_union(s.nbox[s.nleft[p]], s.nbox[s.nright[p]]);
      cur = p; p = s.nparent[cur];
    }
  }
  s.nbox_f.resize((size_t)nn * 6);
  for (int i = 0; i < nn; i++) memcpy(&s.nbox_f[(size_t)i * 6], &s.nbox[i], 24);
  return 0;
}
int o_num_nodes(const o_scene* s) { return (int)s->nbox.size(); }
const uint32_t* o_codes30(const o_scene* s) { return s->codes30.data(); }
const uint64_t* o_codes63(const o_scene* s) { return s->codes63.data(); }
const uint32_t* o_perm(const o_scene* s) { return s->perm.data(); }
const int32_t* o_sorted_refs(const o_scene* s) { return s->srefs_i.data(); }
const float* o_node_bbox(const o_scene* s) { return s->nbox_f.data(); }
const uint32_t* o_node_left(const o_scene* s) { return s->nleft.data(); }
const uint32_t* o_node_right(const o_scene* s) { return s->nright.data(); }
const int32_t* o_node_parent(const o_scene* s) { return s->nparent.data(); }

void o_xorwow_init(unsigned long long seed, unsigned long long subseq, uint32_t* st6) {
  Rng r; xw_init(r, seed, subseq);
  for (int i = 0; i < 5; i++) st6[i] = r.v[i];
  st6[5] = r.d;
}
void o_xorwow_draws(uint32_t* st6, int n, uint32_t* out) {
  Rng r; for (int i = 0; i < 5; i++) r.v[i] = st6[i]; r.d = st6[5]; r.bm_flag = 0; r.bm_extra = 0;
  for (int i = 0; i < n; i++) out[i] = xw_next(r);
  for (int i = 0; i < 5; i++) st6[i] = r.v[i]; st6[5] = r.d;
}

void o_trace(const o_scene* s, int n, const float* org, const float* dir, int32_t* hit_id, float* hit_t, o_counters* cnt) {
  Ctx c(s);
  for (int i = 0; i < n; i++) {
    Ray r; r.eye = mk(org[3 * i], org[3 * i + 1], org[3 * i + 2]); r.dir = mk(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2]); r.bounce = 1;
    Hit h = hit_nearest(c, r);
    hit_id[i] = h.isHit ? h.id : -1;
    if (hit_t) hit_t[i] = h.isHit ? h.distance : INFINITY;
  }
  if (cnt) *cnt = c.cnt;
}

int o_render(const o_scene* s, int W, int H, int aa, int spp_mode, int y0, int y1, int threads, uint8_t* rgba_out,
             int32_t* hit_id, uint8_t* written, float* linear, o_counters* cnt) {
  if (W < 0) W = s->width; if (H < 0) H = s->height; if (aa < 0) aa = s->aa;
  if (y0 < 0) y0 = 0; if (y1 < 0 || y1 > H) y1 = H;
  if (threads < 1) threads = 1;
  const int P = W * H;
  // launch shape of render(), draw.cu:215-239: with aa>=2 only min(P, 4*ceil(P*aa/128)) pixels have a warp (F3)
  long long covered = P;
  if (aa >= 2 && spp_mode == O_SPP_REF) {
    long long total = (long long)P * aa; long long grid = (total + 127) / 128;
    covered = std::min<long long>(P, grid * 4);
  }
  skip_tables();
  // work items of 16 consecutive pixels (not rows: a band of 4 rows would keep 4 threads busy and leave the rest idle)
  const long long p_begin = (long long)y0 * W, p_end = (long long)y1 * W;
  std::atomic<long long> next(p_begin);
  std::vector<o_counters> per((size_t)threads);
  for (auto& p : per) memset(&p, 0, sizeof(p));
  auto work = [&](int tid) {
    Ctx c(s);
    for (;;) {
      const long long b = next.fetch_add(16);
      if (b >= p_end) break;
      for (long long pp = b; pp < std::min(b + 16, p_end); pp++) {
        int pix = (int)pp;
        if (pix >= covered) { if (written) written[pix] = 0; continue; }
        render_pixel(c, W, H, aa, spp_mode, pix, rgba_out + 4 * (size_t)pix, hit_id ? hit_id + pix : nullptr,
                     linear ? linear + 4 * (size_t)pix : nullptr);
        if (written) written[pix] = 1;
      }
    }
    per[(size_t)tid] = c.cnt;
  };
  if (threads == 1) work(0);
  else {
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++) th.emplace_back(work, t);
    for (auto& t : th) t.join();
  }
  if (cnt) { memset(cnt, 0, sizeof(*cnt)); for (auto& p : per) add_counters(cnt, p); }
  return 0;
}

}  // extern "C"
