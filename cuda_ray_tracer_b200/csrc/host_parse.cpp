// host_parse.cpp — TinyRaytracer scene-file parser and host-side precompute.
//
// Replaces parseInput / parseLine (src/parse.cpp:16-222) and the host constructors whose results the
// device code consumes (Triangle: include/object.cuh:177-191, Plane: object.cuh:136-141).  Same
// line-oriented, stateful format: a line is split on whitespace, dispatched on (keyword, exact word
// count) and any other line is an error (the reference prints "One of the lines are not valid." and
// exits, parse.cpp:218-221; here rt_parse_scene_file returns RT_ERR_PARSE with the line number).
// Differences in mechanism only: the file is read once and tokenised in place (no vector<string> per
// line, parse.cpp:31-35), and primitives go straight into the SoA arrays the device upload needs
// (no AoS → SoA de-interleave, src/config_utils.cu:80-133).  Built with -ffp-contract=off like the
// reference's g++ objects (Makefile:12,38-39), so nor/e1/e2 are bit-identical.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/rt_api.h"

void rt_set_error(const std::string& msg);

namespace {
struct V3 { float x, y, z; };
inline V3 mk(float a, float b, float c) { return V3{a, b, c}; }
inline V3 sub(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 add(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 cross(V3 a, V3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline bool fequal(float a, float b) {  // include/vec3.cuh:7-18
  float diff = fabsf(a - b), largest = fmaxf(fabsf(a), fabsf(b));
  if (largest < 1e-6f) return diff < 1e-6f;
  return diff / largest < 1e-6f;
}
inline V3 normalize(V3 a) {  // include/vec3.cuh:72-82
  float mag = std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z);
  if (fequal(mag, 0.0f)) return mk(0.f, 0.f, 0.f);
  float inv = 1.0f / mag;
  return mk(a.x * inv, a.y * inv, a.z * inv);
}
inline void tri_setup(V3 p0, V3 p1, V3 p2, V3& nor, V3& e1, V3& e2) {  // include/object.cuh:177-191
  nor = normalize(cross(sub(p1, p0), sub(p2, p0)));
  V3 a1 = cross(sub(p2, p0), nor), a2 = cross(sub(p1, p0), nor);
  float s1 = 1 / (dot(a1, sub(p1, p0))), s2 = 1 / (dot(a2, sub(p2, p0)));
  e1 = mk(a1.x * s1, a1.y * s1, a1.z * s1);
  e2 = mk(a2.x * s2, a2.y * s2, a2.z * s2);
}
inline void push3(std::vector<float>& v, V3 a) { v.push_back(a.x); v.push_back(a.y); v.push_back(a.z); }
}  // namespace

struct rt_host_scene {
  rt_scene_desc d;
  std::string filename = "file.txt";
  std::vector<float> sc, sr, sm, t[6], tm, suns, bulbs, planes;
  std::vector<int32_t> refs;
  std::vector<V3> verts;
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};  // parse.cpp:28
};

namespace {

void defaults(rt_scene_desc& d) {  // StlConfig defaults, include/config.hpp:29-58
  memset(&d, 0, sizeof(d));
  d.color[0] = d.color[1] = d.color[2] = 1.0f;
  d.bounces = 4;
  d.forward[2] = -1.0f; d.right[0] = 1.0f; d.up[1] = 1.0f; d.target_up[1] = 1.0f;
  d.expose = INFINITY; d.ior = 1.458f;
  for (int k = 0; k < 3; k++) { d.scene_min[k] = INFINITY; d.scene_max[k] = -INFINITY; }  // never assigned (F1)
}
void push_mat(std::vector<float>& v, const rt_scene_desc& d) {  // parse.cpp:136-143,182-189
  for (int k = 0; k < 3; k++) v.push_back(d.color[k]);
  for (int k = 0; k < 3; k++) v.push_back(d.shine[k]);
  for (int k = 0; k < 3; k++) v.push_back(d.trans[k]);
  v.push_back(d.ior); v.push_back(d.rough);
}
void grow(rt_host_scene& h, float lo, float hi, int ax) { h.mn[ax] = fminf(h.mn[ax], lo); h.mx[ax] = fmaxf(h.mx[ax], hi); }

struct Tok { const char* p; size_t n; };
inline bool is(const Tok& t, const char* s) { size_t l = strlen(s); return t.n == l && memcmp(t.p, s, l) == 0; }
// std::stof / std::stoi semantics: longest valid prefix, error when nothing converts
inline bool tof(const Tok& t, float& out) {
  char* end; out = strtof(t.p, &end); return end != t.p;
}
inline bool toi(const Tok& t, int& out) {
  char* end; long v = strtol(t.p, &end, 10); out = (int)v; return end != t.p;
}

int parse_line(rt_host_scene& h, Tok* w, size_t n) {
  rt_scene_desc& d = h.d;
  float f[4]; int iv[3];
  auto F = [&](int cnt) { for (int k = 0; k < cnt; k++) if (!tof(w[1 + k], f[k])) return false; return true; };
  auto I = [&](int cnt) { for (int k = 0; k < cnt; k++) if (!toi(w[1 + k], iv[k])) return false; return true; };
  const Tok& k = w[0];
  if (is(k, "png") && n == 4) {
    if (!I(2)) return -1;
    d.width = iv[0]; d.height = iv[1]; h.filename.assign(w[3].p, w[3].n);
  } else if (is(k, "bounces") && n == 2) { if (!I(1)) return -1; d.bounces = iv[0]; }
  else if ((is(k, "forward") || is(k, "up")) && n == 4) {  // parse.cpp:61-71
    if (!F(3)) return -1;
    V3 fw = mk(d.forward[0], d.forward[1], d.forward[2]), upv = mk(d.up[0], d.up[1], d.up[2]), r;
    if (is(k, "forward")) { fw = mk(f[0], f[1], f[2]); d.forward[0] = f[0]; d.forward[1] = f[1]; d.forward[2] = f[2]; r = normalize(cross(fw, upv)); }
    else { d.target_up[0] = f[0]; d.target_up[1] = f[1]; d.target_up[2] = f[2]; r = normalize(cross(fw, mk(f[0], f[1], f[2]))); }
    V3 u = normalize(cross(r, fw));
    d.right[0] = r.x; d.right[1] = r.y; d.right[2] = r.z; d.up[0] = u.x; d.up[1] = u.y; d.up[2] = u.z;
  } else if (is(k, "eye") && n == 4) { if (!F(3)) return -1; d.eye[0] = f[0]; d.eye[1] = f[1]; d.eye[2] = f[2]; }
  else if (is(k, "expose") && n == 2) { if (!F(1)) return -1; d.expose = f[0]; }
  else if (is(k, "dof") && n == 3) { if (!F(2)) return -1; d.dof_focus = f[0]; d.dof_lens = f[1]; }
  else if (is(k, "aa") && n == 2) { if (!I(1)) return -1; d.aa = iv[0]; }
  else if (is(k, "panorama") && n == 1) d.panorama = 1;
  else if (is(k, "fisheye") && n == 1) d.fisheye = 1;
  else if (is(k, "gi") && n == 2) { if (!I(1)) return -1; d.gi = iv[0]; }
  else if (is(k, "color") && n == 4) { if (!F(3)) return -1; d.color[0] = f[0]; d.color[1] = f[1]; d.color[2] = f[2]; }
  else if (is(k, "roughness") && n == 2) { if (!F(1)) return -1; d.rough = f[0]; }
  else if (is(k, "shininess") && (n == 2 || n == 4)) {
    if (!F((int)n - 1)) return -1;
    d.shine[0] = f[0]; d.shine[1] = n == 2 ? f[0] : f[1]; d.shine[2] = n == 2 ? f[0] : f[2];
  } else if (is(k, "transparency") && (n == 2 || n == 4)) {
    if (!F((int)n - 1)) return -1;
    d.trans[0] = f[0]; d.trans[1] = n == 2 ? f[0] : f[1]; d.trans[2] = n == 2 ? f[0] : f[2];
  } else if (is(k, "ior") && n == 2) { if (!F(1)) return -1; d.ior = f[0]; }
  else if (is(k, "sphere") && n == 5) {  // parse.cpp:130-156
    if (!F(4)) return -1;
    push3(h.sc, mk(f[0], f[1], f[2])); h.sr.push_back(f[3]); push_mat(h.sm, d);
    h.refs.push_back(0); h.refs.push_back((int32_t)(h.sr.size() - 1));
    for (int ax = 0; ax < 3; ax++) {  // AABB(c - r, c + r), include/interval.cuh:55-59
      float lo = f[ax] - f[3], hi = f[ax] + f[3];
      if (lo <= hi) grow(h, lo, hi, ax); else grow(h, hi, lo, ax);
    }
  } else if (is(k, "plane") && n == 5) {  // parse.cpp:157-164, object.cuh:136-141
    if (!F(4)) return -1;
    V3 abc = mk(f[0], f[1], f[2]);
    V3 nor = normalize(abc);
    V3 nd = mk(abc.x * -f[3], abc.y * -f[3], abc.z * -f[3]);  // (-d * vec3(a,b,c))
    float den = (float)(std::pow((double)f[0], 2) + std::pow((double)f[1], 2) + std::pow((double)f[2], 2));
    V3 pt = mk(nd.x / den, nd.y / den, nd.z / den);
    for (int q = 0; q < 4; q++) h.planes.push_back(f[q]);
    push3(h.planes, nor); push3(h.planes, pt); push_mat(h.planes, d);
  } else if (is(k, "xyz") && n == 4) { if (!F(3)) return -1; h.verts.push_back(mk(f[0], f[1], f[2])); }
  else if (is(k, "tri") && n == 4) {  // parse.cpp:174-201
    if (!I(3)) return -1;
    const int size = (int)h.verts.size();
    V3 p[3];
    for (int q = 0; q < 3; q++) {
      // the reference keeps the index in a float (parse.cpp:175-180): exact below 2^24 vertices
      float fi = (iv[q] > 0) ? iv[q] - 1 : size + iv[q];
      if (!(fi >= 0.0f) || fi >= (float)size) return -2;
      p[q] = h.verts[(size_t)fi];
    }
    V3 nor, e1, e2; tri_setup(p[0], p[1], p[2], nor, e1, e2);
    push3(h.t[0], p[0]); push3(h.t[1], p[1]); push3(h.t[2], p[2]); push3(h.t[3], nor); push3(h.t[4], e1); push3(h.t[5], e2);
    push_mat(h.tm, d);
    h.refs.push_back(1); h.refs.push_back((int32_t)(h.tm.size() / 11 - 1));
    const float cx[3][3] = {{p[0].x, p[1].x, p[2].x}, {p[0].y, p[1].y, p[2].y}, {p[0].z, p[1].z, p[2].z}};
    for (int ax = 0; ax < 3; ax++) {  // AABB(p0,p1,p2), include/interval.cuh:61-81
      float lo = std::fmin(std::fmin(cx[ax][0], cx[ax][1]), cx[ax][2]), hi = std::fmax(std::fmax(cx[ax][0], cx[ax][1]), cx[ax][2]);
      if (hi - lo < 0.01f) { lo = lo - 0.01f; hi = hi + 0.01f; }
      grow(h, lo, hi, ax);
    }
  } else if ((is(k, "sun") || is(k, "bulb")) && n == 4) {
    if (!F(3)) return -1;
    std::vector<float>& v = is(k, "sun") ? h.suns : h.bulbs;
    v.push_back(f[0]); v.push_back(f[1]); v.push_back(f[2]); v.push_back(d.color[0]); v.push_back(d.color[1]); v.push_back(d.color[2]);
  } else return -1;
  return 0;
}

}  // namespace

extern "C" int rt_parse_scene_file(const char* path, rt_host_scene** out) {
  if (!path || !out) { rt_set_error("rt_parse_scene_file: null argument"); return RT_ERR_INVALID; }
  *out = nullptr;
  FILE* fp = fopen(path, "rb");
  if (!fp) { rt_set_error(std::string("Error opening file... ") + path); return RT_ERR_IO; }  // parse.cpp:22-25
  fseek(fp, 0, SEEK_END); long sz = ftell(fp); fseek(fp, 0, SEEK_SET);
  std::vector<char> buf((size_t)sz + 1);
  size_t got = fread(buf.data(), 1, (size_t)sz, fp); fclose(fp);
  buf[got] = '\0';
  rt_host_scene* h = new rt_host_scene();
  defaults(h->d);
  char* p = buf.data(); char* end = p + got;
  long lineno = 0;
  Tok words[8];
  while (p < end) {
    char* eol = (char*)memchr(p, '\n', (size_t)(end - p));
    if (!eol) eol = end;
    lineno++;
    size_t nw = 0; bool too_many = false;
    char* q = p;
    while (q < eol) {
      while (q < eol && isspace((unsigned char)*q)) q++;
      if (q >= eol) break;
      char* b = q;
      while (q < eol && !isspace((unsigned char)*q)) q++;
      if (nw < 8) { words[nw].p = b; words[nw].n = (size_t)(q - b); nw++; } else too_many = true;
    }
    char saved = *eol; *eol = '\0';  // strtof/strtol must not run into the next line
    int rc = (nw == 0) ? 0 : (too_many ? -1 : parse_line(*h, words, nw));
    *eol = saved;
    if (rc != 0) {
      rt_set_error(std::string("One of the lines are not valid. (") + path + ":" + std::to_string(lineno) + ")");  // parse.cpp:219
      delete h; return RT_ERR_PARSE;
    }
    p = eol + 1;
  }
  *out = h;
  return RT_OK;
}

extern "C" int rt_host_scene_desc(const rt_host_scene* h, rt_scene_desc* out) {
  if (!h || !out) { rt_set_error("rt_host_scene_desc: null argument"); return RT_ERR_INVALID; }
  rt_scene_desc d = h->d;
  d.num_spheres = (int32_t)h->sr.size(); d.sphere_c = h->sc.data(); d.sphere_r = h->sr.data(); d.sphere_mat = h->sm.data();
  d.num_triangles = (int32_t)(h->tm.size() / 11);
  d.tri_p0 = h->t[0].data(); d.tri_p1 = h->t[1].data(); d.tri_p2 = h->t[2].data(); d.tri_nor = h->t[3].data();
  d.tri_e1 = h->t[4].data(); d.tri_e2 = h->t[5].data(); d.tri_mat = h->tm.data();
  d.num_prims = (int32_t)(h->refs.size() / 2); d.refs = h->refs.data();
  d.num_suns = (int32_t)(h->suns.size() / 6); d.suns = h->suns.data();
  d.num_bulbs = (int32_t)(h->bulbs.size() / 6); d.bulbs = h->bulbs.data();
  d.num_planes = (int32_t)(h->planes.size() / 21); d.planes = h->planes.data();
  *out = d;
  return RT_OK;
}
extern "C" const char* rt_host_scene_filename(const rt_host_scene* h) { return h ? h->filename.c_str() : ""; }
extern "C" void rt_host_scene_bounds(const rt_host_scene* h, float* mn, float* mx) { for (int k = 0; k < 3; k++) { mn[k] = h->mn[k]; mx[k] = h->mx[k]; } }
extern "C" void rt_host_scene_free(rt_host_scene* h) { delete h; }

extern "C" void rt_triangle_setup(int n, const float* p0, const float* p1, const float* p2, float* nor, float* e1, float* e2) {
  for (int i = 0; i < n; i++) {
    V3 a = mk(p0[3 * i], p0[3 * i + 1], p0[3 * i + 2]), b = mk(p1[3 * i], p1[3 * i + 1], p1[3 * i + 2]), c = mk(p2[3 * i], p2[3 * i + 1], p2[3 * i + 2]);
    V3 nn, x1, x2; tri_setup(a, b, c, nn, x1, x2);
    nor[3 * i] = nn.x; nor[3 * i + 1] = nn.y; nor[3 * i + 2] = nn.z;
    e1[3 * i] = x1.x; e1[3 * i + 1] = x1.y; e1[3 * i + 2] = x1.z;
    e2[3 * i] = x2.x; e2[3 * i + 1] = x2.y; e2[3 * i + 2] = x2.z;
  }
}
