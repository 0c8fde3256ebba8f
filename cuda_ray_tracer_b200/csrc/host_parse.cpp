// host_parse.cpp — TinyRaytracer scene-file parser and host-side precompute.
//
// Replaces parseInput / parseLine (src/parse.cpp:16-222) and the host constructors whose results the
// device code consumes (Triangle: include/object.cuh:177-191, Plane: object.cuh:136-141).  Same
// line-oriented, stateful format: a line is split on whitespace, dispatched on (keyword, exact word
// count) and any other line is an error (the reference prints "One of the lines are not valid." and
// exits, parse.cpp:218-221; here rt_parse_scene_file returns RT_ERR_PARSE with the line number).
// Differences in mechanism only: the file is mmap'ed and tokenised in place (no vector<string> per
// line, parse.cpp:31-35), numbers go through std::from_chars, the file is parsed in parallel chunks
// (a 10M-triangle scene is 1.3 GB / 40M lines), and primitives go straight into
// the SoA arrays the device upload needs (no AoS → SoA de-interleave, src/config_utils.cu:80-133).  Built with -ffp-contract=off like the
// reference's g++ objects (Makefile:12,38-39), so nor/e1/e2 are bit-identical.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <utility>
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
inline void put3(float* v, V3 a) { v[0] = a.x; v[1] = a.y; v[2] = a.z; }

// Output arrays are malloc'ed, not std::vector: no zero fill, and the pages are first touched by the worker that
// fills them (10M triangles = 1.2 GB of SoA output).
template <class T>
struct Arr {
  T* p = nullptr; size_t n = 0;
  Arr() = default;
  Arr(const Arr&) = delete;
  Arr& operator=(const Arr&) = delete;
  ~Arr() { free(p); }
  bool alloc(size_t k) { free(p); n = k; p = (T*)malloc((k ? k : 1) * sizeof(T)); return p != nullptr; }
  T* data() const { return p; }
  size_t size() const { return n; }
};
}  // namespace

struct rt_host_scene {
  rt_scene_desc d;
  std::string filename = "file.txt";
  Arr<float> sc, sr, sm, t[6], tm, suns, bulbs, planes;
  Arr<int32_t> refs;
  Arr<V3> verts;
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

// the "current material" part of StlConfig (parse.cpp:104-128): what a primitive line captures
struct MatState { float color[3], shine[3], trans[3], ior, rough; };
inline void put_mat(float* v, const MatState& m) {  // parse.cpp:136-143,182-189
  for (int k = 0; k < 3; k++) { v[k] = m.color[k]; v[3 + k] = m.shine[k]; v[6 + k] = m.trans[k]; }
  v[9] = m.ior; v[10] = m.rough;
}

struct Tok { const char* p; size_t n; };
inline bool is(const Tok& t, const char* s, size_t l) { return t.n == l && memcmp(t.p, s, l) == 0; }
#define KW(t, lit) is(t, lit, sizeof(lit) - 1)

// C-locale isspace, what `istringstream >> word` splits on (parse.cpp:31-35); '\n' ends the line before that
inline bool space(unsigned char c) { return c == ' ' || (c >= 9 && c <= 13); }

// std::stof / std::stoi semantics (strtof / strtol): optional sign, longest valid prefix, failure when nothing
// converts.  Fast path: std::from_chars (correctly rounded, like strtof, so the bits agree); everything it does
// not take (a leading '+', hex floats, out-of-range values) goes through strtof / strtol on a terminated copy.
bool tof_slow(const Tok& t, float& out) {
  std::string z(t.p, t.n);
  char* end; out = strtof(z.c_str(), &end); return end != z.c_str();
}
inline bool tof(const Tok& t, float& out) {
  const char* b = t.p; const char* e = t.p + t.n;
  const char* h = b;
  if (h < e && (*h == '-' || *h == '+')) h++;
  if (*b == '+' || (e - h >= 2 && h[0] == '0' && (h[1] == 'x' || h[1] == 'X'))) return tof_slow(t, out);
  const std::from_chars_result r = std::from_chars(b, e, out);
  if (r.ec == std::errc()) return true;
  if (r.ec == std::errc::result_out_of_range) return tof_slow(t, out);
  return false;
}
bool toi_slow(const Tok& t, int& out) {
  std::string z(t.p, t.n);
  char* end; long v = strtol(z.c_str(), &end, 10); out = (int)v; return end != z.c_str();
}
inline bool toi(const Tok& t, int& out) {
  if (t.n == 0 || *t.p == '+') return toi_slow(t, out);
  const std::from_chars_result r = std::from_chars(t.p, t.p + t.n, out, 10);
  if (r.ec == std::errc()) return true;
  if (r.ec == std::errc::result_out_of_range) return toi_slow(t, out);
  return false;
}

// One line -> up to 8 words.  Returns the start of the next line.
struct Line { Tok w[8]; size_t nw; bool too_many; };
inline const char* next_line(const char* p, const char* end, Line& L) {
  const char* eol = (const char*)memchr(p, '\n', (size_t)(end - p));
  if (!eol) eol = end;
  L.nw = 0; L.too_many = false;
  const char* q = p;
  while (q < eol) {
    while (q < eol && space((unsigned char)*q)) q++;
    if (q >= eol) break;
    const char* b = q;
    while (q < eol && !space((unsigned char)*q)) q++;
    if (L.nw < 8) { L.w[L.nw].p = b; L.w[L.nw].n = (size_t)(q - b); L.nw++; } else L.too_many = true;
  }
  return eol < end ? eol + 1 : end;
}

// (keyword, exact word count) dispatch of parseLine (parse.cpp:43-217)
enum Kind { K_BAD, K_GLOBAL, K_COLOR, K_SHINE, K_TRANS, K_IOR, K_ROUGH, K_XYZ, K_TRI, K_SPHERE, K_PLANE, K_SUN, K_BULB };
inline Kind classify(const Line& L) {
  if (L.too_many) return K_BAD;
  const Tok& k = L.w[0]; const size_t n = L.nw;
  switch (k.p[0]) {
    case 'x': if (KW(k, "xyz") && n == 4) return K_XYZ; break;
    case 't': if (KW(k, "tri") && n == 4) return K_TRI;
              if (KW(k, "transparency") && (n == 2 || n == 4)) return K_TRANS; break;
    case 'c': if (KW(k, "color") && n == 4) return K_COLOR; break;
    case 's': if (KW(k, "sphere") && n == 5) return K_SPHERE;
              if (KW(k, "sun") && n == 4) return K_SUN;
              if (KW(k, "shininess") && (n == 2 || n == 4)) return K_SHINE; break;
    case 'p': if (KW(k, "plane") && n == 5) return K_PLANE;
              if ((KW(k, "png") && n == 4) || (KW(k, "panorama") && n == 1)) return K_GLOBAL; break;
    case 'b': if (KW(k, "bulb") && n == 4) return K_BULB;
              if (KW(k, "bounces") && n == 2) return K_GLOBAL; break;
    case 'i': if (KW(k, "ior") && n == 2) return K_IOR; break;
    case 'r': if (KW(k, "roughness") && n == 2) return K_ROUGH; break;
    case 'f': if ((KW(k, "forward") && n == 4) || (KW(k, "fisheye") && n == 1)) return K_GLOBAL; break;
    case 'u': if (KW(k, "up") && n == 4) return K_GLOBAL; break;
    case 'e': if ((KW(k, "eye") && n == 4) || (KW(k, "expose") && n == 2)) return K_GLOBAL; break;
    case 'd': if (KW(k, "dof") && n == 3) return K_GLOBAL; break;
    case 'a': if (KW(k, "aa") && n == 2) return K_GLOBAL; break;
    case 'g': if (KW(k, "gi") && n == 2) return K_GLOBAL; break;
    default: break;
  }
  return K_BAD;
}
inline bool floats(const Line& L, int cnt, float* f) { for (int k = 0; k < cnt; k++) if (!tof(L.w[1 + k], f[k])) return false; return true; }
inline bool ints(const Line& L, int cnt, int* iv) { for (int k = 0; k < cnt; k++) if (!toi(L.w[1 + k], iv[k])) return false; return true; }

// camera / image lines: order matters only among themselves, so they are replayed sequentially
bool apply_global(const Line& L, rt_scene_desc& d, std::string& filename) {
  const Tok& k = L.w[0]; const size_t n = L.nw;
  float f[4]; int iv[3];
  if (KW(k, "png")) {
    if (!ints(L, 2, iv)) return false;
    d.width = iv[0]; d.height = iv[1]; filename.assign(L.w[3].p, L.w[3].n);
  } else if (KW(k, "bounces")) { if (!ints(L, 1, iv)) return false; d.bounces = iv[0]; }
  else if (KW(k, "forward") || KW(k, "up")) {  // parse.cpp:61-71
    if (!floats(L, 3, f)) return false;
    V3 fw = mk(d.forward[0], d.forward[1], d.forward[2]), upv = mk(d.up[0], d.up[1], d.up[2]), r;
    if (KW(k, "forward")) { fw = mk(f[0], f[1], f[2]); d.forward[0] = f[0]; d.forward[1] = f[1]; d.forward[2] = f[2]; r = normalize(cross(fw, upv)); }
    else { d.target_up[0] = f[0]; d.target_up[1] = f[1]; d.target_up[2] = f[2]; r = normalize(cross(fw, mk(f[0], f[1], f[2]))); }
    V3 u = normalize(cross(r, fw));
    d.right[0] = r.x; d.right[1] = r.y; d.right[2] = r.z; d.up[0] = u.x; d.up[1] = u.y; d.up[2] = u.z;
  } else if (KW(k, "eye")) { if (!floats(L, 3, f)) return false; d.eye[0] = f[0]; d.eye[1] = f[1]; d.eye[2] = f[2]; }
  else if (KW(k, "expose")) { if (!floats(L, 1, f)) return false; d.expose = f[0]; }
  else if (KW(k, "dof")) { if (!floats(L, 2, f)) return false; d.dof_focus = f[0]; d.dof_lens = f[1]; }
  else if (KW(k, "aa")) { if (!ints(L, 1, iv)) return false; d.aa = iv[0]; }
  else if (KW(k, "panorama")) d.panorama = 1;
  else if (KW(k, "fisheye")) d.fisheye = 1;
  else if (KW(k, "gi")) { if (!ints(L, 1, iv)) return false; d.gi = iv[0]; }
  (void)n;
  return true;
}
// material lines (parse.cpp:104-128)
inline bool apply_material(Kind kind, const Line& L, MatState& m) {
  float f[3]; const int cnt = (int)L.nw - 1;
  if (!floats(L, cnt, f)) return false;
  switch (kind) {
    case K_COLOR: m.color[0] = f[0]; m.color[1] = f[1]; m.color[2] = f[2]; break;
    case K_SHINE: m.shine[0] = f[0]; m.shine[1] = cnt == 1 ? f[0] : f[1]; m.shine[2] = cnt == 1 ? f[0] : f[2]; break;
    case K_TRANS: m.trans[0] = f[0]; m.trans[1] = cnt == 1 ? f[0] : f[1]; m.trans[2] = cnt == 1 ? f[0] : f[2]; break;
    case K_IOR: m.ior = f[0]; break;
    default: m.rough = f[0]; break;
  }
  return true;
}

// ---- chunk-parallel parse -------------------------------------------------------------------------------------
// The format is stateful (current material, running vertex list with relative indices, camera lines that depend on
// each other), so the file is cut at line boundaries and parsed in three rounds:
//   scan    (parallel)  per chunk: classify every line, count what it will emit, parse its `xyz` vertices into a
//                       local list, remember the LAST line that sets each material field and every camera line
//   stitch  (serial)    prefix sums give each chunk its output offsets and its first vertex index; the remembered
//                       material lines give each chunk its starting material; camera lines are replayed in order
//   emit    (parallel)  per chunk: copy the vertices to their global place, then (after all copies) replay the
//                       chunk with the right starting state, writing primitives straight into the SoA arrays
// The first invalid line in FILE order is the one reported, whichever round finds it.
struct Chunk {
  const char *b = nullptr, *e = nullptr;
  size_t lines = 0, n_xyz = 0, n_sph = 0, n_tri = 0, n_plane = 0, n_sun = 0, n_bulb = 0;
  std::vector<V3> verts;
  const char* last_mat[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};  // K_COLOR..K_ROUGH
  std::vector<std::pair<const char*, size_t>> globals;                      // (line start, local line number)
  size_t err_line = 0;  // local 1-based line number of the first invalid line, 0 = none
  // stitch
  size_t line0 = 0, v0 = 0, s0 = 0, t0 = 0, p0 = 0, sun0 = 0, bulb0 = 0, prim0 = 0;
  MatState start;
  float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
};

void scan_chunk(Chunk& c) {
  Line L; float f[3];
  const char* p = c.b;
  c.verts.reserve((size_t)(c.e - c.b) / 96);
  while (p < c.e) {
    const char* line = p;
    p = next_line(p, c.e, L);
    c.lines++;
    if (L.nw == 0) continue;
    const Kind k = classify(L);
    switch (k) {
      case K_XYZ:
        if (!floats(L, 3, f)) { c.err_line = c.lines; c.e = line; return; }
        c.verts.push_back(mk(f[0], f[1], f[2])); c.n_xyz++; break;
      case K_TRI: c.n_tri++; break;
      case K_SPHERE: c.n_sph++; break;
      case K_PLANE: c.n_plane++; break;
      case K_SUN: c.n_sun++; break;
      case K_BULB: c.n_bulb++; break;
      case K_COLOR: case K_SHINE: case K_TRANS: case K_IOR: case K_ROUGH: c.last_mat[k - K_COLOR] = line; break;
      case K_GLOBAL: c.globals.emplace_back(line, c.lines); break;
      default: c.err_line = c.lines; c.e = line; return;  // nothing after the first invalid line matters
    }
  }
}

inline void grow(Chunk& c, float lo, float hi, int ax) { c.mn[ax] = fminf(c.mn[ax], lo); c.mx[ax] = fmaxf(c.mx[ax], hi); }

void emit_chunk(Chunk& c, rt_host_scene& h) {
  Line L; float f[4]; int iv[3];
  MatState m = c.start;
  size_t nv = c.v0, is = c.s0, it = c.t0, ip = c.p0, isun = c.sun0, ibulb = c.bulb0, iprim = c.prim0, lineno = 0;
  const V3* verts = h.verts.data();
  const char* p = c.b;
  while (p < c.e) {
    p = next_line(p, c.e, L);
    lineno++;
    if (L.nw == 0) continue;
    const Kind k = classify(L);
    bool ok = true;
    switch (k) {
      case K_XYZ: nv++; break;
      case K_COLOR: case K_SHINE: case K_TRANS: case K_IOR: case K_ROUGH: ok = apply_material(k, L, m); break;
      case K_SPHERE: {  // parse.cpp:130-156
        if (!(ok = floats(L, 4, f))) break;
        put3(h.sc.p + 3 * is, mk(f[0], f[1], f[2])); h.sr.p[is] = f[3]; put_mat(h.sm.p + 11 * is, m);
        h.refs.p[2 * iprim] = 0; h.refs.p[2 * iprim + 1] = (int32_t)is;
        for (int ax = 0; ax < 3; ax++) {  // AABB(c - r, c + r), include/interval.cuh:55-59
          const float lo = f[ax] - f[3], hi = f[ax] + f[3];
          if (lo <= hi) grow(c, lo, hi, ax); else grow(c, hi, lo, ax);
        }
        is++; iprim++;
      } break;
      case K_TRI: {  // parse.cpp:174-201
        if (!(ok = ints(L, 3, iv))) break;
        const int size = (int)nv;
        V3 q[3];
        for (int j = 0; j < 3 && ok; j++) {
          // the reference keeps the index in a float (parse.cpp:175-180): exact below 2^24 vertices
          const float fi = (iv[j] > 0) ? iv[j] - 1 : size + iv[j];
          if (!(fi >= 0.0f) || fi >= (float)size) ok = false; else q[j] = verts[(size_t)fi];
        }
        if (!ok) break;
        V3 nor, e1, e2; tri_setup(q[0], q[1], q[2], nor, e1, e2);
        put3(h.t[0].p + 3 * it, q[0]); put3(h.t[1].p + 3 * it, q[1]); put3(h.t[2].p + 3 * it, q[2]);
        put3(h.t[3].p + 3 * it, nor); put3(h.t[4].p + 3 * it, e1); put3(h.t[5].p + 3 * it, e2);
        put_mat(h.tm.p + 11 * it, m);
        h.refs.p[2 * iprim] = 1; h.refs.p[2 * iprim + 1] = (int32_t)it;
        const float cx[3][3] = {{q[0].x, q[1].x, q[2].x}, {q[0].y, q[1].y, q[2].y}, {q[0].z, q[1].z, q[2].z}};
        for (int ax = 0; ax < 3; ax++) {  // AABB(p0,p1,p2), include/interval.cuh:61-81
          float lo = std::fmin(std::fmin(cx[ax][0], cx[ax][1]), cx[ax][2]), hi = std::fmax(std::fmax(cx[ax][0], cx[ax][1]), cx[ax][2]);
          if (hi - lo < 0.01f) { lo = lo - 0.01f; hi = hi + 0.01f; }
          grow(c, lo, hi, ax);
        }
        it++; iprim++;
      } break;
      case K_PLANE: {  // parse.cpp:157-164, object.cuh:136-141
        if (!(ok = floats(L, 4, f))) break;
        const V3 abc = mk(f[0], f[1], f[2]);
        const V3 nor = normalize(abc);
        const V3 nd = mk(abc.x * -f[3], abc.y * -f[3], abc.z * -f[3]);  // (-d * vec3(a,b,c))
        const float den = (float)(std::pow((double)f[0], 2) + std::pow((double)f[1], 2) + std::pow((double)f[2], 2));
        const V3 pt = mk(nd.x / den, nd.y / den, nd.z / den);
        float* o = h.planes.p + 21 * ip;
        for (int q4 = 0; q4 < 4; q4++) o[q4] = f[q4];
        put3(o + 4, nor); put3(o + 7, pt); put_mat(o + 10, m);
        ip++;
      } break;
      case K_SUN: case K_BULB: {
        if (!(ok = floats(L, 3, f))) break;
        float* o = k == K_SUN ? h.suns.p + 6 * isun++ : h.bulbs.p + 6 * ibulb++;
        o[0] = f[0]; o[1] = f[1]; o[2] = f[2]; o[3] = m.color[0]; o[4] = m.color[1]; o[5] = m.color[2];
      } break;
      default: break;  // K_GLOBAL: replayed by the stitch round
    }
    if (!ok) { c.err_line = lineno; return; }
  }
}

template <class F>
void parallel_for(size_t n, unsigned threads, F fn) {
  if (threads <= 1 || n <= 1) { for (size_t i = 0; i < n; i++) fn(i); return; }
  std::atomic<size_t> next{0};
  auto body = [&]() { for (size_t i; (i = next.fetch_add(1)) < n;) fn(i); };
  std::vector<std::thread> pool;
  const unsigned nt = (unsigned)std::min<size_t>(threads, n);
  for (unsigned t = 1; t < nt; t++) pool.emplace_back(body);
  body();
  for (auto& t : pool) t.join();
}

struct Mapped {  // the whole file, read-only: mmap when it is a regular file, one read otherwise
  const char* p = nullptr; size_t n = 0; bool mapped = false; std::vector<char> buf;
  ~Mapped() { if (mapped) munmap((void*)p, n); }
  bool open_file(const char* path) {
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return false;
    struct stat st;
    if (fstat(fd, &st) == 0 && S_ISREG(st.st_mode) && st.st_size > 0) {
      void* m = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
      if (m != MAP_FAILED) {
        madvise(m, (size_t)st.st_size, MADV_SEQUENTIAL);
        p = (const char*)m; n = (size_t)st.st_size; mapped = true; close(fd); return true;
      }
    }
    char tmp[1 << 16]; ssize_t got;
    while ((got = read(fd, tmp, sizeof(tmp))) > 0) buf.insert(buf.end(), tmp, tmp + got);
    close(fd);
    if (got < 0) return false;  // e.g. a directory
    p = buf.data(); n = buf.size();
    return true;
  }
};

}  // namespace

extern "C" int rt_parse_scene_file(const char* path, rt_host_scene** out) {
  if (!path || !out) { rt_set_error("rt_parse_scene_file: null argument"); return RT_ERR_INVALID; }
  *out = nullptr;
  Mapped file;
  if (!file.open_file(path)) { rt_set_error(std::string("Error opening file... ") + path); return RT_ERR_IO; }  // parse.cpp:22-25
  // cut at line boundaries: >= 1 MiB per chunk, a few chunks per thread for balance (RT_PARSE_THREADS overrides)
  unsigned threads = std::thread::hardware_concurrency();
  if (const char* e = getenv("RT_PARSE_THREADS")) { const int v = atoi(e); if (v >= 1) threads = (unsigned)v; }
  if (threads < 1) threads = 1;
  size_t chunk_bytes = (size_t)1 << 20;
  if (const char* e = getenv("RT_PARSE_CHUNK_BYTES")) { const long long v = atoll(e); if (v >= 1) chunk_bytes = (size_t)v; }  // tests
  size_t nchunks = std::max<size_t>(1, std::min<size_t>(file.n / chunk_bytes, (size_t)threads * 8));
  std::vector<Chunk> chunks(nchunks);
  {
    const char* p = file.p; const char* end = file.p + file.n;
    size_t made = 0;
    for (size_t i = 0; i < nchunks && p < end; i++) {
      const char* stop = end;
      if (i + 1 < nchunks) {
        const char* nominal = file.p + (file.n / nchunks) * (i + 1);
        if (nominal < p) nominal = p;
        const char* nl = (const char*)memchr(nominal, '\n', (size_t)(end - nominal));
        stop = nl ? nl + 1 : end;
      }
      chunks[made].b = p; chunks[made].e = stop; made++;
      p = stop;
    }
    chunks.resize(std::max<size_t>(made, 1));
    nchunks = chunks.size();
    if (made == 0) { chunks[0].b = chunks[0].e = file.p; }
  }
  auto fail_at = [&](size_t line) {
    rt_set_error(std::string("One of the lines are not valid. (") + path + ":" + std::to_string(line) + ")");  // parse.cpp:219
    return RT_ERR_PARSE;
  };

  // ---- scan
  parallel_for(nchunks, threads, [&](size_t i) { scan_chunk(chunks[i]); });
  size_t live = nchunks;  // chunks up to and including the first one with an invalid line
  for (size_t i = 0; i < nchunks; i++) if (chunks[i].err_line) { live = i + 1; break; }

  // ---- stitch
  rt_host_scene* h = new rt_host_scene();
  defaults(h->d);
  MatState m;
  for (int k = 0; k < 3; k++) { m.color[k] = h->d.color[k]; m.shine[k] = h->d.shine[k]; m.trans[k] = h->d.trans[k]; }
  m.ior = h->d.ior; m.rough = h->d.rough;
  size_t line0 = 0, v0 = 0, s0 = 0, t0 = 0, p0 = 0, sun0 = 0, bulb0 = 0;
  size_t first_err = 0;  // global 1-based line number
  auto note = [&](size_t line) { if (line && (!first_err || line < first_err)) first_err = line; };
  Line L;
  for (size_t i = 0; i < live; i++) {
    Chunk& c = chunks[i];
    c.line0 = line0; c.v0 = v0; c.s0 = s0; c.t0 = t0; c.p0 = p0; c.sun0 = sun0; c.bulb0 = bulb0; c.prim0 = s0 + t0; c.start = m;
    if (c.err_line) note(line0 + c.err_line);
    for (auto& g : c.globals) {
      next_line(g.first, file.p + file.n, L);
      if (!apply_global(L, h->d, h->filename)) { note(line0 + g.second); break; }
    }
    for (int k = 0; k < 5; k++)
      if (c.last_mat[k]) { next_line(c.last_mat[k], file.p + file.n, L); apply_material((Kind)(K_COLOR + k), L, m); }  // a bad one is found by emit
    line0 += c.lines; v0 += c.n_xyz; s0 += c.n_sph; t0 += c.n_tri; p0 += c.n_plane; sun0 += c.n_sun; bulb0 += c.n_bulb;
  }
  if (s0 + t0 > 0x7FFFFFFFull || v0 > 0x7FFFFFFFull) { delete h; rt_set_error("rt_parse_scene_file: more than 2^31-1 primitives or vertices"); return RT_ERR_INVALID; }
  bool mem = h->sc.alloc(3 * s0) && h->sr.alloc(s0) && h->sm.alloc(11 * s0) && h->tm.alloc(11 * t0) && h->refs.alloc(2 * (s0 + t0)) &&
             h->suns.alloc(6 * sun0) && h->bulbs.alloc(6 * bulb0) && h->planes.alloc(21 * p0) && h->verts.alloc(v0);
  for (int k = 0; k < 6; k++) mem = mem && h->t[k].alloc(3 * t0);
  if (!mem) { delete h; rt_set_error("rt_parse_scene_file: out of host memory"); return RT_ERR_NOMEM; }

  // ---- emit: vertices to their global place first (a `tri` may point into any earlier chunk), then the replay
  parallel_for(live, threads, [&](size_t i) {
    Chunk& c = chunks[i];
    if (!c.verts.empty()) memcpy(h->verts.p + c.v0, c.verts.data(), c.verts.size() * sizeof(V3));
    std::vector<V3>().swap(c.verts);
  });
  parallel_for(live, threads, [&](size_t i) {
    Chunk& c = chunks[i];
    const size_t scan_err = c.err_line;
    c.err_line = 0;
    emit_chunk(c, *h);
    if (!c.err_line) c.err_line = scan_err;
  });
  for (size_t i = 0; i < live; i++) note(chunks[i].err_line ? chunks[i].line0 + chunks[i].err_line : 0);
  if (first_err) { delete h; return fail_at(first_err); }

  // the primitive order inside refs is file order: spheres and triangles interleave, so prim0 = s0 + t0 at the chunk start
  for (size_t i = 0; i < live; i++)
    for (int ax = 0; ax < 3; ax++) { h->mn[ax] = fminf(h->mn[ax], chunks[i].mn[ax]); h->mx[ax] = fmaxf(h->mx[ax], chunks[i].mx[ax]); }
  for (int k = 0; k < 3; k++) { h->d.color[k] = m.color[k]; h->d.shine[k] = m.shine[k]; h->d.trans[k] = m.trans[k]; }
  h->d.ior = m.ior; h->d.rough = m.rough;
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
