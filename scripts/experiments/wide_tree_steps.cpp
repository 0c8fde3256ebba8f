// CPU experiment for DESIGN.md §8 item 1: how many dependent steps does a ray take in the reference's binary LBVH, and in
// the 4-wide tree made of every node's grandchildren (same leaves, same boxes)?  Near-first closest-hit traversal with
// distance culling, Moeller-Trumbore triangle test (step COUNTS are what is measured here, not the reference's exact
// arithmetic).  Input: a binary dump written by wide_tree_steps.py.  Output: one JSON line.
// build: g++ -O2 -std=c++17 -o wide_tree_steps wide_tree_steps.cpp
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <vector>
struct V { float x, y, z; };
static V sub(V a, V b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static V cross(V a, V b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
static float dot(V a, V b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static int n; static std::vector<uint32_t> L, R, perm; static std::vector<float> bb, tri;  // bb: xmin,xmax,ymin,ymax,zmin,zmax per node
static bool slab(uint32_t node, V o, V inv, float tmax, float& enter) {
  const float* b = &bb[6 * (size_t)node];
  float t1 = (b[0] - o.x) * inv.x, t2 = (b[1] - o.x) * inv.x, tn = std::fmin(t1, t2), tf = std::fmax(t1, t2);
  t1 = (b[2] - o.y) * inv.y; t2 = (b[3] - o.y) * inv.y; tn = std::fmax(tn, std::fmin(t1, t2)); tf = std::fmin(tf, std::fmax(t1, t2));
  t1 = (b[4] - o.z) * inv.z; t2 = (b[5] - o.z) * inv.z; tn = std::fmax(tn, std::fmin(t1, t2)); tf = std::fmin(tf, std::fmax(t1, t2));
  enter = tn;
  return tn < tf && tn <= tmax && tf > 1e-4f;
}
static bool hit_tri(uint32_t leaf, V o, V d, float& t) {
  const float* p = &tri[9 * (size_t)perm[leaf]];
  V p0{p[0], p[1], p[2]}, e1 = sub(V{p[3], p[4], p[5]}, p0), e2 = sub(V{p[6], p[7], p[8]}, p0), h = cross(d, e2);
  float a = dot(e1, h); if (std::fabs(a) < 1e-12f) return false;
  float f = 1.f / a; V s = sub(o, p0); float u = f * dot(s, h); if (u < -1e-3f || u > 1.001f) return false;
  V q = cross(s, e1); float v = f * dot(d, q); if (v < -1e-3f || u + v > 1.001f) return false;
  t = f * dot(e2, q); return t > 1e-3f;
}
struct Stat { std::vector<uint32_t> steps; double nodes = 0, boxes = 0, leaves = 0; };
static void report(const char* name, Stat& s, bool last) {
  std::sort(s.steps.begin(), s.steps.end());
  const size_t m = s.steps.size(); double sum = 0; for (uint32_t v : s.steps) sum += v;
  printf("\"%s\": {\"node_visits_per_ray\": %.2f, \"box_tests_per_ray\": %.2f, \"leaf_tests_per_ray\": %.2f, \"dependent_steps_per_ray\": {\"mean\": %.2f, \"p50\": %u, \"p99\": %u, \"p999\": %u, \"max\": %u}}%s",
         name, s.nodes / m, s.boxes / m, s.leaves / m, sum / m, s.steps[m / 2], s.steps[(size_t)(m * 0.99)], s.steps[(size_t)(m * 0.999)], s.steps[m - 1], last ? "" : ", ");
}
int main(int argc, char** argv) {
  FILE* f = fopen(argv[1], "rb"); if (!f) return 1;
  int m; if (fread(&n, 4, 1, f) != 1 || fread(&m, 4, 1, f) != 1) return 1;
  L.resize(n - 1); R.resize(n - 1); perm.resize(n); bb.resize(6 * (size_t)(2 * n - 1)); tri.resize(9 * (size_t)n);
  std::vector<float> rays(6 * (size_t)m);
  if (fread(L.data(), 4, n - 1, f) != (size_t)n - 1 || fread(R.data(), 4, n - 1, f) != (size_t)n - 1 || fread(perm.data(), 4, n, f) != (size_t)n ||
      fread(bb.data(), 4, bb.size(), f) != bb.size() || fread(tri.data(), 4, tri.size(), f) != tri.size() || fread(rays.data(), 4, rays.size(), f) != rays.size()) return 1;
  const uint32_t NI = (uint32_t)n - 1;  // node index < NI: internal, else leaf NI + k
  Stat s2, s4; s2.steps.resize(m); s4.steps.resize(m);
  std::vector<std::pair<float, uint32_t>> st; st.reserve(256);
  for (int r = 0; r < m; r++) {
    V o{rays[6 * r], rays[6 * r + 1], rays[6 * r + 2]}, d{rays[6 * r + 3], rays[6 * r + 4], rays[6 * r + 5]}, inv{1.f / d.x, 1.f / d.y, 1.f / d.z};
    {  // binary, near child first
      float tmax = INFINITY; uint32_t steps = 0; st.clear(); st.push_back({0.f, 0u});
      while (!st.empty()) {
        auto [key, cur] = st.back(); st.pop_back();
        if (key > tmax) continue;
        while (true) {
          steps++;
          if (cur >= NI) { s2.leaves++; float t; if (hit_tri(cur - NI, o, d, t) && t < tmax) tmax = t; break; }
          s2.nodes++; s2.boxes += 2;
          float el, er; const bool hl = slab(L[cur], o, inv, tmax, el), hr = slab(R[cur], o, inv, tmax, er);
          if (hl && hr) { const bool rf = er < el; st.push_back({rf ? el : er, rf ? L[cur] : R[cur]}); cur = rf ? R[cur] : L[cur]; }
          else if (hl) cur = L[cur]; else if (hr) cur = R[cur]; else break;
        }
      }
      s2.steps[r] = steps;
    }
    {  // 4-wide: the children of a step are the node's grandchildren (a leaf child stays a child)
      float tmax = INFINITY; uint32_t steps = 0; st.clear(); st.push_back({0.f, 0u});
      while (!st.empty()) {
        auto [key, cur] = st.back(); st.pop_back();
        if (key > tmax) continue;
        while (true) {
          steps++;
          if (cur >= NI) { s4.leaves++; float t; if (hit_tri(cur - NI, o, d, t) && t < tmax) tmax = t; break; }
          s4.nodes++;
          uint32_t ch[4]; int nc = 0;
          for (uint32_t c : {L[cur], R[cur]}) { if (c >= NI) ch[nc++] = c; else { ch[nc++] = L[c]; ch[nc++] = R[c]; } }
          std::pair<float, uint32_t> hit[4]; int nh = 0;
          for (int k = 0; k < nc; k++) { float e; s4.boxes++; if (slab(ch[k], o, inv, tmax, e)) hit[nh++] = {e, ch[k]}; }
          if (nh == 0) break;
          std::sort(hit, hit + nh);
          for (int k = nh - 1; k >= 1; k--) st.push_back(hit[k]);
          cur = hit[0].second;
        }
      }
      s4.steps[r] = steps;
    }
  }
  printf("{\"triangles\": %d, \"rays\": %d, ", n, m);
  report("binary_lbvh", s2, false); report("four_wide_grandchildren", s4, true);
  printf("}\n");
  return 0;
}
