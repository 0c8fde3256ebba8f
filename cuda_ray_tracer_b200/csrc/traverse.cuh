// traverse.cuh — closest-hit / any-hit BVH traversal over the 64-byte two-child nodes.
//
// Replaces traverse_lbvh, hit_aabb_adapted and intersect_leaf_primitives (src/bvh_traversal.cu:11-183),
// checkSphereIntersectionSoA / checkTriangleIntersectionSoA (src/struct.cu:64-163), checkPlane
// (src/draw.cu:581-615) and the BVH-vs-plane choice of hitNearest (src/draw.cu:292-318).
//
// One node fetch (4 x 16 B, one 64-byte line half) brings both child boxes and both child references,
// so a ray never loads a "leaf node": the reference reads 2 x 44 B per internal step plus 44 B + 8 B
// per leaf.  Leaves hold one primitive (lbvh_builder.cu:69-70); its 48-byte trace record sits at the
// leaf's sorted position.
//
// Three lanes (template MODE):
//  * RT_LANE_REF: left child first, right pushed, popped entries not re-tested, strict `enter < tmax`
//    (bvh_traversal.cu:136-179) — the reference's exact sequence of box and primitive tests, hence its
//    exact answer.  Shadow rays stop at the first accepted primitive closer than the light: up to that
//    point the traversal is the reference's, and "the closest hit is closer than the light"
//    (draw.cu:350-353,365-368) is true iff some accepted hit is, because accepted distances only decrease.
//  * RT_LANE_EXACT (default): nearer child first with distance culling, and still the reference's answer.
//    Call a leaf a *candidate* when every box on its root path passes the order-independent part of the box
//    test (enter < exit, exit > 1e-4) and the primitive test accepts it.  The reference's left-first sweep
//    returns the last leaf it accepts; accepted distances decrease, so its answer is the lexicographic
//    minimum (t, sorted position) over candidates PROVIDED that leaf is reached, and it is reached whenever
//    all boxes on its path are entered strictly before its own hit (enter_B < t): the reference's running
//    tmax never drops below the minimum.  So: (1) find the lexicographic minimum over ALL candidates with a
//    near-first search that culls a subtree only when even its slack-inflated box (aux word of the node
//    half: how far outside its AABB an accepted hit of that subtree can lie, build.cuh) is entered beyond
//    the current hit; (2) carry the maximum exact entry distance along the path; (3) if the winner's path
//    maximum is not strictly below its distance — a hit that lies outside one of its own boxes, ~1e-4 of
//    rays on meshes — re-trace that ray in RT_LANE_REF.  Shadow rays to a sun need no distance test at all
//    (with tmax = +INF the reference visits every candidate until the first acceptance, so "occluded" ==
//    "a candidate exists"); shadow rays to a bulb accept a candidate closer than the bulb only when it is
//    consistent in the sense of (3), else they fall back as well.
//  * RT_LANE_NEAR: plain near-first with the un-inflated culling test; differs from the reference in
//    exactly the inconsistent cases above.  Kept to measure what exactness costs.
#pragma once
#include "common.cuh"
#include "rt_math.cuh"

#define RT_TRAV_THREADS 128
#define RT_STACK_SHARED 12   // entries per thread kept in shared memory
#define RT_LANE_REF 0
#define RT_LANE_NEAR 1
#define RT_LANE_EXACT 2
#define RT_STACK_TOTAL 64    // TRAVERSAL_STACK_SIZE, bvh_traversal.cu:8 (deeper pushes are dropped there too)

struct RtHit { float t; int id; };  // id >= 0: sorted position | RT_SPHERE_BIT;  -1 miss;  -2-p plane p

// hit_aabb_adapted, bvh_traversal.cu:11-44.  h0 = (xmin,ymin,zmin,xmax), h1 = (ymax,zmax,..)
template <bool INCLUSIVE>
__device__ __forceinline__ bool rt_slab(const float4 h0, const float4 h1, const V3 o, const V3 inv, const float tmax, float& enter) {
  const float tx1 = (h0.x - o.x) * inv.x, tx2 = (h0.w - o.x) * inv.x;
  const float tnx = fminf(tx1, tx2), tfx = fmaxf(tx1, tx2);
  const float ty1 = (h0.y - o.y) * inv.y, ty2 = (h1.x - o.y) * inv.y;
  const float tny = fminf(ty1, ty2), tfy = fmaxf(ty1, ty2);
  const float tz1 = (h0.z - o.z) * inv.z, tz2 = (h1.y - o.z) * inv.z;
  const float tnz = fminf(tz1, tz2), tfz = fmaxf(tz1, tz2);
  const float t_enter = fmaxf(fmaxf(tnx, tny), tnz);
  const float t_exit = fminf(fminf(tfx, tfy), tfz);
  enter = t_enter;
  return t_enter < t_exit && (INCLUSIVE ? t_enter <= tmax : t_enter < tmax) && t_exit > 0.0001f;
}

// primitive test at sorted position k; returns true and the distance when the reference's test reports a hit
__device__ __forceinline__ bool rt_prim_test(const float4* __restrict__ tv, const unsigned ref, const V3 o, const V3 d, float& th) {
  const unsigned k = ref & RT_LEAF_MASK;
  const float4 a = __ldg(tv + 3 * (size_t)k);
  if (ref & RT_SPHERE_BIT) {  // struct.cu:64-109
    const V3 c = mk3(a.x, a.y, a.z);
    const float r2 = a.w * a.w;
    const V3 cr0 = c - o;
    const bool inside = dot(cr0, cr0) < r2;
    const float tc = dot(cr0, d);
    if (!inside && tc < 0.0f) return false;
    const V3 dv = o + (tc * d) - c;
    const float d2 = dot(dv, dv);
    if (!inside && r2 < d2) return false;
    const float toff = sqrtf(r2 - d2);
    th = inside ? tc + toff : tc - toff;
    return true;
  }
  const float4 b = __ldg(tv + 3 * (size_t)k + 1);  // struct.cu:111-163
  const V3 p0 = mk3(a.x, a.y, a.z), nor = mk3(a.w, b.x, b.y);
  const float den = dot(d, nor);
  if (fabsf(den) < 1e-9f) return false;
  const float ti = dot(p0 - o, nor) / den;
  if (ti <= 0.001f) return false;
  const float4 c = __ldg(tv + 3 * (size_t)k + 2);
  const V3 e1 = mk3(b.z, b.w, c.x), e2 = mk3(c.y, c.z, c.w);
  const V3 pp = (ti * d + o) - p0;
  const float b1 = dot(e1, pp), b2 = dot(e2, pp);
  const float b0 = 1.0f - b1 - b2;
  if (!((b0 >= -0.001f) && (b1 >= -0.001f) && (b2 >= -0.001f))) return false;
  th = ti;
  return true;
}

// entry distance of the node half's box inflated by its hit slack plus a few ulps of the coordinates
// involved (rounding of t*d + o and of the slab products)
__device__ __forceinline__ float rt_enter_inflated(const float4 h0, const float4 h1, const V3 o, const V3 inv) {
  const float s = h1.w, k = 4e-7f;
  const float ax = fabsf(o.x) * k + s, ay = fabsf(o.y) * k + s, az = fabsf(o.z) * k + s;
  const float x0 = h0.x - (fabsf(h0.x) * k + ax), x1 = h0.w + (fabsf(h0.w) * k + ax);
  const float y0 = h0.y - (fabsf(h0.y) * k + ay), y1 = h1.x + (fabsf(h1.x) * k + ay);
  const float z0 = h0.z - (fabsf(h0.z) * k + az), z1 = h1.y + (fabsf(h1.y) * k + az);
  const float tnx = fminf((x0 - o.x) * inv.x, (x1 - o.x) * inv.x);
  const float tny = fminf((y0 - o.y) * inv.y, (y1 - o.y) * inv.y);
  const float tnz = fminf((z0 - o.z) * inv.z, (z1 - o.z) * inv.z);
  return fmaxf(fmaxf(tnx, tny), tnz);
}

struct RtStack {  // per-thread traversal stack: first RT_STACK_SHARED entries in shared memory, the rest in local memory
  unsigned (*ref)[RT_TRAV_THREADS];
  float (*key)[RT_TRAV_THREADS];    // entry distance used for culling at pop time
  float (*path)[RT_TRAV_THREADS];   // RT_LANE_EXACT: max exact entry distance on the path to the entry
};

// hitNearest for a ray with bounce != 0 (draw.cu:292-318).  anyhit: shadow ray — only "is there a hit closer
// than tlimit" is wanted (the reference runs a closest-hit search and looks at isHit / distance < |bulbDir|,
// draw.cu:348-367).
template <int MODE, bool COUNT>
__device__ __forceinline__ RtHit rt_trace_ray(const RtDeviceScene& sc, const V3 o, const V3 d, const float tlimit, const bool anyhit,
                                              const RtStack& S, unsigned long long& n_nodes, unsigned long long& n_leaves) {
  constexpr bool REF = MODE == RT_LANE_REF, EXACT = MODE == RT_LANE_EXACT;
  // reference order: tmax evolves from +INF exactly as in traverse_lbvh, tlimit only ends a shadow ray early
  float tmax = (REF && anyhit) ? INFINITY : tlimit;
  int best = -1;
  float best_path = -INFINITY;
  if (sc.n > 0) {
    const V3 inv = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);  // bvh_traversal.cu:106
    unsigned loc_ref[RT_STACK_TOTAL - RT_STACK_SHARED];
    float loc_key[REF ? 1 : RT_STACK_TOTAL - RT_STACK_SHARED], loc_path[EXACT ? RT_STACK_TOTAL - RT_STACK_SHARED : 1];
    int sp = 0;
    unsigned cur = (unsigned)sc.root;
    float path = -INFINITY;
    const int tid = threadIdx.x;
    const bool want_key = !(anyhit && tlimit == INFINITY);  // a sun ray never culls by distance
    while (true) {
      bool pop = false;
      if (cur & RT_LEAF_BIT) {
        if (COUNT) n_leaves++;
        float th;
        if (rt_prim_test(sc.sorted_tv, cur, o, d, th) && th > 1e-6f) {  // bvh_traversal.cu:80-86
          bool take = th < tmax;
          if (!REF && !anyhit && th == tmax && best >= 0 && (cur & RT_LEAF_MASK) < ((unsigned)best & RT_LEAF_MASK)) take = true;
          if (take) {
            if (EXACT && anyhit && tlimit != INFINITY && !(path < th)) { best = -3; break; }  // inconsistent occluder: decide in reference order
            tmax = th; best = (int)(cur & ~RT_LEAF_BIT); best_path = path;
            if (anyhit && th < tlimit) break;
          }
        }
        pop = true;
      } else {
        if (COUNT) n_nodes++;
        const float4* np = sc.nodes + 4 * (size_t)cur;
        const float4 l0 = __ldg(np), l1 = __ldg(np + 1), r0 = __ldg(np + 2), r1 = __ldg(np + 3);
        float el, er;
        bool hl = rt_slab<!REF>(l0, l1, o, inv, EXACT ? INFINITY : tmax, el);
        bool hr = rt_slab<!REF>(r0, r1, o, inv, EXACT ? INFINITY : tmax, er);
        float kl = el, kr = er;  // culling keys
        if (EXACT && want_key) {  // distance part of the test: exact entry first, the slack-inflated entry only when that fails
          const float tcut = tmax * 1.000001f;  // rounding of the slab products
          if (hl && !(el <= tmax)) { kl = rt_enter_inflated(l0, l1, o, inv); hl = kl <= tcut; }
          if (hr && !(er <= tmax)) { kr = rt_enter_inflated(r0, r1, o, inv); hr = kr <= tcut; }
        }
        const unsigned refl = __float_as_uint(l1.z), refr = __float_as_uint(r1.z);
        if (hl && hr) {
          unsigned far_ref = refr; float far_e = er, far_k = kr, near_e = el;
          cur = refl;
          if (!REF && er < el) { cur = refr; far_ref = refl; far_e = el; far_k = kl; near_e = er; }
          if (EXACT && want_key && far_k == far_e) far_k = far_ref == refl ? rt_enter_inflated(l0, l1, o, inv) : rt_enter_inflated(r0, r1, o, inv);
          const float far_path = fmaxf(path, far_e);
          if (sp < RT_STACK_SHARED) { S.ref[sp][tid] = far_ref; if (!REF) S.key[sp][tid] = far_k; if (EXACT) S.path[sp][tid] = far_path; }
          else if (sp < RT_STACK_TOTAL) { loc_ref[sp - RT_STACK_SHARED] = far_ref; if (!REF) loc_key[sp - RT_STACK_SHARED] = far_k; if (EXACT) loc_path[sp - RT_STACK_SHARED] = far_path; }
          if (sp < RT_STACK_TOTAL) sp++;  // bvh_traversal.cu:157-164: a push beyond 64 entries is dropped
          else if (EXACT) { best = -3; break; }  // the reference drops a different entry: let the reference-order lane decide
          if (EXACT) path = fmaxf(path, near_e);
        } else if (hl) { cur = refl; if (EXACT) path = fmaxf(path, el); }
        else if (hr) { cur = refr; if (EXACT) path = fmaxf(path, er); }
        else pop = true;
      }
      if (pop) {
        bool got = false;
        while (sp > 0) {
          sp--;
          const bool in_sh = sp < RT_STACK_SHARED;
          const float key = REF ? 0.f : (in_sh ? S.key[sp][tid] : loc_key[sp - RT_STACK_SHARED]);
          if (REF || !(key > (EXACT ? tmax * 1.000001f : tmax))) {
            cur = in_sh ? S.ref[sp][tid] : loc_ref[sp - RT_STACK_SHARED];
            if (EXACT) path = in_sh ? S.path[sp][tid] : loc_path[sp - RT_STACK_SHARED];
            got = true; break;
          }
        }
        if (!got) break;
      }
    }
  }
  if (EXACT) {
    // certificate: the winner must be entered strictly before its own hit on every box of its path
    if (best == -3 || (best >= 0 && !(best_path < tmax))) return rt_trace_ray<RT_LANE_REF, COUNT>(sc, o, d, tlimit, anyhit, S, n_nodes, n_leaves);
  }
  if (REF && anyhit && !(tmax < tlimit)) best = -1;  // closest hit is not closer than the light
  RtHit h;
  h.t = tmax; h.id = best;
  if (anyhit && best >= 0) return h;
  // checkPlane, draw.cu:581-615 — planes live outside the BVH
  float t_sol = INFINITY; int pidx = -1;
  for (int p = 0; p < sc.num_planes; p++) {
    const float* pl = reinterpret_cast<const float*>(sc.planes + p);
    const V3 nor = mk3(__ldg(pl), __ldg(pl + 1), __ldg(pl + 2)), pt = mk3(__ldg(pl + 3), __ldg(pl + 4), __ldg(pl + 5));
    const float t = dot(pt - o, nor) / (dot(d, nor));
    if (t <= 1e-6f) continue;
    if (t < t_sol && t > 0.001f) { t_sol = t; pidx = p; }
  }
  const bool plane_hit = !(t_sol >= 2147483648.0f);  // t_sol >= INT_MAX - 10 (int -> float), draw.cu:609
  if (plane_hit) {
    // draw.cu:304-315: the BVH hit wins only when strictly closer
    const bool bvh_wins = best >= 0 && tmax < t_sol;
    if (!bvh_wins && (!anyhit || t_sol < tlimit)) { h.t = t_sol; h.id = -2 - pidx; }
  }
  return h;
}

// ray record: q[2r] = (origin.xyz, tlimit), q[2r+1] = (dir.xyz, bits: sample | anyhit << 31)
#define RT_RAY_ANYHIT 0x80000000u

// Persistent kernel: a fixed grid (a multiple of the SM count) of warps, each fetching 32 consecutive
// rays at a time from the queue with one atomic per warp.
template <int MODE, bool COUNT>
__global__ void __launch_bounds__(RT_TRAV_THREADS) traverse_kernel(const RtDeviceScene sc, const float4* __restrict__ rays, const unsigned n,
                                                                  float2* __restrict__ hits, unsigned* fetch_counter,
                                                                  unsigned long long* stats) {
  __shared__ unsigned s_ref[RT_STACK_SHARED][RT_TRAV_THREADS];
  __shared__ float s_key[MODE == RT_LANE_REF ? 1 : RT_STACK_SHARED][RT_TRAV_THREADS];
  __shared__ float s_path[MODE == RT_LANE_EXACT ? RT_STACK_SHARED : 1][RT_TRAV_THREADS];
  RtStack S; S.ref = s_ref; S.key = s_key; S.path = s_path;
  const unsigned lane = threadIdx.x & 31;
  unsigned long long n_nodes = 0, n_leaves = 0;
  while (true) {
    unsigned base = 0;
    if (lane == 0) base = atomicAdd(fetch_counter, 32u);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= n) break;
    const unsigned r = base + lane;
    if (r < n) {
      const float4 q0 = __ldcs(rays + 2 * (size_t)r), q1 = __ldcs(rays + 2 * (size_t)r + 1);
      const RtHit h = rt_trace_ray<MODE, COUNT>(sc, mk3(q0.x, q0.y, q0.z), mk3(q1.x, q1.y, q1.z), q0.w,
                                                (__float_as_uint(q1.w) & RT_RAY_ANYHIT) != 0u, S, n_nodes, n_leaves);
      __stcs(hits + r, make_float2(h.t, __int_as_float(h.id)));
    }
    __syncwarp();
  }
  if (COUNT) {
    for (int o = 16; o > 0; o >>= 1) { n_nodes += __shfl_xor_sync(0xffffffffu, n_nodes, o); n_leaves += __shfl_xor_sync(0xffffffffu, n_leaves, o); }
    if (lane == 0) { atomicAdd(stats, n_nodes); atomicAdd(stats + 1, n_leaves); }
  }
}
