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
// Two visiting orders:
//  * REF_ORDER (default): left child first, right pushed, popped entries not re-tested, strict
//    `enter < tmax` (bvh_traversal.cu:136-179) — the reference's exact sequence of box and primitive
//    tests, hence its exact answer.  Shadow rays stop at the first accepted primitive closer than the
//    light: up to that point the traversal is the reference's, and "the closest hit is closer than the
//    light" (draw.cu:350-353,365-368) is true iff some accepted hit is, because accepted distances only
//    decrease.
//  * RT_TRAVERSE_NEAR_FIRST: nearer child first, popped entries dropped when their entry distance
//    exceeds the current hit, box test `enter <= tmax`, hit choice argmin (t, sorted position) — the
//    same answer as the reference's first-visited-wins rule whenever a hit lies inside its own boxes
//    (DESIGN.md §6).
#pragma once
#include "common.cuh"
#include "rt_math.cuh"

#define RT_TRAV_THREADS 128
#define RT_STACK_SHARED 12   // entries per thread kept in shared memory
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

// hitNearest for a ray with bounce != 0 (draw.cu:292-318).  anyhit: stop at the first primitive or plane
// closer than tlimit (shadow rays; the reference runs a closest-hit search and only looks at isHit /
// distance < |bulbDir|, draw.cu:348-367).
template <bool REF_ORDER, bool COUNT>
__device__ __forceinline__ RtHit rt_trace_ray(const RtDeviceScene& sc, const V3 o, const V3 d, const float tlimit, const bool anyhit,
                                              uint2 (*sh)[RT_TRAV_THREADS], unsigned long long& n_nodes, unsigned long long& n_leaves) {
  // reference order: tmax evolves from +INF exactly as in traverse_lbvh, tlimit only ends a shadow ray early
  float tmax = (REF_ORDER && anyhit) ? INFINITY : tlimit;
  int best = -1;
  if (sc.n > 0) {
    const V3 inv = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);  // bvh_traversal.cu:106
    uint2 loc[RT_STACK_TOTAL - RT_STACK_SHARED];
    int sp = 0;
    unsigned cur = (unsigned)sc.root;
    const int tid = threadIdx.x;
    while (true) {
      bool pop = false;
      if (cur & RT_LEAF_BIT) {
        if (COUNT) n_leaves++;
        float th;
        if (rt_prim_test(sc.sorted_tv, cur, o, d, th) && th > 1e-6f) {  // bvh_traversal.cu:80-86
          bool take = th < tmax;
          if (!REF_ORDER && !anyhit && th == tmax && best >= 0 && (cur & RT_LEAF_MASK) < ((unsigned)best & RT_LEAF_MASK)) take = true;
          if (take) {
            tmax = th; best = (int)(cur & ~RT_LEAF_BIT);
            if (anyhit && th < tlimit) break;
          }
        }
        pop = true;
      } else {
        if (COUNT) n_nodes++;
        const float4* np = sc.nodes + 4 * (size_t)cur;
        const float4 l0 = __ldg(np), l1 = __ldg(np + 1), r0 = __ldg(np + 2), r1 = __ldg(np + 3);
        float el, er;
        const bool hl = rt_slab<!REF_ORDER>(l0, l1, o, inv, tmax, el);
        const bool hr = rt_slab<!REF_ORDER>(r0, r1, o, inv, tmax, er);
        const unsigned refl = __float_as_uint(l1.z), refr = __float_as_uint(r1.z);
        if (hl && hr) {
          unsigned far_ref = refr; float far_t = er;
          cur = refl;
          if (!REF_ORDER && er < el) { cur = refr; far_ref = refl; far_t = el; }
          if (sp < RT_STACK_SHARED) sh[sp][tid] = make_uint2(far_ref, __float_as_uint(far_t));
          else if (sp < RT_STACK_TOTAL) loc[sp - RT_STACK_SHARED] = make_uint2(far_ref, __float_as_uint(far_t));
          if (sp < RT_STACK_TOTAL) sp++;  // bvh_traversal.cu:157-164: a push beyond 64 entries is dropped
        } else if (hl) cur = refl;
        else if (hr) cur = refr;
        else pop = true;
      }
      if (pop) {
        bool got = false;
        while (sp > 0) {
          sp--;
          const uint2 e = sp < RT_STACK_SHARED ? sh[sp][tid] : loc[sp - RT_STACK_SHARED];
          if (REF_ORDER || !(__uint_as_float(e.y) > tmax)) { cur = e.x; got = true; break; }
        }
        if (!got) break;
      }
    }
  }
  if (REF_ORDER && anyhit && !(tmax < tlimit)) best = -1;  // closest hit is not closer than the light
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
template <bool REF_ORDER, bool COUNT>
__global__ void __launch_bounds__(RT_TRAV_THREADS) traverse_kernel(const RtDeviceScene sc, const float4* __restrict__ rays, const unsigned n,
                                                                  float2* __restrict__ hits, unsigned* fetch_counter,
                                                                  unsigned long long* stats) {
  __shared__ uint2 s_stack[RT_STACK_SHARED][RT_TRAV_THREADS];
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
      const RtHit h = rt_trace_ray<REF_ORDER, COUNT>(sc, mk3(q0.x, q0.y, q0.z), mk3(q1.x, q1.y, q1.z), q0.w,
                                                     (__float_as_uint(q1.w) & RT_RAY_ANYHIT) != 0u, s_stack, n_nodes, n_leaves);
      __stcs(hits + r, make_float2(h.t, __int_as_float(h.id)));
    }
    __syncwarp();
  }
  if (COUNT) {
    for (int o = 16; o > 0; o >>= 1) { n_nodes += __shfl_xor_sync(0xffffffffu, n_nodes, o); n_leaves += __shfl_xor_sync(0xffffffffu, n_leaves, o); }
    if (lane == 0) { atomicAdd(stats, n_nodes); atomicAdd(stats + 1, n_leaves); }
  }
}
