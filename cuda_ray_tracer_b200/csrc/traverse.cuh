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
//    the current hit; (2) entry distances grow down a path, so the path condition is the leaf box's own
//    `enter < t`; (3) if an accepted hit fails it — a hit that lies outside its own box, ~1e-4 of rays on
//    meshes — re-trace that ray in RT_LANE_REF.  Shadow rays to a sun need no distance test at all
//    (with tmax = +INF the reference visits every candidate until the first acceptance, so "occluded" ==
//    "a candidate exists"); shadow rays to a bulb accept a candidate closer than the bulb only when it is
//    consistent in the sense of (3), else they fall back as well.
//  * RT_LANE_NEAR: plain near-first with the un-inflated culling test; differs from the reference in
//    exactly the inconsistent cases above.  Kept to measure what exactness costs.
#pragma once
#include "common.cuh"
#include "rt_math.cuh"

#ifndef RT_TRAV_THREADS
#define RT_TRAV_THREADS 128
#endif
#ifndef RT_STACK_SHARED
#define RT_STACK_SHARED 12   // entries per thread kept in shared memory
#endif
#define RT_LANE_REF 0
#define RT_LANE_NEAR 1
#define RT_LANE_EXACT 2
#define RT_STACK_TOTAL 64    // TRAVERSAL_STACK_SIZE, bvh_traversal.cu:8 (deeper pushes are dropped there too)

struct RtHit { float t; int id; };  // id >= 0: sorted position | RT_SPHERE_BIT;  -1 miss;  -2-p plane p

// hit_aabb_adapted, bvh_traversal.cu:11-44.  h0 = (xmin,ymin,xmax,ymax), h1 = (zmin,zmax,..)
// tn = per-axis near distances (the inflated key of the exact lane is derived from them)
template <bool INCLUSIVE>
__device__ __forceinline__ bool rt_slab_finish(const float tx1, const float tx2, const float ty1, const float ty2, const float tz1, const float tz2,
                                               const float tmax, float& enter, V3& tn) {
  const float tnx = fminf(tx1, tx2), tfx = fmaxf(tx1, tx2);
  const float tny = fminf(ty1, ty2), tfy = fmaxf(ty1, ty2);
  const float tnz = fminf(tz1, tz2), tfz = fmaxf(tz1, tz2);
  const float t_enter = fmaxf(fmaxf(tnx, tny), tnz);
  const float t_exit = fminf(fminf(tfx, tfy), tfz);
  enter = t_enter; tn = mk3(tnx, tny, tnz);
  return t_enter < t_exit && (INCLUSIVE ? t_enter <= tmax : t_enter < tmax) && t_exit > 0.0001f;
}
template <bool INCLUSIVE>
__device__ __forceinline__ bool rt_slab_tn(const float4 h0, const float4 h1, const V3 o, const V3 inv, const float tmax, float& enter, V3& tn) {
  return rt_slab_finish<INCLUSIVE>((h0.x - o.x) * inv.x, (h0.z - o.x) * inv.x, (h0.y - o.y) * inv.y, (h0.w - o.y) * inv.y,
                                   (h1.x - o.z) * inv.z, (h1.y - o.z) * inv.z, tmax, enter, tn);
}
template <bool INCLUSIVE>
__device__ __forceinline__ bool rt_slab(const float4 h0, const float4 h1, const V3 o, const V3 inv, const float tmax, float& enter) {
  V3 tn;
  return rt_slab_tn<INCLUSIVE>(h0, h1, o, inv, tmax, enter, tn);
}

// The same six plane distances with Blackwell's packed fp32 pairs (FADD2 / FMUL2 on sm_100a: two IEEE
// round-to-nearest operations per instruction, bit-identical to the scalar ones).  The kernel is bound by
// instruction issue, so six packed instructions instead of twelve per child box is a direct saving.  The node half
// is laid out for it: (xmin,ymin) (xmax,ymax) (zmin,zmax) are register pairs as loaded.
struct RtRay2 { unsigned long long oxy, ozz, ixy, izz; };  // (o.x,o.y) (o.z,o.z) (inv.x,inv.y) (inv.z,inv.z)
__device__ __forceinline__ unsigned long long rt_pack2(const float a, const float b) {
  unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r;
}
__device__ __forceinline__ void rt_unpack2(const unsigned long long v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ unsigned long long rt_submul2(const unsigned long long b, const unsigned long long o, const unsigned long long inv) {
  unsigned long long d, r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(b), "l"(o));
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(d), "l"(inv));
  return r;
}
__device__ __forceinline__ RtRay2 rt_ray2(const V3 o, const V3 inv) {
  RtRay2 q; q.oxy = rt_pack2(o.x, o.y); q.ozz = rt_pack2(o.z, o.z); q.ixy = rt_pack2(inv.x, inv.y); q.izz = rt_pack2(inv.z, inv.z); return q;
}
__device__ __forceinline__ V3 rt_ray2_o(const RtRay2& q) { float x, y, z, z2; rt_unpack2(q.oxy, x, y); rt_unpack2(q.ozz, z, z2); return mk3(x, y, z); }
__device__ __forceinline__ V3 rt_ray2_inv(const RtRay2& q) { float x, y, z, z2; rt_unpack2(q.ixy, x, y); rt_unpack2(q.izz, z, z2); return mk3(x, y, z); }
template <bool INCLUSIVE>
__device__ __forceinline__ bool rt_slab_pairs(const unsigned long long lo_xy, const unsigned long long hi_xy, const unsigned long long zz, const RtRay2& q,
                                              const float tmax, float& enter, V3& tn) {
  float tx1, ty1, tx2, ty2, tz1, tz2;
  rt_unpack2(rt_submul2(lo_xy, q.oxy, q.ixy), tx1, ty1);
  rt_unpack2(rt_submul2(hi_xy, q.oxy, q.ixy), tx2, ty2);
  rt_unpack2(rt_submul2(zz, q.ozz, q.izz), tz1, tz2);
  return rt_slab_finish<INCLUSIVE>(tx1, tx2, ty1, ty2, tz1, tz2, tmax, enter, tn);
}
template <bool INCLUSIVE>
__device__ __forceinline__ bool rt_slab_tn2(const float4 h0, const float4 h1, const RtRay2& q, const float tmax, float& enter, V3& tn) {
  return rt_slab_pairs<INCLUSIVE>(rt_pack2(h0.x, h0.y), rt_pack2(h0.z, h0.w), rt_pack2(h1.x, h1.y), q, tmax, enter, tn);
}
// RT_WIDE_BF16: the 4-wide node stores every child plane as a bfloat16 OFFSET from the node's min corner (96-byte nodes,
// three 32-byte fetches for four boxes instead of four).  A pair of offsets packed in one word becomes an f32x2 pair with
// a shift and a mask, and one packed add of the corner gives fp32 plane coordinates: lo = fl(org + off).  The builder
// (collapse4_kernel) rounds the offsets outward and CHECKS them with this very addition, so the decoded box contains the
// child's exact box as fp32 numbers and everything downstream is the exact test's own arithmetic on a slightly larger
// box — the monotonicity the exact lane relies on holds as it is.
#ifndef RT_WIDE_BF16
#define RT_WIDE_BF16 0
#endif
__device__ __forceinline__ unsigned long long rt_bf16pair_plus(const unsigned w, const unsigned long long org2) {
  unsigned long long p, r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(p) : "r"(w << 16), "r"(w & 0xFFFF0000u));
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(p), "l"(org2));
  return r;
}

// 256-bit read-only load (LDG.E.256 on sm_100a): one instruction per 32-byte node half / leaf record half
#ifndef RT_USE_LD256
#define RT_USE_LD256 1
#endif
__device__ __forceinline__ void rt_ld256(const float4* p, float4& a, float4& b) {
#if !RT_USE_LD256
  a = __ldg(p); b = __ldg(p + 1);
  return;
#endif
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
               : "l"(p));
}

// the same load executed only where `on` (the destination registers keep whatever they held otherwise)
__device__ __forceinline__ void rt_ld256_if(const float4* p, const bool on, float4& a, float4& b) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %9, 0;\n\t@p ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n\t}"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
               : "l"(p), "r"((unsigned)on));
}
#ifndef RT_PAIR_FETCH
#define RT_PAIR_FETCH 0
#endif


// primitive test at sorted position k; returns true and the distance when the reference's test reports a hit.
// Leaf records are 64 bytes: (p0.xyz nor.x | nor.yz e1.xy) (e1.z e2.xyz | file-order index, 0, 0, 0)
__device__ __forceinline__ bool rt_prim_test(const float4* __restrict__ tv, const unsigned ref, const V3 o, const V3 d, float& th) {
  const unsigned k = ref & RT_LEAF_MASK;
  float4 a, b;
  rt_ld256(tv + RT_TV_STRIDE * (size_t)k, a, b);
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
  // struct.cu:111-163.  The reference returns at the first failed condition; here every quantity is computed and
  // the conditions are combined at the end, so the lanes of a leaf phase stay together (a lane that fails early
  // would only idle).  den == 0 gives ti = +-INF or NaN, which the |den| condition masks.
  const float4 c = __ldg(tv + RT_TV_STRIDE * (size_t)k + 2);
  const V3 p0 = mk3(a.x, a.y, a.z), nor = mk3(a.w, b.x, b.y);
  const float den = dot(d, nor);
  const float ti = dot(p0 - o, nor) / den;
  const V3 e1 = mk3(b.z, b.w, c.x), e2 = mk3(c.y, c.z, c.w);
  const V3 pp = (ti * d + o) - p0;
  const float b1 = dot(e1, pp), b2 = dot(e2, pp);
  const float b0 = 1.0f - b1 - b2;
  th = ti;
  return !(fabsf(den) < 1e-9f) && !(ti <= 0.001f) && (b0 >= -0.001f) && (b1 >= -0.001f) && (b2 >= -0.001f);
}

// Lower bound of the entry distance of the node half's box inflated by S = aux + omk on every side, where aux
// (node half word 7, written by the build) = hit slack of the subtree + 4e-7 * largest |coordinate| of the box
// and omk = RT_ORIGIN_SLACK * largest |coordinate| of the ray origin: the slack covers accepted hits outside the box,
// the coordinate terms the rounding of t*d + o inside the primitive test (4e-7 per unit of coordinate for the triangle
// test).  The sphere test needs more from the origin term: t = tc - sqrt(r^2 - d^2) with c - o, tc and d^2 rounded at the
// magnitude M = max(|o|, |c|) puts the accepted point up to ~(6 eps M + eta^2 / r) / 2 <= 1e-6 M outside the sphere's box
// even at grazing incidence (the error of t itself is first order there, sqrt(6 eps r M), but it moves the point ALONG
// a tangent, which leaves the box only in second order), whatever r is.  Hence 2e-6 per unit of |o| here and 2e-6 per
// unit of |c| + r in the sphere's own slack (scene.cu); at |o| = 3.5 that inflates every box by 7e-6.
#define RT_ORIGIN_SLACK 2e-6f
// Per axis the inflated slab is entered at tn - S/|d| in real arithmetic.  tn (the exact test's own near
// distance, fl(fl(b - o) * inv)) carries a relative error of at most 3 ulp, inv = fl(1/d) half an ulp, so
// key = fma(-S * 1.000001, |inv|, tn) (the factor is folded into the stored aux and into omk) satisfies key <= t * (1 + 3e-7) for every accepted hit t of the subtree;
// the caller compares it with tmax * 1.000001.  Rounding is monotone, so key <= tn: the inflated test never
// culls more than the exact one.  Axis-parallel rays (inv = +-INF): tn is -INF inside the slab (key -INF),
// +INF outside (key NaN, dropped by fmaxf: no constraint from that axis, and the exact test fails anyway).
__device__ __forceinline__ float rt_key_inflated(const V3 tn, const float aux, const V3 inv, const float omk) {
  const float S = aux + omk;  // both already carry the factor 1.000001 (build_tree.cu store_half; ray fetch)
  return fmaxf(fmaxf(__fmaf_rn(-S, fabsf(inv.x), tn.x), __fmaf_rn(-S, fabsf(inv.y), tn.y)), __fmaf_rn(-S, fabsf(inv.z), tn.z));
}

struct RtStack {  // per-thread traversal stack: first RT_STACK_SHARED entries in shared memory, the rest in local memory
  unsigned (*ref)[RT_TRAV_THREADS];
  float (*key)[RT_TRAV_THREADS];    // near-first lanes: entry distance used for culling at pop time
};

// planes live outside the BVH (checkPlane, draw.cu:581-615) and the BVH hit wins only when strictly closer
// (hitNearest, draw.cu:304-315)
__device__ __forceinline__ RtHit rt_finish_with_planes(const RtDeviceScene& sc, const V3 o, const V3 d, const float tlimit, const bool anyhit,
                                                       const float tmax, const int best) {
  RtHit h;
  h.t = tmax; h.id = best;
  if (anyhit && best >= 0) return h;
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
    const bool bvh_wins = best >= 0 && tmax < t_sol;
    if (!bvh_wins && (!anyhit || t_sol < tlimit)) { h.t = t_sol; h.id = -2 - pidx; }
  }
  return h;
}

// Hit point and facing normal of hit `id` at distance t along (o, d): the ObjectInfo fields the reference's intersection
// routines return (struct.cu:76-108 sphere, :124,160 triangle; draw.cu:597-611 plane).  The same expressions as
// reconstruct_hit (render.cu), so that a shadow ray generated here starts where the shade kernel would start it.
__device__ __forceinline__ void rt_hit_point_normal(const RtDeviceScene& sc, const int id, const float t, const V3 o, const V3 d, V3& point, V3& normal) {
  point = t * d + o;
  if (id >= 0) {
    const unsigned k = (unsigned)id & RT_LEAF_MASK;
    const float4 a = __ldg(sc.sorted_tv + RT_TV_STRIDE * (size_t)k);
    if ((unsigned)id & RT_SPHERE_BIT) {
      const V3 c = mk3(a.x, a.y, a.z);
      const V3 cr0 = c - o;
      const bool inside = dot(cr0, cr0) < a.w * a.w;
      normal = normalize3(inside ? (c - point) : (point - c));
    } else {
      const float4 b = __ldg(sc.sorted_tv + RT_TV_STRIDE * (size_t)k + 1);
      const V3 nor = mk3(a.w, b.x, b.y);
      normal = (dot(d, nor) < 0.0f) ? nor : -nor;
    }
  } else {
    const float* pl = reinterpret_cast<const float*>(sc.planes + (-2 - id));
    const V3 nor = mk3(__ldg(pl), __ldg(pl + 1), __ldg(pl + 2));
    normal = (dot(nor, d) < 0.0f) ? nor : -nor;
  }
}
// origin and direction of the shadow ray towards sun 0 from that hit (diffuseLight, draw.cu:329-353: origin offset by
// the unperturbed normal * 1e-3, Ray(eye, dir, 1) normalises the direction, object.cuh:69)
__device__ __forceinline__ void rt_sun_shadow_ray(const RtDeviceScene& sc, const int id, const float t, const V3 o, const V3 d, V3& so, V3& sd) {
  V3 pt, nrm;
  rt_hit_point_normal(sc, id, t, o, d, pt, nrm);
  so = pt + nrm * 0.001f;
  const LightSun& L = sc.suns[0];
  sd = normalize3(mk3(L.dir[0], L.dir[1], L.dir[2]));
}

// ---- RT_LANE_REF: traverse_lbvh's own loop (bvh_traversal.cu:92-183) over the 64-byte nodes ----------------
// anyhit: shadow ray — only "is there a hit closer than tlimit" is wanted (the reference runs a closest-hit
// search and looks at isHit / distance < |bulbDir|, draw.cu:348-367); tmax still evolves from +INF.
template <bool COUNT>
__device__ __forceinline__ RtHit rt_trace_ref(const RtDeviceScene& sc, const V3 o, const V3 d, const float tlimit, const bool anyhit,
                                           const RtStack& S, unsigned long long& n_nodes, unsigned long long& n_leaves) {
  float tmax = INFINITY;
  int best = -1;
  if (sc.n > 0) {
    const V3 inv = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);  // bvh_traversal.cu:106
    unsigned loc_ref[RT_STACK_TOTAL - RT_STACK_SHARED];
    int sp = 0;
    unsigned cur = (unsigned)__ldg(sc.root);
    const int tid = threadIdx.x;
    while (true) {
      bool pop = false;
      if (cur & RT_LEAF_BIT) {
        if (COUNT) n_leaves++;
        float th;
        if (rt_prim_test(sc.sorted_tv, cur, o, d, th) && th > 1e-6f && th < tmax) {  // bvh_traversal.cu:80-86
          tmax = th; best = (int)(cur & ~RT_LEAF_BIT);
          if (anyhit && th < tlimit) break;
        }
        pop = true;
      } else {
        if (COUNT) n_nodes++;
        const float4* np = sc.nodes + 4 * (size_t)cur;
        float4 l0, l1, r0, r1;
        rt_ld256(np, l0, l1); rt_ld256(np + 2, r0, r1);
        float el, er;
        const bool hl = rt_slab<false>(l0, l1, o, inv, tmax, el);
        const bool hr = rt_slab<false>(r0, r1, o, inv, tmax, er);
        const unsigned refl = __float_as_uint(l1.z), refr = __float_as_uint(r1.z);
        if (hl && hr) {  // left first, right pushed (bvh_traversal.cu:149-165)
          cur = refl;
          if (sp < RT_STACK_SHARED) S.ref[sp][tid] = refr;
          else if (sp < RT_STACK_TOTAL) loc_ref[sp - RT_STACK_SHARED] = refr;
          if (sp < RT_STACK_TOTAL) sp++;  // a push beyond 64 entries is dropped (bvh_traversal.cu:157-164)
        } else if (hl) cur = refl;
        else if (hr) cur = refr;
        else pop = true;
      }
      if (pop) {
        if (sp == 0) break;
        sp--;
        cur = sp < RT_STACK_SHARED ? S.ref[sp][tid] : loc_ref[sp - RT_STACK_SHARED];
      }
    }
  }
  if (anyhit && !(tmax < tlimit)) best = -1;  // the closest hit is not closer than the light
  return rt_finish_with_planes(sc, o, d, tlimit, anyhit, tmax, best);
}

// ray record: q[2r] = (origin.xyz, tlimit), q[2r+1] = (dir.xyz, bits: sample | anyhit << 31)
#define RT_RAY_ANYHIT 0x80000000u
#define RT_RAY_DEAD 0x7FFFFFFFu   // tag of a queue slot that holds no ray (generate_kernel: the primary ray of sample i sits at slot i)

#ifndef RT_REFILL_MIN
#define RT_REFILL_MIN 8      // idle lanes that trigger a refill of the warp
#endif
#ifndef RT_LEAF_MIN
#define RT_LEAF_MIN 8        // lanes waiting with a leaf that trigger the warp's leaf phase
#endif
#ifndef RT_CHUNK
#define RT_CHUNK 64          // rays a warp reserves from the queue with one atomic (swept 32..1024 on the 1M soup)
#endif

// ---- reference-order kernel: validation lane, and the exact lane's fallback list -------------------------
// list == nullptr: the n rays of the queue from slot `first`, 32 at a time per warp.  list != nullptr: the rays whose indices the
// near-first kernel appended (their hits are written here).
template <bool COUNT>
__global__ void __launch_bounds__(RT_TRAV_THREADS) traverse_ref_kernel(const RtDeviceScene sc, const float4* __restrict__ rays, const unsigned first_in,
                                                                      const unsigned n_in, float2* __restrict__ hits, unsigned* fetch_counter,
                                                                      unsigned long long* stats, const unsigned* __restrict__ list,
                                                                      const unsigned* __restrict__ list_count,
                                                                      const unsigned* __restrict__ n_dev, const unsigned cap_back,
                                                                      unsigned* __restrict__ hits2) {
  // hits2 != nullptr (inline sun shadow rays, see traverse_near_kernel): a ray that hits something is followed by its
  // shadow ray towards sun 0, traced here in the same order; hits2[r] = 1 when that ray is occluded
  // n_dev != nullptr: the ray count is read on the device (a pass queued without host round trips, render.cu);
  // cap_back != 0: the rays are the back end of the queue, slots [cap_back - n, cap_back)
  const unsigned n = n_dev ? *n_dev : n_in, first = (n_dev && cap_back) ? cap_back - n : first_in;
  __shared__ unsigned s_ref[RT_STACK_SHARED][RT_TRAV_THREADS];
  RtStack S; S.ref = s_ref; S.key = nullptr;
  const unsigned lane = threadIdx.x & 31;
  const unsigned total = list ? *list_count : n;
  unsigned long long n_nodes = 0, n_leaves = 0;
  while (true) {
    unsigned base = 0;
    if (lane == 0) base = atomicAdd(fetch_counter, 32u);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= total) break;
    const unsigned i = base + lane;
    if (i < total) {
      const unsigned r = list ? list[i] : first + i;
      const float4 q0 = __ldcs(rays + 2 * (size_t)r), q1 = __ldcs(rays + 2 * (size_t)r + 1);
      if (__float_as_uint(q1.w) != RT_RAY_DEAD) {
        const RtHit h = rt_trace_ref<COUNT>(sc, mk3(q0.x, q0.y, q0.z), mk3(q1.x, q1.y, q1.z), q0.w,
                                            (__float_as_uint(q1.w) & RT_RAY_ANYHIT) != 0u, S, n_nodes, n_leaves);
        __stcs(hits + r, make_float2(h.t, __int_as_float(h.id)));
        if (hits2) {
          unsigned occ = 0u;
          if (h.id != -1) {
            V3 so, sd;
            rt_sun_shadow_ray(sc, h.id, h.t, mk3(q0.x, q0.y, q0.z), mk3(q1.x, q1.y, q1.z), so, sd);
            occ = rt_trace_ref<COUNT>(sc, so, sd, INFINITY, true, S, n_nodes, n_leaves).id != -1 ? 1u : 0u;
          }
          __stcs(hits2 + r, occ);
        }
      }
    }
    __syncwarp();
  }
  if (COUNT) {
    for (int o2 = 16; o2 > 0; o2 >>= 1) { n_nodes += __shfl_xor_sync(0xffffffffu, n_nodes, o2); n_leaves += __shfl_xor_sync(0xffffffffu, n_leaves, o2); }
    if (lane == 0) { atomicAdd(stats, n_nodes); atomicAdd(stats + 1, n_leaves); }
  }
}

// ---- near-first kernel (RT_LANE_EXACT / RT_LANE_NEAR) ------------------------------------------------------
// Persistent warps over a fixed grid (a multiple of the SM count).  The kernel is instruction-issue bound (ncu:
// 75 % of issue slots, ALU pipe 65 %, DRAM 1 %), so the structure is chosen for few instructions per step and
// many lanes per instruction:
//  * rays differ a lot in length (a miss leaves after a handful of nodes, a ray through the soup takes a
//    hundred), so a warp reserves RT_CHUNK consecutive rays with one atomic and hands them to its lanes as they
//    finish (refill when RT_REFILL_MIN lanes are idle);
//  * every lane is a step-by-step state machine (stack in shared memory) whose whole position is `cur`: an
//    internal node slot, or a leaf reference (bit 31).  Leaf children are not special in a node step: the nearer
//    hit child becomes `cur`, the other is pushed with its culling key;
//  * a primitive test is as long as a node step and happens once in ten steps, so lanes that arrive at a leaf
//    wait and the warp runs the tests together once RT_LEAF_MIN lanes are there (or nobody can continue);
//  * the certificate needs the entry distance of the winning leaf's own box: the 96-byte leaf record carries the
//    box in its third sector, read only when a hit is accepted (about once per ray).
// Exact lane: rays whose winning hit fails the certificate are appended to fb_list for traverse_ref_kernel.
// SUN: the queue end that holds only sun shadow rays (any hit, no distance limit).  Their answer is "does a
// candidate exist", whatever the order, so this instantiation carries no culling keys at all: the exact box test
// decides, the stack holds references only, a pop is a pop.
// EXPERIMENT, off by default (-DRT_INLINE_SUN=1): trace the sun shadow ray INLINE in the lane that has just found a primary
// hit (see the hits2 comment in traverse_near_kernel).  Bit-identical images (tests/test_render_gpu.py::
// test_inline_sun_passes_equal_two_launch_passes), one traverse launch and one drain per pass instead of two — and slower:
// 3801 vs 4107 Mrays/s on the 4K frame, 6.42 vs 6.23 ms for one rank's step of the 8-GPU frame (profiles/inline_sun_r02.txt).
// The fused launch takes 19.6 ms where the two launches take 10.6 + 6.4: the shadow rays lose the key-free SUN
// instantiation (4-byte stack entries, no culling keys) and, mixed into the same warps, dilute the coherence the
// pixel-major queue gives the primary rays.
#ifndef RT_INLINE_SUN
#define RT_INLINE_SUN 0
#endif
#ifndef RT_TRAV_MIN_CTAS
#define RT_TRAV_MIN_CTAS (RT_INLINE_SUN ? 9 : 1)   // the inline-sun experiment takes the kernel from 56 to 62 registers unless it is held to 9 CTAs per SM (55, no spills)
#endif
#ifndef RT_WIDE_MIN_CTAS
#define RT_WIDE_MIN_CTAS 9   // 4-wide instantiations (experimental): resident CTAs per SM the compiler must allow (56 registers; measured
                             // 3445 Mrays/s against 3397 at 8 CTAs / 64 registers and 3325 at 7 / 70)
#endif
enum : unsigned { F_ANYHIT = 1u, F_WANTKEY = 2u, F_FALLBACK = 4u, F_INCONS = 8u, F_SHADOW = 16u /* the inline sun shadow ray of slot r */ };  // per-ray flags of the near-first kernels
// Certificate of the exact lane (closest-hit rays).  The winner m = (t*, k*) is the reference's answer when every
// box on its path is entered before the reference's running tmax at that moment.  Entry distances grow down a
// path, so that is implied by e_m < t* (e_m = entry of m's own leaf box; "consistent").  When e_m >= t* (the hit
// lies outside its own box) it is still implied when every OTHER candidate c has t_c > e_m: the reference's tmax
// is then +INF or some t_c > e_m >= enter_B on the way to m, m is accepted (t* <= e_m < t_c) and nothing later
// beats it.  To know that, culling compares with tcull = max(t*, e_m) instead of t*, and t2 tracks the smallest
// distance among tested candidates other than the current best (a replaced best included).  A candidate that was
// never tested was culled with key > 1.000001 * max(t_b, e_b) of the best b of that moment: if b = m then
// t_c > e_m; if b != m then t_c > t_b >= t2.  So t2 > e_m proves the claim; otherwise the ray goes to fb_list.
#define RT_LEAF_TEST(ref_)                                                                                        \
do {                                                                                                            \
    if (COUNT) n_leaves++;                                                                                        \
    float th_;                                                                                                    \
    const V3 o = rt_ray2_o(q2);                                                                                   \
    float e_own_ = 0.f; bool own_ok_ = true;                                                                      \
    if (LEAF_BOX_FIRST) { /* compressed wide nodes: the leaf's OWN exact box decides whether it is a candidate */   \
      float4 b0_, b1_; rt_ld256(sc.sorted_tv + RT_TV_STRIDE * (size_t)((ref_) & RT_LEAF_MASK) + 4, b0_, b1_);      \
      V3 tn_; own_ok_ = rt_slab_tn2<true>(b0_, b1_, q2, INFINITY, e_own_, tn_);                                   \
    }                                                                                                             \
    if (own_ok_ && rt_prim_test(sc.sorted_tv, ref_, o, d, th_) && th_ > 1e-6f) { /* bvh_traversal.cu:80-86 */    \
      const bool anyhit_ = SUN || (fl & F_ANYHIT);                                                                \
      bool take_ = th_ < tmax;                                                                                    \
      if (!anyhit_ && th_ == tmax && best >= 0 && ((ref_) & RT_LEAF_MASK) < ((unsigned)best & RT_LEAF_MASK)) take_ = true; \
      if (take_) {                                                                                                \
        if (EXACT && !SUN && (fl & F_WANTKEY)) {                                                                  \
          /* entry distance of the leaf's own box, with the exact test's arithmetic (third sector of the record) */ \
          float e_ = e_own_;                                                                                      \
          if (!LEAF_BOX_FIRST) {                                                                                  \
            float4 b0_, b1_; rt_ld256(sc.sorted_tv + RT_TV_STRIDE * (size_t)((ref_) & RT_LEAF_MASK) + 4, b0_, b1_); \
            V3 tn_; rt_slab_tn2<true>(b0_, b1_, q2, INFINITY, e_, tn_);                                           \
          }                                                                                                       \
          const bool cons_ = e_ < th_;                                                                            \
          if (anyhit_) { if (!cons_) fl |= F_FALLBACK; } /* bulb shadow ray: first consistent occluder */         \
          else {                                                                                                  \
            if (best >= 0) t2 = fminf(t2, tmax);                                                                  \
            fl = cons_ ? (fl & ~F_INCONS) : (fl | F_INCONS); tcull = cons_ ? th_ : fmaxf(th_, e_);                \
          }                                                                                                       \
        } else tcull = th_;                                                                                       \
        tcut = tcull * 1.000001f;                                                                                 \
        tmax = th_; best = (int)((ref_) & ~RT_LEAF_BIT);                                                          \
      } else if (EXACT && !SUN && !anyhit_) t2 = fminf(t2, th_);                                                  \
    }                                                                                                             \
  } while (0)

template <bool EXACT, bool COUNT, bool SUN, bool WIDE>
__global__ void __launch_bounds__(RT_TRAV_THREADS, WIDE ? RT_WIDE_MIN_CTAS : RT_TRAV_MIN_CTAS) traverse_near_kernel(const RtDeviceScene sc, const float4* __restrict__ rays, const unsigned first_in,
                                                                       const unsigned n_in, float2* __restrict__ hits, unsigned* fetch_counter,
                                                                       unsigned long long* stats, unsigned* __restrict__ fb_list,
                                                                       unsigned* fb_count, const unsigned* __restrict__ n_dev, const unsigned cap_back,
                                                                       unsigned* __restrict__ hits2) {
  // hits2 != nullptr: INLINE SUN SHADOW RAYS (scenes without mirrors, glass, GI or rough surfaces, lit by one sun).  The lane
  // that has just finished a primary ray with a hit writes the hit, builds the shadow ray towards the sun from it and
  // traces it at once (an any-hit ray without a distance limit: tcull stays +INF, nothing is culled, no certificate);
  // hits2[r] = 1 when it is occluded.  A frame is then ONE traverse launch per pass instead of two: one drain of long
  // rays instead of two, and the shadow rays never travel through HBM as queue records (render.cu, shade_inline_kernel).
  // the two values are needed once per chunk of rays: kept in shared memory, not in two registers of every thread
  // (56 registers = 9 CTAs per SM; 59 = 8)
  constexpr bool LEAF_BOX_FIRST = WIDE && RT_WIDE_BF16;
  __shared__ unsigned s_range[2];
  if (threadIdx.x == 0) { const unsigned nn = n_dev ? *n_dev : n_in; s_range[0] = nn; s_range[1] = (n_dev && cap_back) ? cap_back - nn : first_in; }
  __syncthreads();
  // stack entry = (reference, culling key): one 64-bit shared-memory access per push / pop (SUN: references only)
  __shared__ unsigned s_ref[SUN ? RT_STACK_SHARED : 1][RT_TRAV_THREADS];
  __shared__ uint2 s_ent[SUN ? 1 : RT_STACK_SHARED][RT_TRAV_THREADS];
  const unsigned lane = threadIdx.x & 31;
  const int tid = threadIdx.x;
#ifdef RT_LANE_STATS
  unsigned long long ls_iter[2] = {0, 0}, ls_lanes[2] = {0, 0}, ls_parked = 0, ls_idle = 0;
#endif
  unsigned long long n_nodes = 0, n_leaves = 0;
  unsigned loc_ref[SUN ? RT_STACK_TOTAL - RT_STACK_SHARED : 1];
  uint2 loc_ent[SUN ? 1 : RT_STACK_TOTAL - RT_STACK_SHARED];
  // lane state
  enum { IDLE = 0, ACTIVE = 1, DONE = 2 };  // ACTIVE: at `cur`; DONE: result ready to be written
  int st = IDLE;
  unsigned fl = 0u, r = 0, cur = 0; int sp = 0, best = -1;
  V3 d = mk3(0, 0, 0);
  float tlimit = 0.f, tmax = 0.f, omk = 0.f;
  RtRay2 q2 = rt_ray2(d, d);  // origin and reciprocal direction live only as the packed pairs of the slab test
  // exact lane, closest-hit rays: tcull = max(tmax, entry distance of the best hit's own leaf box) is what culling
  // compares with, t2 = smallest accepted distance among the tested primitives other than the best (see below)
  float tcull = 0.f, t2 = 0.f, tcut = 0.f;  // tcut = tcull * 1.000001 (rounding of the key's products), kept in step
  // warp state (uniform)
  unsigned chunk_next = 0, chunk_end = 0; bool exhausted = false;

#define RT_PUSH(ref_, key_)                                                                                       \
  do {                                                                                                            \
    if (sp < RT_STACK_SHARED) { if (SUN) s_ref[sp][tid] = (ref_); else s_ent[sp][tid] = make_uint2((ref_), __float_as_uint(key_)); sp++; } \
    else if (sp < RT_STACK_TOTAL) { if (SUN) loc_ref[sp - RT_STACK_SHARED] = (ref_); else loc_ent[sp - RT_STACK_SHARED] = make_uint2((ref_), __float_as_uint(key_)); sp++; } \
    else { fl |= F_FALLBACK; st = DONE; } /* the reference drops an entry here (bvh_traversal.cu:157-164): let its own order decide */ \
  } while (0)
  while (true) {
    bool act = st == ACTIVE, atleaf = act && (cur & RT_LEAF_BIT);
    unsigned actm = __ballot_sync(0xffffffffu, act), leafm = __ballot_sync(0xffffffffu, atleaf);
    // ---- retire finished rays and refill from the warp's chunk, RT_REFILL_MIN lanes at a time ----
    const unsigned free_m = ~actm;
    if (free_m && (__popc(free_m) >= RT_REFILL_MIN || actm == 0u)) {
      unsigned idle_m = free_m;  // lanes that can take a new ray
      if (st == DONE) {
        st = IDLE;
        if (EXACT && ((fl & F_FALLBACK) || ((fl & F_INCONS) && !(t2 > tcull)))) fb_list[atomicAdd(fb_count, 1u)] = r;  // (inline mode: the reference-order pass redoes ray AND shadow ray)
        else {
          const RtHit h = rt_finish_with_planes(sc, rt_ray2_o(q2), d, tlimit, SUN || (fl & F_ANYHIT), tmax, best);
          if (RT_INLINE_SUN && !SUN && hits2 && (fl & F_SHADOW)) __stcs(hits2 + r, h.id != -1 ? 1u : 0u);
          else {
            __stcs(hits + r, make_float2(h.t, __int_as_float(h.id)));
            if (RT_INLINE_SUN && !SUN && hits2) {
              if (h.id != -1 && sc.n == 0) {  // planes only: nothing to traverse
                V3 so, sd;
                rt_sun_shadow_ray(sc, h.id, h.t, rt_ray2_o(q2), d, so, sd);
                __stcs(hits2 + r, rt_finish_with_planes(sc, so, sd, INFINITY, true, INFINITY, -1).id != -1 ? 1u : 0u);
              } else if (h.id != -1) {  // this lane goes on with the shadow ray of its hit
                V3 so, sd;
                rt_sun_shadow_ray(sc, h.id, h.t, rt_ray2_o(q2), d, so, sd);
                d = sd; tlimit = INFINITY;
                const V3 inv = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
                omk = (RT_ORIGIN_SLACK * 1.000001f) * fmaxf(fmaxf(fabsf(so.x), fabsf(so.y)), fabsf(so.z));
                q2 = rt_ray2(so, inv);
                tmax = INFINITY; tcull = INFINITY; tcut = INFINITY; t2 = INFINITY; best = -1; sp = 0; cur = (unsigned)__ldg(sc.root);
                fl = F_ANYHIT | F_SHADOW;
                st = ACTIVE;
              } else __stcs(hits2 + r, 0u);
            }
          }
        }
      }
      if (RT_INLINE_SUN && !SUN && hits2) idle_m = __ballot_sync(0xffffffffu, st == IDLE);  // converted lanes are busy again
      if (!exhausted) {
        if (chunk_next == chunk_end) {
          // (Guided chunk sizes — ever smaller pieces towards the end of the queue — were measured: no change of the
          // per-launch fixed cost, which is the latency of the longest single rays, scripts/launch_fixed_cost.py.)
          unsigned base = 0;
          if (lane == 0) base = atomicAdd(fetch_counter, (unsigned)RT_CHUNK);
          base = __shfl_sync(0xffffffffu, base, 0);
          const unsigned n = s_range[0], first = s_range[1];
          if (base >= n) exhausted = true;
          else { chunk_next = first + base; chunk_end = first + min(base + (unsigned)RT_CHUNK, n); }
        }
        if (!exhausted) {
          const unsigned avail = chunk_end - chunk_next, want = (unsigned)__popc(idle_m);
          const unsigned take = min(avail, want), rank = (unsigned)__popc(idle_m & ((1u << lane) - 1u));
          bool fetch = st == IDLE && rank < take;
          float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f), q1 = q0;
          // both halves of the record in one go (ncu: testing the tag before loading q0 put two memory latencies in a row
          // into every refill, 2 % of the kernel's stall samples); a DEAD slot — no ray — is rare
          if (fetch) { r = chunk_next + rank; q0 = __ldcs(rays + 2 * (size_t)r); q1 = __ldcs(rays + 2 * (size_t)r + 1); fetch = __float_as_uint(q1.w) != RT_RAY_DEAD; }
          if (fetch) {
            const V3 o = mk3(q0.x, q0.y, q0.z); d = mk3(q1.x, q1.y, q1.z); tlimit = q0.w;
            const bool anyhit = SUN || (__float_as_uint(q1.w) & RT_RAY_ANYHIT) != 0u;
            const V3 inv = mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);  // bvh_traversal.cu:106
            omk = (RT_ORIGIN_SLACK * 1.000001f) * fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z));
            q2 = rt_ray2(o, inv);
            tmax = tlimit; tcull = tlimit; tcut = tlimit * 1.000001f; t2 = INFINITY; best = -1; sp = 0; cur = sc.n > 0 ? (unsigned)__ldg(sc.root) : 0u;
            // a sun ray never culls by distance (tcull stays +INF) and needs no certificate
            fl = (anyhit ? F_ANYHIT : 0u) | ((!SUN && !(anyhit && tlimit == INFINITY)) ? F_WANTKEY : 0u);
            st = sc.n > 0 ? ACTIVE : DONE;  // n == 0: planes only.  n == 1: the root is the leaf
          }
          chunk_next += take;
        }
      }
      act = st == ACTIVE; atleaf = act && (cur & RT_LEAF_BIT);
      actm = __ballot_sync(0xffffffffu, act); leafm = __ballot_sync(0xffffffffu, atleaf);
      if (actm == 0u) {  // everybody was retired and nothing came in
        if (exhausted) break;
        continue;
      }
    }
    bool need_pop = false;
#ifdef RT_LANE_STATS  // experiment build only (make EXTRA=-DRT_LANE_STATS, scripts/lane_stats.py): lane occupancy of the two phases
    if (COUNT && lane == 0) {
      const bool lp = __popc(leafm) >= RT_LEAF_MIN || leafm == actm;
      ls_iter[lp ? 1 : 0]++; ls_lanes[lp ? 1 : 0] += lp ? __popc(leafm) : __popc(actm & ~leafm);
      if (!lp) ls_parked += __popc(leafm);
      ls_idle += 32 - __popc(actm);
    }
#endif
    // (Serving leaf lanes AND node lanes in every iteration once the queue is empty — latency over throughput in the
    // drain — was measured: no change of the per-launch fixed cost, scripts/launch_fixed_cost.py.)
    const bool leaf_phase = __popc(leafm) >= RT_LEAF_MIN || leafm == actm;
    if (leaf_phase) {
      // ---- leaf phase ----
      if (atleaf) {
        RT_LEAF_TEST(cur);
        if (((SUN || (fl & F_ANYHIT)) && best >= 0) || (fl & F_FALLBACK)) st = DONE; else need_pop = true;
      }
    }
#if RT_PAIR_FETCH
    else {
      // ---- node phase, node fetched by LANE PAIRS ----
      // EXPERIMENT, off by default.  Measured on the 4K soup frame: 2579-2813 Mrays/s against 3410 (-18 % .. -25 %) at
      // 72 / 64 / 56 registers: halving the L1TEX wavefronts does not pay for 9 shuffles + 16 selects per node step
      // and 16 more live registers in a kernel that issues at 70 %.
      // L1TEX charges one wavefront per distinct 128-byte line per load instruction, so 32 lanes reading 32 different
      // nodes cost 32 wavefronts per 32-byte half.  Here lanes 2k and 2k+1 read the two halves of ONE node in the same
      // instruction (first the even lane's node, then the odd lane's): 16 lines per instruction, half the wavefronts
      // per node (scripts/microbench/l2bw.cu: 14.7 instead of 8.2 TB/s of 64-byte records).  Each lane keeps the half
      // of its own node that it loaded itself (child A: the left child on even lanes, the right child on odd lanes)
      // and receives the other half (child B) from its partner through eight shuffles.
      const bool step = act && !atleaf;
      const unsigned odd = lane & 1u;
      const unsigned mine = step ? cur : 0xFFFFFFFFu;
      const unsigned theirs = __shfl_xor_sync(0xffffffffu, mine, 1);
      const unsigned ce = odd ? theirs : mine, co = odd ? mine : theirs;
      float4 x0, x1, y0, y1;
      rt_ld256_if(sc.nodes + 4 * (size_t)ce + 2 * odd, ce != 0xFFFFFFFFu, x0, x1);
      rt_ld256_if(sc.nodes + 4 * (size_t)co + 2 * odd, co != 0xFFFFFFFFu, y0, y1);
      float4 a0 = odd ? y0 : x0, a1 = odd ? y1 : x1, b0 = odd ? x0 : y0, b1 = odd ? x1 : y1;
      b0.x = __shfl_xor_sync(0xffffffffu, b0.x, 1); b0.y = __shfl_xor_sync(0xffffffffu, b0.y, 1);
      b0.z = __shfl_xor_sync(0xffffffffu, b0.z, 1); b0.w = __shfl_xor_sync(0xffffffffu, b0.w, 1);
      b1.x = __shfl_xor_sync(0xffffffffu, b1.x, 1); b1.y = __shfl_xor_sync(0xffffffffu, b1.y, 1);
      b1.z = __shfl_xor_sync(0xffffffffu, b1.z, 1);
      if (EXACT && !SUN) b1.w = __shfl_xor_sync(0xffffffffu, b1.w, 1);
      if (step) {
        if (COUNT) n_nodes++;
        float ea, eb;
        V3 tna, tnb;
        bool ha = rt_slab_tn2<true>(a0, a1, q2, (EXACT || SUN) ? INFINITY : tmax, ea, tna);
        bool hb = rt_slab_tn2<true>(b0, b1, q2, (EXACT || SUN) ? INFINITY : tmax, eb, tnb);
        const unsigned refa = __float_as_uint(a1.z), refb = __float_as_uint(b1.z);
        float ka = ea, kb = eb;
        if (EXACT && !SUN) {
          const V3 inv = rt_ray2_inv(q2);
          ka = rt_key_inflated(tna, a1.w, inv, omk); kb = rt_key_inflated(tnb, b1.w, inv, omk);
          ha = ha && !(ka > tcut); hb = hb && !(kb > tcut);
        }
        // the order of the scalar kernel: right child first iff it is hit and (the left is not, or right enters strictly
        // earlier).  Even lanes: A = left, B = right.  Odd lanes: A = right, B = left.
        const bool b_first = hb && (!ha || (odd ? !(ea < eb) : (eb < ea)));
        if (ha && hb) RT_PUSH(b_first ? refa : refb, b_first ? ka : kb);
        cur = b_first ? refb : refa;
        need_pop = !(ha || hb);
      }
    }
#else
    if (WIDE && !leaf_phase && act && !atleaf) {
      // ---- node phase, 4-wide tree (RT_BUILD_WIDE4): one step = the four grandchildren of a binary node ----
      // Half the dependent steps per ray for the same box tests (profiles/wide_tree_steps_r01.json).  Every child is
      // tested exactly like a child of the binary step (order-independent box test, slack-inflated culling key), so
      // the set of leaves reached differs from the binary lane's only by subtrees the intermediate node would have
      // culled — and the exact lane's answer does not depend on the order or the tree (its definition is on leaves).
      // Places 0,1 hold the children of the binary node's left child, places 2,3 those of its right child (a leaf child
      // sits alone in its pair; the empty place holds a NaN box, which fails the test).
      // ORDER: what two binary steps would do — the pair whose nearer hit child is nearer first, inside a pair the
      // nearer child first — from three comparisons, no sorting network: child i gets a 2-bit visiting rank, the hit
      // children are written to their stack slots directly (slot = sp + number of hit children visited after it), the
      // first visited one becomes the next position.  No (distance, key, reference) triples move between registers.
      if (COUNT) n_nodes++;
      float e[4], k[4]; unsigned rf[4]; bool h[4];
#if RT_WIDE_BF16
      // 96-byte node: (org.xyz, -) (ref0..3) | 12 words of bf16 offset pairs, child c = words 3c..3c+2 = (xlo,ylo) (xhi,yhi) (zlo,zhi) | aux0..3
      const float4* np = sc.wnodes + 6 * (size_t)cur;
      float4 n0, n1, w0, w1, w2, ax;
      rt_ld256(np, n0, n1); rt_ld256(np + 2, w0, w1); rt_ld256(np + 4, w2, ax);
      const unsigned long long oxy = rt_pack2(n0.x, n0.y), ozz = rt_pack2(n0.z, n0.z);
      const unsigned wd[12] = {__float_as_uint(w0.x), __float_as_uint(w0.y), __float_as_uint(w0.z), __float_as_uint(w0.w), __float_as_uint(w1.x), __float_as_uint(w1.y),
                               __float_as_uint(w1.z), __float_as_uint(w1.w), __float_as_uint(w2.x), __float_as_uint(w2.y), __float_as_uint(w2.z), __float_as_uint(w2.w)};
      const float axv[4] = {ax.x, ax.y, ax.z, ax.w};
      rf[0] = __float_as_uint(n1.x); rf[1] = __float_as_uint(n1.y); rf[2] = __float_as_uint(n1.z); rf[3] = __float_as_uint(n1.w);
#pragma unroll
      for (int c = 0; c < 4; c++) {
        V3 tn;
        h[c] = rt_slab_pairs<true>(rt_bf16pair_plus(wd[3 * c], oxy), rt_bf16pair_plus(wd[3 * c + 1], oxy), rt_bf16pair_plus(wd[3 * c + 2], ozz), q2,
                                   (EXACT || SUN) ? INFINITY : tmax, e[c], tn);  // NaN offsets (unused place) fail
        k[c] = e[c];
        if (EXACT && !SUN) { k[c] = rt_key_inflated(tn, axv[c], rt_ray2_inv(q2), omk); h[c] = h[c] && !(k[c] > tcut); }
        if (!h[c]) e[c] = INFINITY;
      }
#else
      const float4* np = sc.wnodes + 8 * (size_t)cur;
#pragma unroll
      for (int c = 0; c < 4; c++) {
        float4 a, b;
        rt_ld256(np + 2 * c, a, b);
        V3 tn;
        h[c] = rt_slab_tn2<true>(a, b, q2, (EXACT || SUN) ? INFINITY : tmax, e[c], tn);  // a NaN box (unused place) fails
        rf[c] = __float_as_uint(b.z);
        k[c] = e[c];
        if (EXACT && !SUN) { k[c] = rt_key_inflated(tn, b.w, rt_ray2_inv(q2), omk); h[c] = h[c] && !(k[c] > tcut); }
        if (!h[c]) e[c] = INFINITY;  // a box that passes is never entered at +INF
      }
#endif
      const unsigned fl_ = e[1] < e[0] ? 1u : 0u, fr_ = e[3] < e[2] ? 1u : 0u;
      const unsigned ft_ = fminf(e[2], e[3]) < fminf(e[0], e[1]) ? 1u : 0u;
      const unsigned rk0 = (ft_ << 1) | fl_, rk1 = (ft_ << 1) | (fl_ ^ 1u), rk2 = ((ft_ ^ 1u) << 1) | fr_, rk3 = ((ft_ ^ 1u) << 1) | (fr_ ^ 1u);
      const unsigned hr = ((unsigned)h[0] << rk0) | ((unsigned)h[1] << rk1) | ((unsigned)h[2] << rk2) | ((unsigned)h[3] << rk3);  // hits in visiting order
      const int nhit = __popc(hr);
      need_pop = nhit == 0;
      if (nhit > 0) {
        const int b0_ = __popc(hr >> (rk0 + 1)), b1_ = __popc(hr >> (rk1 + 1)), b2_ = __popc(hr >> (rk2 + 1)), b3_ = __popc(hr >> (rk3 + 1));  // hit children visited after child i
        const int first = nhit - 1;
        if (h[0] && b0_ == first) cur = rf[0];
        if (h[1] && b1_ == first) cur = rf[1];
        if (h[2] && b2_ == first) cur = rf[2];
        if (h[3] && b3_ == first) cur = rf[3];
        if (sp + first <= RT_STACK_SHARED) {  // all in shared memory: one predicated store per child
          if (SUN) {
            if (h[0] && b0_ != first) s_ref[sp + b0_][tid] = rf[0];
            if (h[1] && b1_ != first) s_ref[sp + b1_][tid] = rf[1];
            if (h[2] && b2_ != first) s_ref[sp + b2_][tid] = rf[2];
            if (h[3] && b3_ != first) s_ref[sp + b3_][tid] = rf[3];
          } else {
            if (h[0] && b0_ != first) s_ent[sp + b0_][tid] = make_uint2(rf[0], __float_as_uint(k[0]));
            if (h[1] && b1_ != first) s_ent[sp + b1_][tid] = make_uint2(rf[1], __float_as_uint(k[1]));
            if (h[2] && b2_ != first) s_ent[sp + b2_][tid] = make_uint2(rf[2], __float_as_uint(k[2]));
            if (h[3] && b3_ != first) s_ent[sp + b3_][tid] = make_uint2(rf[3], __float_as_uint(k[3]));
          }
          sp += first;
        } else {  // deep stack (rare): one entry at a time through the shared / local / overflow cases, last visited first
          for (int want = 0; want < first; want++) {
            if (h[0] && b0_ == want) RT_PUSH(rf[0], k[0]);
            if (h[1] && b1_ == want) RT_PUSH(rf[1], k[1]);
            if (h[2] && b2_ == want) RT_PUSH(rf[2], k[2]);
            if (h[3] && b3_ == want) RT_PUSH(rf[3], k[3]);
          }
        }
      }
    }
    if (!WIDE && !leaf_phase && act && !atleaf) {
      // ---- node phase ----
      if (COUNT) n_nodes++;
      const float4* np = sc.nodes + 4 * (size_t)cur;
      float4 l0, l1, r0, r1;
      rt_ld256(np, l0, l1); rt_ld256(np + 2, r0, r1);
      float el, er;
      V3 tnl, tnr;
      // exact lane: order-independent part of the box test here, distance part below; otherwise the full inclusive test
      bool hl = rt_slab_tn2<true>(l0, l1, q2, (EXACT || SUN) ? INFINITY : tmax, el, tnl);
      bool hr = rt_slab_tn2<true>(r0, r1, q2, (EXACT || SUN) ? INFINITY : tmax, er, tnr);
      const unsigned refl = __float_as_uint(l1.z), refr = __float_as_uint(r1.z);
      // (Prefetching both children's records into L1 once the queue is empty, so that the next step of the last, slowest
      // rays overlaps its fetch with this step's arithmetic, was measured: per-launch fixed cost 0.50-0.52 ms against
      // 0.49-0.50 without, 4 more registers.)
      float kl = el, kr = er;  // culling keys: what the distance test, now and at pop time, compares with tmax
      if (EXACT && !SUN) {
        // a subtree may be skipped only if even its slack-inflated box is entered beyond the current hit
        // (sun shadow rays: tcull stays +INF, nothing is culled by distance)
        const V3 inv = rt_ray2_inv(q2);
        kl = rt_key_inflated(tnl, l1.w, inv, omk); kr = rt_key_inflated(tnr, r1.w, inv, omk);
        hl = hl && !(kl > tcut); hr = hr && !(kr > tcut);
      }
      // nearer hit child next (leaf or not), the other one on the stack
      const bool right_first = hr && (!hl || er < el);
      if (hl && hr) RT_PUSH(right_first ? refl : refr, right_first ? kl : kr);
      cur = right_first ? refr : refl;
      need_pop = !(hl || hr);
    }
#endif
    // ---- next position from the stack, skipping entries the current hit has made unreachable ----
    if (need_pop) {
      st = DONE;
      if (SUN) {
        if (sp > 0) { sp--; cur = sp < RT_STACK_SHARED ? s_ref[sp][tid] : loc_ref[sp - RT_STACK_SHARED]; st = ACTIVE; }
      } else {
        const float tpop = EXACT ? tcut : tmax;
        while (sp > 0) {
          sp--;
          const uint2 e = sp < RT_STACK_SHARED ? s_ent[sp][tid] : loc_ent[sp - RT_STACK_SHARED];
          if (!(__uint_as_float(e.y) > tpop)) { cur = e.x; st = ACTIVE; break; }
        }
      }
    }
  }
#undef RT_PUSH
  if (COUNT) {
    for (int o2 = 16; o2 > 0; o2 >>= 1) { n_nodes += __shfl_xor_sync(0xffffffffu, n_nodes, o2); n_leaves += __shfl_xor_sync(0xffffffffu, n_leaves, o2); }
    if (lane == 0) { atomicAdd(stats, n_nodes); atomicAdd(stats + 1, n_leaves); }
#ifdef RT_LANE_STATS
    if (lane == 0) {
      atomicAdd(stats + 3, ls_iter[0]); atomicAdd(stats + 4, ls_iter[1]); atomicAdd(stats + 5, ls_lanes[0]); atomicAdd(stats + 6, ls_lanes[1]);
      atomicAdd(stats + 7, ls_parked); atomicAdd(stats + 8, ls_idle);
    }
#endif
  }
}

// ---- packet kernel: one warp walks the tree ONCE for 32 consecutive rays ----------------------------------
// The queue order is pixel-major (render.cu, generate_kernel): 32 consecutive rays are the samples of one or two
// pixels — the reference's own mapping, one warp = one pixel's 32 samples (draw.cu:135-213) — and the shadow rays
// of their hits.  Such rays visit almost the same nodes, so the warp keeps ONE position and ONE stack (shared
// memory, warp-uniform pushes and pops, no per-lane state machine, no refill logic) and every node and leaf record is
// fetched once per warp: all lanes read the same address, one L1TEX wavefront instead of up to 32.
//
// Exactness.  The exact lane's answer is defined on leaves (see the top of this file): a ray's candidates are the
// leaves whose OWN box passes the order-independent test and whose primitive test accepts, whatever the order of the
// visits and whatever else is visited.  A lane therefore runs its own box tests (the same arithmetic as everywhere)
// and takes part in a leaf test only when the leaf's box — a child half of the node just tested — passed for its own
// ray and is not culled by its own key; an inner child is entered when ANY lane wants it.  A lane that did not want
// a subtree is masked out below it (the mask travels with the stack entry).  Culling at pop time uses the smallest
// key among the lanes that wanted the entry against the largest cut-off distance among the live lanes: skipped only
// when every lane alone would have skipped it.  Near-first order is decided by majority.  Certificate, t2 and the
// fallback list work per lane exactly as in traverse_near_kernel (any superset of tested candidates keeps the
// argument: t2 only gets smaller).
#ifndef RT_PKT_THREADS
#define RT_PKT_THREADS 128
#endif
#ifndef RT_PKT_MIN_CTAS
#define RT_PKT_MIN_CTAS 1
#endif
__device__ __forceinline__ unsigned rt_warp_min_u32(unsigned v) { return __reduce_min_sync(0xffffffffu, v); }
__device__ __forceinline__ unsigned rt_warp_max_u32(unsigned v) { return __reduce_max_sync(0xffffffffu, v); }

template <bool EXACT, bool COUNT, bool SUN>
__global__ void __launch_bounds__(RT_PKT_THREADS, RT_PKT_MIN_CTAS) traverse_packet_kernel(const RtDeviceScene sc, const float4* __restrict__ rays, const unsigned first_in,
                                                                       const unsigned n_in, float2* __restrict__ hits, unsigned* fetch_counter,
                                                                       unsigned long long* stats, unsigned* __restrict__ fb_list,
                                                                       unsigned* fb_count, const unsigned* __restrict__ n_dev, const unsigned cap_back) {
  __shared__ unsigned s_range[2];
  if (threadIdx.x == 0) { const unsigned nn = n_dev ? *n_dev : n_in; s_range[0] = nn; s_range[1] = (n_dev && cap_back) ? cap_back - nn : first_in; }
  __syncthreads();
  constexpr bool LEAF_BOX_FIRST = false;  // binary nodes hold the leaves' exact boxes
  // warp-uniform stack entry: (reference, smallest culling key among the lanes that want it, their lane mask)
  __shared__ uint4 s_stk[RT_PKT_THREADS / 32][RT_STACK_TOTAL];
  uint4* const stk = s_stk[threadIdx.x >> 5];
  const unsigned lane = threadIdx.x & 31;
  const unsigned n = s_range[0], first = s_range[1];
  unsigned long long n_nodes = 0, n_leaves = 0, n_steps = 0;
  const unsigned root = sc.n > 0 ? (unsigned)__ldg(sc.root) : 0u;
  while (true) {
    unsigned base = 0;
    if (lane == 0) base = atomicAdd(fetch_counter, 32u);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= n) break;
    const unsigned r = first + base + lane;
    bool valid = base + lane < n;
    float4 q1 = make_float4(0.f, 0.f, 0.f, 0.f), q0 = q1;
    if (valid) { q1 = __ldcs(rays + 2 * (size_t)r + 1); valid = __float_as_uint(q1.w) != RT_RAY_DEAD; }
    if (valid) q0 = __ldcs(rays + 2 * (size_t)r);
    const V3 d = mk3(q1.x, q1.y, q1.z);
    const float tlimit = q0.w;
    const RtRay2 q2 = rt_ray2(mk3(q0.x, q0.y, q0.z), mk3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z));  // bvh_traversal.cu:106
    const float omk = (RT_ORIGIN_SLACK * 1.000001f) * fmaxf(fmaxf(fabsf(q0.x), fabsf(q0.y)), fabsf(q0.z));
    const bool anyhit = SUN || (__float_as_uint(q1.w) & RT_RAY_ANYHIT) != 0u;
    unsigned fl = (anyhit ? F_ANYHIT : 0u) | ((!SUN && !(anyhit && tlimit == INFINITY)) ? F_WANTKEY : 0u);
    float tmax = tlimit, tcull = tlimit, tcut = tlimit * 1.000001f, t2 = INFINITY;
    int best = -1;
    bool alive = valid;               // still searching
    bool part = alive;                // wants the current node
    unsigned cur = root; int sp = 0;
    bool overflow = false;
    // largest cut-off among the live lanes, as ordered bits (distances are >= 0): what a stack entry's key must not exceed
    unsigned cutmax = SUN ? 0x7F800000u : rt_warp_max_u32(alive ? __float_as_uint(tcut) : 0u);
    if (__ballot_sync(0xffffffffu, alive) != 0u && sc.n > 0) {
      if (cur & RT_LEAF_BIT) {  // n == 1: the root is the leaf, tested without a box as in traverse_near_kernel
        if (alive) { RT_LEAF_TEST(cur); }
      } else
      while (true) {
        // ---- node step: both child boxes for every lane's own ray ----
        if (COUNT) { n_nodes += part ? 1u : 0u; n_steps += lane == 0 ? 1u : 0u; }
        const float4* np = sc.nodes + 4 * (size_t)cur;
        float4 l0, l1, r0, r1;
        rt_ld256(np, l0, l1); rt_ld256(np + 2, r0, r1);
        float el, er; V3 tnl, tnr;
        bool hl = rt_slab_tn2<true>(l0, l1, q2, (EXACT || SUN) ? INFINITY : tmax, el, tnl) && part;
        bool hr = rt_slab_tn2<true>(r0, r1, q2, (EXACT || SUN) ? INFINITY : tmax, er, tnr) && part;
        const unsigned refl = __float_as_uint(l1.z), refr = __float_as_uint(r1.z);
        float kl = el, kr = er;
        if (EXACT && !SUN) {
          const V3 inv = rt_ray2_inv(q2);
          kl = rt_key_inflated(tnl, l1.w, inv, omk); kr = rt_key_inflated(tnr, r1.w, inv, omk);
          hl = hl && !(kl > tcut); hr = hr && !(kr > tcut);
        }
        const unsigned ml = __ballot_sync(0xffffffffu, hl), mr = __ballot_sync(0xffffffffu, hr);
        // nearer child first, by majority of the lanes that hit anything
        const unsigned vr = __ballot_sync(0xffffffffu, hr && (!hl || er < el));
        const bool right_first = __popc(vr) > __popc((ml | mr) & ~vr);
        // ---- leaf children: tested now, by the lanes whose own test passed the leaf's box ----
        bool took = false;
#pragma unroll 1
        for (int k = 0; k < 2; k++) {
          const bool isr = (k == 0) == right_first;
          const unsigned ref = isr ? refr : refl, m = isr ? mr : ml;
          if ((ref & RT_LEAF_BIT) && m) {  // warp-uniform
            const bool h = isr ? hr : hl; const float key = isr ? kr : kl;
            if (h && alive && !(EXACT && !SUN && key > tcut)) {
              const int before = best; const float tbefore = tmax;
              RT_LEAF_TEST(ref);
              took = took || best != before || tmax != tbefore || (fl & F_FALLBACK);
              if (((SUN || (fl & F_ANYHIT)) && best >= 0) || (fl & F_FALLBACK)) alive = false;
            }
          }
        }
        if (__ballot_sync(0xffffffffu, took)) {
          if (__ballot_sync(0xffffffffu, alive) == 0u) break;  // every ray has its answer
          if (!SUN) cutmax = rt_warp_max_u32(alive ? __float_as_uint(tcut) : 0u);
        }
        // ---- inner children ----
        const bool il = !(refl & RT_LEAF_BIT) && ml, ir = !(refr & RT_LEAF_BIT) && mr;  // warp-uniform
        bool pop = false;
        if (il && ir) {
          const unsigned refp = right_first ? refl : refr, mp = right_first ? ml : mr;
          unsigned kp = 0u;
          if (!SUN) kp = rt_warp_min_u32((right_first ? hl : hr) ? __float_as_uint(fmaxf(right_first ? kl : kr, 0.0f)) : 0x7F800000u);
          if (sp < RT_STACK_TOTAL) { stk[sp] = make_uint4(refp, kp, mp, 0u); sp++; }
          else overflow = true;  // the reference drops an entry here (bvh_traversal.cu:157-164): let its own order decide
          cur = right_first ? refr : refl; part = (right_first ? hr : hl) && alive;
        } else if (il) { cur = refl; part = hl && alive; }
        else if (ir) { cur = refr; part = hr && alive; }
        else pop = true;
        if (pop) {
          bool found = false;
          while (sp > 0) {
            sp--;
            const uint4 e = stk[sp];
            if (SUN || !(e.y > cutmax)) { cur = e.x; part = ((e.z >> lane) & 1u) != 0u && alive; found = true; break; }
          }
          if (!found) break;
        }
      }
    }
    if (valid) {
      if (overflow) fl |= F_FALLBACK;
      if (EXACT && ((fl & F_FALLBACK) || ((fl & F_INCONS) && !(t2 > tcull)))) fb_list[atomicAdd(fb_count, 1u)] = r;
      else {
        const RtHit h = rt_finish_with_planes(sc, rt_ray2_o(q2), d, tlimit, SUN || (fl & F_ANYHIT), tmax, best);
        __stcs(hits + r, make_float2(h.t, __int_as_float(h.id)));
      }
    }
  }
  if (COUNT) {
    for (int o2 = 16; o2 > 0; o2 >>= 1) {
      n_nodes += __shfl_xor_sync(0xffffffffu, n_nodes, o2); n_leaves += __shfl_xor_sync(0xffffffffu, n_leaves, o2); n_steps += __shfl_xor_sync(0xffffffffu, n_steps, o2);
    }
    if (lane == 0) { atomicAdd(stats, n_nodes); atomicAdd(stats + 1, n_leaves); atomicAdd(stats + 9, n_steps); }
  }
}
#undef RT_LEAF_TEST
