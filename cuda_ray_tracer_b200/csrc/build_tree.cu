// build_tree.cu — hierarchy + AABB refit (K3..K6) and rt_build_lbvh.
//
// Replaces initialize_leaf_nodes_kernel, generate_internal_nodes_karas_kernel,
// initialize_visited_counters_kernel and set_aabb_kernel_adapted (src/lbvh_builder.cu:59-71,224-322,
// 389-397,324-387) and the orchestration in build_lbvh_karas (:401-521).
//
// The tree is the binary radix tree of Karras 2012 over the augmented keys (code, sorted position):
// delta(i,j) = clz(code_i ^ code_j), or bits + clz(i ^ j) for equal codes (lbvh_builder.cu:76-101).
// Two builders produce the same tree:
//  * default, fused bottom-up (Apetrei 2014): one thread per leaf climbs; at every step the node
//    [l,r] joins the neighbour it shares the longer prefix with (delta(r,r+1) > delta(l-1,l) ⇒ it is
//    the LEFT child of the node that splits after r, else the RIGHT child of the node that splits
//    after l-1).  The first child to arrive deposits its 32-byte half (box + reference) in the
//    parent's slot and leaves; the second unions the two boxes and continues.  Topology and refit
//    are one kernel, one atomic per node, and only adjacent deltas are evaluated.  Nodes are stored
//    at slot = split position; the reference's numbering (internal index = split of the parent for a
//    left child, split+1 for a right child, root 0) is recovered on export.
//  * RT_BUILD_TOPDOWN: per-node range and split search exactly as the reference does it, in the
//    reference's numbering, followed by the atomic bottom-up refit with the fence the reference
//    lacks.
#include <cstdlib>
#include <cstring>
#include "build.cuh"

namespace {

template <typename KeyT>
struct Delta {
  const KeyT* keys; int n;
  __device__ inline int adj(int i) const {  // delta(i, i+1); -1 outside (lbvh_builder.cu:84-87)
    if (i < 0 || i >= n - 1) return -1;
    return pair(i, i + 1);
  }
  __device__ inline int pair(int a, int b) const {
    if (a < 0 || a >= n || b < 0 || b >= n) return -1;
    KeyT ka = keys[a], kb = keys[b];
    if (ka == kb) return (int)(8 * sizeof(KeyT)) + __clz((unsigned)a ^ (unsigned)b);  // lbvh_builder.cu:92-98
    return sizeof(KeyT) == 4 ? __clz((unsigned)(ka ^ kb)) : __clzll((long long)(ka ^ kb));
  }
};

// aux = (hit slack of the subtree (build.cuh, rt_leaf_slack) + 4e-7 * largest |coordinate| of the box) * 1.000001:
// the amount by which the traversal inflates the box before it culls the subtree by distance, with the rounding
// allowance of its key already applied (traverse.cuh, rt_key_inflated)
__device__ inline void store_half(float4* nodes, int slot, int side, const RtBox& b, uint32_t ref, float slack) {
  float4* h = nodes + 4 * (size_t)slot + 2 * side;
  const float cmax = fmaxf(fmaxf(fmaxf(fabsf(b.xmin), fabsf(b.xmax)), fmaxf(fabsf(b.ymin), fabsf(b.ymax))), fmaxf(fabsf(b.zmin), fabsf(b.zmax)));
  __stcg(h, make_float4(b.xmin, b.ymin, b.xmax, b.ymax));
  __stcg(h + 1, make_float4(b.zmin, b.zmax, __uint_as_float(ref), (slack + 4e-7f * cmax) * 1.000001f));
}
__device__ inline float load_half_box(const float4* nodes, int slot, int side, RtBox& b) {
  const float4* h = nodes + 4 * (size_t)slot + 2 * side;
  float4 u = __ldcg(h), v = __ldcg(h + 1);
  b.xmin = u.x; b.ymin = u.y; b.xmax = u.z; b.ymax = u.w; b.zmin = v.x; b.zmax = v.y;
  return v.w;
}
__device__ inline RtBox union_box(const RtBox& l, const RtBox& r) {  // AABB(left,right), interval.cuh:83-88
  RtBox o;
  o.xmin = fminf(l.xmin, r.xmin); o.xmax = fmaxf(l.xmax, r.xmax);
  o.ymin = fminf(l.ymin, r.ymin); o.ymax = fmaxf(l.ymax, r.ymax);
  o.zmin = fminf(l.zmin, r.zmin); o.zmax = fmaxf(l.zmax, r.zmax);
  return o;
}

// Arrival handshake of the bottom-up passes: one exchange with acquire-release ordering at device scope.  The
// release half publishes this thread's 32-byte node half (written before it), the acquire half makes the sibling's
// visible to the loads after it — the ordering the reference's refit lacks — without the two full
// membar.gl fences (which on Blackwell also invalidate L1) that __threadfence() would cost per level.
__device__ __forceinline__ int exch_acq_rel(int* addr, int v) {
  int old;
  asm volatile("atom.acq_rel.gpu.global.exch.b32 %0, [%1], %2;" : "=r"(old) : "l"(addr), "r"(v) : "memory");
  return old;
}

__device__ __forceinline__ int ld_acquire(const int* addr) {
  int v;
  asm volatile("ld.acquire.gpu.global.b32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}

// ---- fused bottom-up builder ------------------------------------------------------------------
template <typename KeyT>
__global__ void __launch_bounds__(256) bottom_up_kernel(const KeyT* __restrict__ keys, const uint32_t* __restrict__ perm,
                                                       const float4* __restrict__ bv, const float4* __restrict__ tv,
                                                       float4* __restrict__ sorted_tv, float4* nodes, int* flags, int n,
                                                       int* root_out, float* root_box) {
  rt_pdl_wait();
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const uint32_t p = perm[k];
  const float4 a = __ldg(bv + RT_PRIM_STRIDE * (size_t)p), b = __ldg(bv + RT_PRIM_STRIDE * (size_t)p + 1), c = __ldg(bv + RT_PRIM_STRIDE * (size_t)p + 2);
  // gather the trace view into Morton order (leaf k <-> record k)
  const float4 t0 = __ldg(tv + RT_PRIM_STRIDE * (size_t)p), t1 = __ldg(tv + RT_PRIM_STRIDE * (size_t)p + 1), t2 = __ldg(tv + RT_PRIM_STRIDE * (size_t)p + 2);
  RtBox box = rt_prim_box(a, b, c);
  {
    float4* rec = sorted_tv + RT_TV_STRIDE * (size_t)k;
    rec[0] = t0; rec[1] = t1; rec[2] = t2; rec[3] = make_float4(__uint_as_float(p), 0.f, 0.f, 0.f);
    rec[4] = make_float4(box.xmin, box.ymin, box.xmax, box.ymax); rec[5] = make_float4(box.zmin, box.zmax, 0.f, 0.f);
  }
  float slack = c.y;  // hit slack of the primitive, computed with the triangle setup (scene.cu, build.cuh)
  uint32_t ref = RT_LEAF_BIT | (c.w == RT_TAG_SPHERE ? RT_SPHERE_BIT : 0u) | (uint32_t)k;
  Delta<KeyT> D{keys, n};
  int l = k, r = k;
  while (true) {
    if (l == 0 && r == n - 1) {  // root
      *root_out = (int)ref;
      root_box[0] = box.xmin; root_box[1] = box.xmax; root_box[2] = box.ymin; root_box[3] = box.ymax; root_box[4] = box.zmin; root_box[5] = box.zmax;
      return;
    }
    const int dl = D.adj(l - 1), dr = D.adj(r);
    int s, side, mine;
    if (dr > dl) { s = r; side = 0; mine = l; } else { s = l - 1; side = 1; mine = r; }
    store_half(nodes, s, side, box, ref, slack);
    const int other = exch_acq_rel(flags + s, mine);
    if (other == -1) return;  // sibling not there yet: it will finish this node
    RtBox sib;
    slack = fmaxf(slack, load_half_box(nodes, s, side ^ 1, sib));
    if (side == 0) { r = other; box = union_box(box, sib); } else { l = other; box = union_box(sib, box); }
    ref = (uint32_t)s;
  }
}

// ---- fused bottom-up builder, lower levels inside the warp (default since round 2) -----------------------------------
// ncu on bottom_up_kernel (profiles/ncu_build_r02.txt): 9 of 32 lanes per instruction, 14 % of the issue slots, long_scoreboard
// 36 per issue: a warp stays resident until its last lane has finished a climb of dependent device-scope exchanges, and
// most of that time most of its lanes are dead.  Here the 32 consecutive leaves of a warp first meet INSIDE the warp: in
// every round each live lane (a finished subtree [l, r]) names its parent s and side exactly as above; two live lanes that
// name the same parent are its two children (the left one looks into the lane where its sibling must sit), the left one takes the right one's half by shuffle, writes
// the finished 64-byte node and goes on with the union, the right one retires.  No flag, no atomic, no global read for a
// node whose children both lie in the warp (about 26 of 31).  Rounds stop when none pairs up; what is left (subtrees whose
// sibling reaches beyond the warp's 32 leaves) continues with the global protocol of bottom_up_kernel.  Same tree, same
// boxes: fminf / fmaxf unions do not depend on who forms them (tests/test_build_gpu.py runs every size with both kernels).
// Measured (profiles/build_climb_r02.txt) against the round-1 kernel, with the survivor compaction and the look before the
// exchange below: 1M 0.295 vs 0.302 ms, 10M 2.117 vs 2.304 ms (-8 %).  The split the leaf-only run gives at 1M: codes + sort
// 0.136, leaf records 0.054, climb 0.105 ms - of which ~0.04 is the chain to the root (the whole climb of a 100K build,
// one wave of blocks, 17 levels, takes 0.037 ms) - so what happens to the lower levels moves little.
// mode bit 0: TIMING EXPERIMENT ONLY (scripts/build_ab.py, RT_BU_LEAFONLY): stop after the leaf records, no tree.
template <typename KeyT>
__global__ void __launch_bounds__(256) bottom_up_warp_kernel(const KeyT* __restrict__ keys, const uint32_t* __restrict__ perm,
                                                            const float4* __restrict__ bv, const float4* __restrict__ tv,
                                                            float4* __restrict__ sorted_tv, float4* nodes, int* flags, int n,
                                                            int* root_out, float* root_box, const int mode) {
  rt_pdl_wait();  // last link of the build's chain (build.cuh): the sort's last pass must be complete
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const unsigned lane = threadIdx.x & 31;
  const int k0 = k - (int)lane, klast = min(k0 + 31, n - 1);  // the warp's leaves
  bool active = k < n;
  RtBox box; box.xmin = box.ymin = box.zmin = box.xmax = box.ymax = box.zmax = 0.f;
  float slack = 0.f; uint32_t ref = 0u;
  if (active) {
    const uint32_t p = perm[k];
    const float4 a = __ldg(bv + RT_PRIM_STRIDE * (size_t)p), b = __ldg(bv + RT_PRIM_STRIDE * (size_t)p + 1), c = __ldg(bv + RT_PRIM_STRIDE * (size_t)p + 2);
    const float4 t0 = __ldg(tv + RT_PRIM_STRIDE * (size_t)p), t1 = __ldg(tv + RT_PRIM_STRIDE * (size_t)p + 1), t2 = __ldg(tv + RT_PRIM_STRIDE * (size_t)p + 2);
    box = rt_prim_box(a, b, c);
    float4* rec = sorted_tv + RT_TV_STRIDE * (size_t)k;
    rec[0] = t0; rec[1] = t1; rec[2] = t2; rec[3] = make_float4(__uint_as_float(p), 0.f, 0.f, 0.f);
    rec[4] = make_float4(box.xmin, box.ymin, box.xmax, box.ymax); rec[5] = make_float4(box.zmin, box.zmax, 0.f, 0.f);
    slack = c.y;
    ref = RT_LEAF_BIT | (c.w == RT_TAG_SPHERE ? RT_SPHERE_BIT : 0u) | (uint32_t)k;
  }
  if (mode & 1) return;
  Delta<KeyT> D{keys, n};
  int l = k, r = k, s = -1, side = 0, dl = -1, dr = -1;
  bool need = true;  // (s, side) must be (re)computed: the range has changed
  while (true) {     // ---- rounds inside the warp ----
    if (active && need) {
      if (l == 0 && r == n - 1) {  // root (n <= 32, or the first warp of a tiny tree)
        *root_out = (int)ref;
        root_box[0] = box.xmin; root_box[1] = box.xmax; root_box[2] = box.ymin; root_box[3] = box.ymax; root_box[4] = box.zmin; root_box[5] = box.zmax;
        active = false;
      } else {
        dl = D.adj(l - 1); dr = D.adj(r);
        if (dr > dl) { s = r; side = 0; } else { s = l - 1; side = 1; }
        need = false;
      }
    }
    // A finished subtree [l, r] is always held by the lane of its leftmost leaf (the left child goes on, the right one
    // retires), so the right sibling of a left child [l, s], if it is finished, sits in lane s + 1 - k0 and says so: a right
    // child announces its parent.  One shuffle and one warp-wide OR per round; the MATCH.ANY this replaces resolves one
    // distinct value after the other (see rt_match_digit8, build_sort.cu).
    const bool left_cand = active && side == 0 && s + 1 <= klast;
    const int want = left_cand ? s + 1 - k0 : (int)lane;
    const int tag = (active && side == 1) ? s : -2;
    const int ptag = __shfl_sync(0xffffffffu, tag, want);  // every lane takes part: not behind the && below
    const bool pair_left = left_cand && ptag == s;
    const unsigned taken = __reduce_or_sync(0xffffffffu, pair_left ? 1u << want : 0u);
    if (taken == 0u) break;
    const bool pair = pair_left || ((taken >> lane) & 1u);
    const int partner = pair_left ? want : (int)lane;
    RtBox pb;
    pb.xmin = __shfl_sync(0xffffffffu, box.xmin, partner); pb.ymin = __shfl_sync(0xffffffffu, box.ymin, partner); pb.zmin = __shfl_sync(0xffffffffu, box.zmin, partner);
    pb.xmax = __shfl_sync(0xffffffffu, box.xmax, partner); pb.ymax = __shfl_sync(0xffffffffu, box.ymax, partner); pb.zmax = __shfl_sync(0xffffffffu, box.zmax, partner);
    const float pslack = __shfl_sync(0xffffffffu, slack, partner);
    const uint32_t pref = __shfl_sync(0xffffffffu, ref, partner);
    const int pr = __shfl_sync(0xffffffffu, r, partner);
    if (pair) {
      if (side == 0) {  // the left child writes the finished node and goes on
        store_half(nodes, s, 0, box, ref, slack);
        store_half(nodes, s, 1, pb, pref, pslack);
        box = union_box(box, pb); slack = fmaxf(slack, pslack); r = pr; ref = (uint32_t)s; need = true;
      } else active = false;  // the right child has handed its half over
    }
  }
  // here every live lane knows dl = delta(l-1, l), dr = delta(r, r+1) of its subtree, and none is the root
  if (mode & 2) {  // the block's survivors (about 9 per warp) move into its first warps, the other warps leave the SM
    __shared__ int s_cnt;
    __shared__ float s_state[12][256];
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    const unsigned m = __ballot_sync(0xffffffffu, active);
    int base = 0;
    if (lane == 0 && m) base = atomicAdd(&s_cnt, __popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (active) {
      const int q = base + __popc(m & ((1u << lane) - 1u));
      s_state[0][q] = box.xmin; s_state[1][q] = box.ymin; s_state[2][q] = box.zmin; s_state[3][q] = box.xmax; s_state[4][q] = box.ymax; s_state[5][q] = box.zmax;
      s_state[6][q] = slack; s_state[7][q] = __uint_as_float(ref); s_state[8][q] = __int_as_float(l); s_state[9][q] = __int_as_float(r);
      s_state[10][q] = __int_as_float(dl); s_state[11][q] = __int_as_float(dr);
    }
    __syncthreads();
    const int q = threadIdx.x;
    active = q < s_cnt;
    if (active) {
      box.xmin = s_state[0][q]; box.ymin = s_state[1][q]; box.zmin = s_state[2][q]; box.xmax = s_state[3][q]; box.ymax = s_state[4][q]; box.zmax = s_state[5][q];
      slack = s_state[6][q]; ref = __float_as_uint(s_state[7][q]); l = __float_as_int(s_state[8][q]); r = __float_as_int(s_state[9][q]);
      dl = __float_as_int(s_state[10][q]); dr = __float_as_int(s_state[11][q]);
    }
  }
  if (!active) return;
  while (true) {  // ---- what is left: the global protocol of bottom_up_kernel, one dependent round trip shorter per level ----
    int mine;
    if (dr > dl) { s = r; side = 0; mine = l; } else { s = l - 1; side = 1; mine = r; }
    store_half(nodes, s, side, box, ref, slack);
    int other = -1;
    // look before the exchange: a child that finds its sibling's mark needs no release of its own half (nobody reads it
    // through this flag) and skips MEMBAR.ALL.GPU + the atomic's round trip; the chain that ends at the root is the late
    // child of every level.  -3 % at 1M, -2 % at 10M (profiles/build_climb_r02.txt)
    if (mode & 4) other = ld_acquire(flags + s);
    if (other == -1) other = exch_acq_rel(flags + s, mine);
    if (other == -1) return;
    // both loads that depend on the exchange go out together: the sibling's half and the two keys at the new far edge
    // (the near edge's delta is unchanged)
    RtBox sib;
    const float sslack = load_half_box(nodes, s, side ^ 1, sib);
    const int dnew = side == 0 ? D.adj(other) : D.adj(other - 1);
    slack = fmaxf(slack, sslack);
    if (side == 0) { r = other; dr = dnew; box = union_box(box, sib); } else { l = other; dl = dnew; box = union_box(sib, box); }
    ref = (uint32_t)s;
    if (l == 0 && r == n - 1) {
      *root_out = (int)ref;
      root_box[0] = box.xmin; root_box[1] = box.xmax; root_box[2] = box.ymin; root_box[3] = box.ymax; root_box[4] = box.zmin; root_box[5] = box.zmax;
      return;
    }
  }
}

// ---- fused bottom-up builder, block-local lower levels (RT_BUILD_LOCAL_CLIMB) ---------------------
// The climb above costs one device-scope exchange and one dependent 32-byte read per level: ~20 L2 round trips
// in a row per surviving thread.  Here a block owns RT_BU_CHUNK consecutive leaves.  A node whose leaves all lie
// in the chunk ("local") is completed by this block alone, so its arrival at a parent whose split also lies
// inside the chunk goes through SHARED memory: deposit the half, one shared exchange, the second child reads the
// sibling from shared memory and writes the finished 64-byte node to global memory in one piece.  Everything
// else uses the global protocol of bottom_up_kernel:
//   * a local node whose parent splits at the chunk border arrives globally right after the local phase;
//   * a local node left waiting in shared memory when the local phase ends has a sibling that reaches beyond the
//     chunk (that sibling can only be finished through global arrivals): the thread that owns the slot re-issues
//     the arrival globally;
//   * once a thread's node reaches beyond the chunk it stays on the global protocol (ranges only grow).
// Both children of a split therefore always meet in the same place, and only O(log) nodes per chunk (~3 % at
// 512 leaves) touch the global flags.  Same tree, same boxes (tests/test_build_gpu.py runs both builders).
// MEASURED (B200, round 1): 0.320 ms vs 0.309 ms at 1M triangles, 2.36 vs 2.31 ms at 10M — no gain, so the global
// climb stays the default.  The kernel is bound by the random 2 x 48-byte gather of the primitive records through
// the sorted permutation (HBM random-gather rate, profiles/l2_peaks_r01.json), not by the arrival atomics.
#define RT_BU_CHUNK 512

__device__ __forceinline__ float rt_node_aux(const RtBox& b, float slack) {
  const float cmax = fmaxf(fmaxf(fmaxf(fabsf(b.xmin), fabsf(b.xmax)), fmaxf(fabsf(b.ymin), fabsf(b.ymax))), fmaxf(fabsf(b.zmin), fabsf(b.zmax)));
  return (slack + 4e-7f * cmax) * 1.000001f;
}

template <typename KeyT>
__device__ __forceinline__ void climb_global(const Delta<KeyT>& D, float4* nodes, int* flags, int n, int* root_out, float* root_box,
                                             int l, int r, RtBox box, uint32_t ref, float slack) {
  while (true) {
    if (l == 0 && r == n - 1) {  // root
      *root_out = (int)ref;
      root_box[0] = box.xmin; root_box[1] = box.xmax; root_box[2] = box.ymin; root_box[3] = box.ymax; root_box[4] = box.zmin; root_box[5] = box.zmax;
      return;
    }
    const int dl = D.adj(l - 1), dr = D.adj(r);
    int s, side, mine;
    if (dr > dl) { s = r; side = 0; mine = l; } else { s = l - 1; side = 1; mine = r; }
    store_half(nodes, s, side, box, ref, slack);
    const int other = exch_acq_rel(flags + s, mine);
    if (other == -1) return;  // sibling not there yet: it will finish this node
    RtBox sib;
    slack = fmaxf(slack, load_half_box(nodes, s, side ^ 1, sib));
    if (side == 0) { r = other; box = union_box(box, sib); } else { l = other; box = union_box(sib, box); }
    ref = (uint32_t)s;
  }
}

template <typename KeyT>
__global__ void __launch_bounds__(RT_BU_CHUNK) bottom_up_local_kernel(const KeyT* __restrict__ keys, const uint32_t* __restrict__ perm,
                                                                     const float4* __restrict__ bv, const float4* __restrict__ tv,
                                                                     float4* __restrict__ sorted_tv, float4* nodes, int* flags, int n,
                                                                     int* root_out, float* root_box) {
  __shared__ int s_delta[RT_BU_CHUNK + 1];      // s_delta[i] = delta(c0 - 1 + i, c0 + i)
  __shared__ int s_flag[RT_BU_CHUNK];           // split after leaf c0 + i: -1 nobody, (range end << 1 | side) one child waits, -2 finished
  __shared__ float4 s_half[RT_BU_CHUNK][2][2];  // [split][side]: (xmin ymin xmax ymax | zmin zmax ref raw-slack)
  const int t = threadIdx.x, c0 = blockIdx.x * RT_BU_CHUNK, k = c0 + t;
  const int c1 = min(c0 + RT_BU_CHUNK, n) - 1;  // last leaf of the chunk
  Delta<KeyT> D{keys, n};
  s_delta[t] = D.adj(k - 1);
  if (t == 0) s_delta[RT_BU_CHUNK] = D.adj(c0 + RT_BU_CHUNK - 1);
  s_flag[t] = -1;
  int l = k, r = k;
  RtBox box; uint32_t ref = 0; float slack = 0.f;
  if (k < n) {
    const uint32_t p = perm[k];
    const float4 a = __ldg(bv + RT_PRIM_STRIDE * (size_t)p), b = __ldg(bv + RT_PRIM_STRIDE * (size_t)p + 1), c = __ldg(bv + RT_PRIM_STRIDE * (size_t)p + 2);
    const float4 t0 = __ldg(tv + RT_PRIM_STRIDE * (size_t)p), t1 = __ldg(tv + RT_PRIM_STRIDE * (size_t)p + 1), t2 = __ldg(tv + RT_PRIM_STRIDE * (size_t)p + 2);
    box = rt_prim_box(a, b, c);
    float4* rec = sorted_tv + RT_TV_STRIDE * (size_t)k;  // the trace view in Morton order (leaf k <-> record k)
    rec[0] = t0; rec[1] = t1; rec[2] = t2; rec[3] = make_float4(__uint_as_float(p), 0.f, 0.f, 0.f);
    rec[4] = make_float4(box.xmin, box.ymin, box.xmax, box.ymax); rec[5] = make_float4(box.zmin, box.zmax, 0.f, 0.f);
    slack = c.y;  // hit slack of the primitive, computed with the triangle setup (scene.cu, build.cuh)
    ref = RT_LEAF_BIT | (c.w == RT_TAG_SPHERE ? RT_SPHERE_BIT : 0u) | (uint32_t)k;
  }
  __syncthreads();
  bool go_global = false;
  if (k < n) {
    while (true) {
      if (l == 0 && r == n - 1) {  // root of a tree that fits one chunk
        *root_out = (int)ref;
        root_box[0] = box.xmin; root_box[1] = box.xmax; root_box[2] = box.ymin; root_box[3] = box.ymax; root_box[4] = box.zmin; root_box[5] = box.zmax;
        break;
      }
      const int dl = s_delta[l - c0], dr = s_delta[r - c0 + 1];
      int s, side, mine;
      if (dr > dl) { s = r; side = 0; mine = l; } else { s = l - 1; side = 1; mine = r; }
      if (s < c0 || s >= c1) { go_global = true; break; }  // the parent splits at the chunk border
      const int q = s - c0;
      const float4 h0 = make_float4(box.xmin, box.ymin, box.xmax, box.ymax);
      s_half[q][side][0] = h0;
      s_half[q][side][1] = make_float4(box.zmin, box.zmax, __uint_as_float(ref), slack);
      __threadfence_block();
      const int other = atomicExch(&s_flag[q], (mine << 1) | side);
      if (other == -1) break;  // sibling not there yet (or never, in this phase: see the hand-over below)
      __threadfence_block();
      const float4 g0 = s_half[q][side ^ 1][0], g1 = s_half[q][side ^ 1][1];
      RtBox sib; sib.xmin = g0.x; sib.ymin = g0.y; sib.xmax = g0.z; sib.ymax = g0.w; sib.zmin = g1.x; sib.zmax = g1.y;
      float4* nd = nodes + 4 * (size_t)s;  // the finished node, both halves, one 64-byte line
      __stcg(nd + 2 * side, h0);
      __stcg(nd + 2 * side + 1, make_float4(box.zmin, box.zmax, __uint_as_float(ref), rt_node_aux(box, slack)));
      __stcg(nd + 2 * (side ^ 1), g0);
      __stcg(nd + 2 * (side ^ 1) + 1, make_float4(g1.x, g1.y, g1.z, rt_node_aux(sib, g1.w)));
      s_flag[q] = -2;
      slack = fmaxf(slack, g1.w);
      if (side == 0) { r = other >> 1; box = union_box(box, sib); } else { l = other >> 1; box = union_box(sib, box); }
      ref = (uint32_t)s;
    }
  }
  __syncthreads();
  if (go_global) climb_global(D, nodes, flags, n, root_out, root_box, l, r, box, ref, slack);
  const int f = s_flag[t];
  if (f >= 0) {  // a local child of split c0 + t is still waiting: its sibling reaches beyond the chunk
    const int side = f & 1, end = f >> 1, s = c0 + t;
    const float4 g0 = s_half[t][side][0], g1 = s_half[t][side][1];
    RtBox wb; wb.xmin = g0.x; wb.ymin = g0.y; wb.xmax = g0.z; wb.ymax = g0.w; wb.zmin = g1.x; wb.zmax = g1.y;
    if (side == 0) climb_global(D, nodes, flags, n, root_out, root_box, end, s, wb, __float_as_uint(g1.z), g1.w);
    else climb_global(D, nodes, flags, n, root_out, root_box, s + 1, end, wb, __float_as_uint(g1.z), g1.w);
  }
}

// ---- top-down Karras builder (reference numbering) --------------------------------------------
template <typename KeyT>
__global__ void __launch_bounds__(256) karras_topdown_kernel(const KeyT* __restrict__ keys, int n, float4* nodes, int* parent) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  Delta<KeyT> D{keys, n};
  // determine_range_adapted, lbvh_builder.cu:103-182
  const int dl = D.pair(i, i - 1), dr = D.pair(i, i + 1);
  int d, dmin;
  if (dr > dl) { d = 1; dmin = dl; } else { d = -1; dmin = dr; }
  unsigned lmax = 1;
  int nb = i + (int)lmax * d;
  int cur = D.pair(i, nb);
  while (cur > dmin) {
    lmax <<= 1;
    nb = i + (int)lmax * d;
    if (nb < 0 || nb >= n) break;
    cur = D.pair(i, nb);
  }
  unsigned len = 0;
  for (unsigned t = lmax >> 1; t > 0; t >>= 1) {
    int nbb = i + (int)(len + t) * d;
    if (nbb >= 0 && nbb < n && D.pair(i, nbb) > dmin) len += t;
  }
  const int j = i + (int)len * d;
  const int first = min(i, j), last = max(i, j);
  // find_split_adapted, lbvh_builder.cu:186-221
  int split = first;
  if (first != last) {
    const int common = D.pair(first, last);
    int step = last - first;
    do {
      step = (step + 1) >> 1;
      const int ns = split + step;
      if (ns < last && D.pair(first, ns) > common) split = ns;
    } while (step > 1);
  }
  // children, lbvh_builder.cu:264-286 (the delta re-checks there are always true)
  const bool lleaf = split == first, rleaf = split + 1 == last;
  // leaf references are completed (sphere bit, box) by the refit kernel; internal children get their slot now
  float4* h = nodes + 4 * (size_t)i;
  h[1].z = __uint_as_float(lleaf ? (RT_LEAF_BIT | (uint32_t)split) : (uint32_t)split);
  h[3].z = __uint_as_float(rleaf ? (RT_LEAF_BIT | (uint32_t)(split + 1)) : (uint32_t)(split + 1));
  // parent records: [0,n-1) internal nodes, [n-1, 2n-1) leaves; value = parent*2 + side
  if (lleaf) parent[n - 1 + split] = 2 * i; else parent[split] = 2 * i;
  if (rleaf) parent[n - 1 + split + 1] = 2 * i + 1; else parent[split + 1] = 2 * i + 1;
  if (i == 0) parent[0] = -1;
}

__global__ void __launch_bounds__(256) refit_kernel(const uint32_t* __restrict__ perm, const float4* __restrict__ bv,
                                                   const float4* __restrict__ tv, float4* __restrict__ sorted_tv, float4* nodes,
                                                   const int* __restrict__ parent, int* counters, int n, int* root_out, float* root_box) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const uint32_t p = perm[k];
  const float4 a = __ldg(bv + RT_PRIM_STRIDE * (size_t)p), b = __ldg(bv + RT_PRIM_STRIDE * (size_t)p + 1), c = __ldg(bv + RT_PRIM_STRIDE * (size_t)p + 2);
  const float4 t0 = __ldg(tv + RT_PRIM_STRIDE * (size_t)p), t1 = __ldg(tv + RT_PRIM_STRIDE * (size_t)p + 1), t2 = __ldg(tv + RT_PRIM_STRIDE * (size_t)p + 2);
  RtBox box = rt_prim_box(a, b, c);
  {
    float4* rec = sorted_tv + RT_TV_STRIDE * (size_t)k;
    rec[0] = t0; rec[1] = t1; rec[2] = t2; rec[3] = make_float4(__uint_as_float(p), 0.f, 0.f, 0.f);
    rec[4] = make_float4(box.xmin, box.ymin, box.xmax, box.ymax); rec[5] = make_float4(box.zmin, box.zmax, 0.f, 0.f);
  }
  float slack = c.y;  // hit slack of the primitive, computed with the triangle setup (scene.cu, build.cuh)
  uint32_t ref = RT_LEAF_BIT | (c.w == RT_TAG_SPHERE ? RT_SPHERE_BIT : 0u) | (uint32_t)k;
  if (n == 1) { *root_out = (int)ref; root_box[0] = box.xmin; root_box[1] = box.xmax; root_box[2] = box.ymin; root_box[3] = box.ymax; root_box[4] = box.zmin; root_box[5] = box.zmax; return; }
  int pi = parent[n - 1 + k];
  while (true) {  // set_aabb_kernel_adapted, lbvh_builder.cu:353-386, plus the missing fences (F9)
    const int par = pi >> 1, side = pi & 1;
    store_half(nodes, par, side, box, ref, slack);
    if (exch_acq_rel(counters + par, 1) == -1) return;  // counters were initialised to -1
    RtBox sib;
    slack = fmaxf(slack, load_half_box(nodes, par, side ^ 1, sib));
    box = side == 0 ? union_box(box, sib) : union_box(sib, box);
    ref = (uint32_t)par;
    pi = parent[par];
    if (pi == -1) {
      *root_out = 0;
      root_box[0] = box.xmin; root_box[1] = box.xmax; root_box[2] = box.ymin; root_box[3] = box.ymax; root_box[4] = box.zmin; root_box[5] = box.zmax;
      return;
    }
  }
}

// ---- 4-wide tree over the same leaves (RT_BUILD_WIDE4, experimental) ---------------------------------------
// Wide node i (same slot as binary node i) holds up to four 32-byte halves in two pairs: for each child of node i, the
// child's own half if it is a leaf (alone in its pair), else the two halves of the child node.  Unused places hold NaN boxes (never hit).  One thread
// per slot; a slot that no wide step can reach (odd depth) is built as well, it costs nothing to have.
__global__ void __launch_bounds__(256) collapse4_kernel(const float4* __restrict__ nodes, int n, float4* __restrict__ wnodes) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n - 1) return;
  const float nan = __int_as_float(0x7fc00000);
  float4 out[8];
#pragma unroll
  for (int k = 0; k < 8; k++) out[k] = make_float4(nan, nan, nan, nan);
  // places 0,1: the left child's children (or the left child alone, a leaf); places 2,3: the right child's.  The
  // traversal orders the pairs and the children inside a pair by entry distance (traverse.cuh).
  for (int side = 0; side < 2; side++) {
    const float4 h0 = nodes[4 * (size_t)i + 2 * side], h1 = nodes[4 * (size_t)i + 2 * side + 1];
    const uint32_t ref = __float_as_uint(h1.z);
    const int c = 2 * side;
    if (ref & RT_LEAF_BIT) { out[2 * c] = h0; out[2 * c + 1] = h1; out[2 * c + 3].z = __uint_as_float(0u); }  // ref word of the unused place
    else {
      const float4* g = nodes + 4 * (size_t)ref;
      out[2 * c] = g[0]; out[2 * c + 1] = g[1];
      out[2 * c + 2] = g[2]; out[2 * c + 3] = g[3];
    }
  }
#if RT_WIDE_BF16
  // 96-byte node: every plane as a bfloat16 offset from the node's min corner, rounded OUTWARD and checked with the very
  // addition the traversal decodes it with (traverse.cuh, rt_bf16pair_plus): fl(org + lo_off) <= lo and fl(org + hi_off) >= hi
  // as fp32 numbers, so the decoded box contains the child's exact box.  Leaves keep their exact box in the leaf record;
  // the traversal tests it before the primitive.
  float org[3] = {INFINITY, INFINITY, INFINITY};
#pragma unroll
  for (int c = 0; c < 4; c++) {
    const float4 h0 = out[2 * c], h1 = out[2 * c + 1];
    if (h0.x == h0.x) { org[0] = fminf(org[0], h0.x); org[1] = fminf(org[1], h0.y); org[2] = fminf(org[2], h1.x); }
  }
  auto off_lo = [](float lo, float o) -> unsigned {  // largest bf16 offset with fl(o + off) <= lo
    unsigned b = __float_as_uint(fmaxf(lo - o, 0.0f)) & 0xFFFF0000u;
    while (b != 0u && !(o + __uint_as_float(b) <= lo)) b -= 0x10000u;
    return b >> 16;
  };
  auto off_hi = [](float hi, float o) -> unsigned {  // smallest bf16 offset with fl(o + off) >= hi
    unsigned b = (__float_as_uint(fmaxf(hi - o, 0.0f)) + 0xFFFFu) & 0xFFFF0000u;
    while (b < 0x7F800000u && !(o + __uint_as_float(b) >= hi)) b += 0x10000u;
    return b >> 16;
  };
  unsigned w[12]; float aux[4]; unsigned refs[4];
#pragma unroll
  for (int c = 0; c < 4; c++) {
    const float4 h0 = out[2 * c], h1 = out[2 * c + 1];  // xmin ymin xmax ymax | zmin zmax ref aux
    refs[c] = __float_as_uint(h1.z); aux[c] = h1.w;
    if (h0.x == h0.x) {
      w[3 * c + 0] = off_lo(h0.x, org[0]) | (off_lo(h0.y, org[1]) << 16);
      w[3 * c + 1] = off_hi(h0.z, org[0]) | (off_hi(h0.w, org[1]) << 16);
      w[3 * c + 2] = off_lo(h1.x, org[2]) | (off_hi(h1.y, org[2]) << 16);
    } else { w[3 * c + 0] = w[3 * c + 1] = w[3 * c + 2] = 0x7FC07FC0u; refs[c] = 0u; aux[c] = 0.f; }  // unused place: NaN offsets never pass
  }
  float4* o6 = wnodes + 6 * (size_t)i;
  o6[0] = make_float4(org[0], org[1], org[2], 0.f);
  o6[1] = make_float4(__uint_as_float(refs[0]), __uint_as_float(refs[1]), __uint_as_float(refs[2]), __uint_as_float(refs[3]));
  o6[2] = make_float4(__uint_as_float(w[0]), __uint_as_float(w[1]), __uint_as_float(w[2]), __uint_as_float(w[3]));
  o6[3] = make_float4(__uint_as_float(w[4]), __uint_as_float(w[5]), __uint_as_float(w[6]), __uint_as_float(w[7]));
  o6[4] = make_float4(__uint_as_float(w[8]), __uint_as_float(w[9]), __uint_as_float(w[10]), __uint_as_float(w[11]));
  o6[5] = make_float4(aux[0], aux[1], aux[2], aux[3]);
#else
#pragma unroll
  for (int k = 0; k < 8; k++) wnodes[8 * (size_t)i + k] = out[k];
#endif
}

// ---- export in the reference's layouts --------------------------------------------------------
struct RefNode44 {  // include/lbvh.cuh:6-28
  float bbox[6]; uint32_t left, right, prim; uint16_t nleaf; uint16_t pad; uint32_t counter;
};
static_assert(sizeof(RefNode44) == 44, "LBVHNode layout");

__global__ void karras_index_kernel(const float4* nodes, int n, int root, int* karras) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n - 1) return;
  const uint32_t lr = __float_as_uint(nodes[4 * (size_t)s + 1].z), rr = __float_as_uint(nodes[4 * (size_t)s + 3].z);
  if (!(lr & RT_LEAF_BIT)) karras[lr] = s;       // left child's index = split of the parent
  if (!(rr & RT_LEAF_BIT)) karras[rr] = s + 1;   // right child's index = split + 1
  if (s == root) karras[s] = 0;
}
__global__ void export_nodes_kernel(const float4* nodes, int n, int root, const int* karras, const float* root_box, RefNode44* out) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (n == 1) {
    if (s == 0) { RefNode44 o; memset(&o, 0, sizeof(o)); for (int q = 0; q < 6; q++) o.bbox[q] = root_box[q]; o.nleaf = 1; out[0] = o; }
    return;
  }
  if (s >= n - 1) return;
  const int me = karras ? karras[s] : s;
  uint32_t child_idx[2];
  for (int side = 0; side < 2; side++) {
    const float4 u = nodes[4 * (size_t)s + 2 * side], v = nodes[4 * (size_t)s + 2 * side + 1];
    const uint32_t ref = __float_as_uint(v.z);
    uint32_t idx;
    RefNode44 o; memset(&o, 0, sizeof(o));
    if (ref & RT_LEAF_BIT) { idx = (uint32_t)(n - 1) + (ref & RT_LEAF_MASK); o.prim = ref & RT_LEAF_MASK; o.nleaf = 1; }
    else { idx = (uint32_t)(karras ? karras[ref] : (int)ref); }
    child_idx[side] = idx;
    float bb[6] = {u.x, u.z, u.y, u.w, v.x, v.y};  // xmin,xmax,ymin,ymax,zmin,zmax
    if (ref & RT_LEAF_BIT) { for (int q = 0; q < 6; q++) o.bbox[q] = bb[q]; out[idx] = o; }
    else { for (int q = 0; q < 6; q++) out[idx].bbox[q] = bb[q]; }
  }
  out[me].left = child_idx[0]; out[me].right = child_idx[1]; out[me].prim = 0; out[me].nleaf = 0; out[me].pad = 0; out[me].counter = 2;
  if (s == root) for (int q = 0; q < 6; q++) out[me].bbox[q] = root_box[q];
}
__global__ void export_refs_kernel(const int2* refs, const uint32_t* perm, int n, int2* out) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n) out[k] = refs[perm[k]];
}

}  // namespace

extern "C" int rt_build_lbvh(rt_scene* s, int morton_bits, unsigned flags, void* stream, float* ms_out) {
  if (!s) { rt_set_error("rt_build_lbvh: null scene"); return RT_ERR_INVALID; }
  if (morton_bits != 30 && morton_bits != 63) { rt_set_error("rt_build_lbvh: morton_bits must be 30 or 63"); return RT_ERR_INVALID; }
  if (s->n == 0) { rt_set_error("rt_build_lbvh: scene has no primitives"); return RT_ERR_INVALID; }  // main.cu:44 guards N>0
  RT_CUDA(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  uint32_t launches = 0;
  int rc = rt_arm_fork(s, st);
  if (rc) return rc;
  RT_CUDA(cudaEventRecord(s->ev0, st));
  rc = rt_build_codes_and_sort(s, morton_bits, flags, st, &launches);
  if (rc) return rc;
  const int n = s->n, grid = (n + 255) / 256;
    // RT_BU_LEAFONLY=1: timing experiment of scripts/build_ab.py (leaf records only, NO tree) - never set it otherwise
  const int bu_mode = (flags & RT_BUILD_GLOBAL_CLIMB) ? 0 : 1;
  const int bu_bits = (getenv("RT_BU_LEAFONLY") ? 1 : 0) | (getenv("RT_BU_NO_COMPACT") ? 0 : 2) | (getenv("RT_BU_NO_PEEK") ? 0 : 4);  // env switches: scripts/build_ab.py only
  int* root_dev = (int*)(s->bounds + 8);
  float* root_box = s->bounds + 10;
  if (flags & RT_BUILD_TOPDOWN) {
    if (n > 1) {
      if (morton_bits == 30) karras_topdown_kernel<uint32_t><<<(n - 1 + 255) / 256, 256, 0, st>>>(s->sorted_keys32, n, s->nodes, s->parent);
      else karras_topdown_kernel<uint64_t><<<(n - 1 + 255) / 256, 256, 0, st>>>(s->sorted_keys64, n, s->nodes, s->parent);
    }
    refit_kernel<<<grid, 256, 0, st>>>(s->sorted_perm, s->prim_bv, s->prim_tv, s->sorted_tv, s->nodes, s->parent, s->node_flags, n, root_dev, root_box);
  } else if (!(flags & RT_BUILD_LOCAL_CLIMB)) {
    if (morton_bits == 30)
      if (bu_mode == 0) RT_CUDA(rt_launch_chained(rt_pdl_for(n), bottom_up_kernel<uint32_t>, dim3(grid), dim3(256), 0, st, s->sorted_keys32, s->sorted_perm, s->prim_bv, s->prim_tv, s->sorted_tv, s->nodes, s->node_flags, n, root_dev, root_box));
      else RT_CUDA(rt_launch_chained(rt_pdl_for(n), bottom_up_warp_kernel<uint32_t>, dim3(grid), dim3(256), 0, st, s->sorted_keys32, s->sorted_perm, s->prim_bv, s->prim_tv, s->sorted_tv, s->nodes, s->node_flags, n, root_dev, root_box, bu_bits));
    else
      if (bu_mode == 0) RT_CUDA(rt_launch_chained(rt_pdl_for(n), bottom_up_kernel<uint64_t>, dim3(grid), dim3(256), 0, st, s->sorted_keys64, s->sorted_perm, s->prim_bv, s->prim_tv, s->sorted_tv, s->nodes, s->node_flags, n, root_dev, root_box));
      else RT_CUDA(rt_launch_chained(rt_pdl_for(n), bottom_up_warp_kernel<uint64_t>, dim3(grid), dim3(256), 0, st, s->sorted_keys64, s->sorted_perm, s->prim_bv, s->prim_tv, s->sorted_tv, s->nodes, s->node_flags, n, root_dev, root_box, bu_bits));
  } else {
    const int gl = (n + RT_BU_CHUNK - 1) / RT_BU_CHUNK;
    if (morton_bits == 30)
      bottom_up_local_kernel<uint32_t><<<gl, RT_BU_CHUNK, 0, st>>>(s->sorted_keys32, s->sorted_perm, s->prim_bv, s->prim_tv, s->sorted_tv, s->nodes, s->node_flags, n, root_dev, root_box);
    else
      bottom_up_local_kernel<uint64_t><<<gl, RT_BU_CHUNK, 0, st>>>(s->sorted_keys64, s->sorted_perm, s->prim_bv, s->prim_tv, s->sorted_tv, s->nodes, s->node_flags, n, root_dev, root_box);
  }
  s->built_wide = false;
  if ((flags & RT_BUILD_WIDE4) && n > 1) {
    collapse4_kernel<<<(n - 1 + 255) / 256, 256, 0, st>>>(s->nodes, n, s->wnodes);
    s->built_wide = true;
  }
  RT_CUDA(cudaGetLastError());
  s->built = true; s->built_bits = morton_bits; s->built_topdown = (flags & RT_BUILD_TOPDOWN) != 0;
  RT_CUDA(cudaEventRecord(s->ev1, st));
  if (ms_out) {
    RT_CUDA(cudaEventSynchronize(s->ev1));
    RT_CUDA(cudaEventElapsedTime(ms_out, s->ev0, s->ev1));
  }
  return RT_OK;
}

extern "C" int rt_scene_build_ms(rt_scene* s, float* ms_out) {
  if (!s || !ms_out) { rt_set_error("rt_scene_build_ms: null argument"); return RT_ERR_INVALID; }
  if (!s->built) { rt_set_error("rt_scene_build_ms: rt_build_lbvh has not run"); return RT_ERR_NOT_BUILT; }
  RT_CUDA(cudaSetDevice(s->device));
  RT_CUDA(cudaEventSynchronize(s->ev1));
  RT_CUDA(cudaEventElapsedTime(ms_out, s->ev0, s->ev1));
  return RT_OK;
}

extern "C" int rt_export_reference_layout(rt_scene* s, uint32_t* codes, uint64_t* codes64, void* refs8, void* nodes44, uint32_t* perm) {
  if (!s) { rt_set_error("rt_export_reference_layout: null scene"); return RT_ERR_INVALID; }
  if (!s->built) { rt_set_error("rt_export_reference_layout: rt_build_lbvh has not run"); return RT_ERR_NOT_BUILT; }
  RT_CUDA(cudaSetDevice(s->device));
  RT_CUDA(cudaDeviceSynchronize());
  s->fork_armed = false;
  const int n = s->n;
  if (codes) {
    if (!s->sorted_keys32) { rt_set_error("rt_export_reference_layout: built with 63-bit codes, pass codes64"); return RT_ERR_INVALID; }
    RT_CUDA(cudaMemcpy(codes, s->sorted_keys32, (size_t)n * 4, cudaMemcpyDeviceToHost));
  }
  if (codes64) {
    if (!s->sorted_keys64) { rt_set_error("rt_export_reference_layout: built with 30-bit codes, pass codes"); return RT_ERR_INVALID; }
    RT_CUDA(cudaMemcpy(codes64, s->sorted_keys64, (size_t)n * 8, cudaMemcpyDeviceToHost));
  }
  if (perm) RT_CUDA(cudaMemcpy(perm, s->sorted_perm, (size_t)n * 4, cudaMemcpyDeviceToHost));
  if (refs8) {
    int2* tmp; RT_CUDA(cudaMalloc(&tmp, (size_t)n * 8));
    export_refs_kernel<<<(n + 255) / 256, 256>>>(s->refs, s->sorted_perm, n, tmp);
    cudaError_t e = cudaMemcpy(refs8, tmp, (size_t)n * 8, cudaMemcpyDeviceToHost);
    cudaFree(tmp);
    RT_CUDA(e);
  }
  if (nodes44) {
    int root = 0;
    RT_CUDA(cudaMemcpy(&root, s->bounds + 8, 4, cudaMemcpyDeviceToHost));
    RefNode44* tmp; const size_t nn = (size_t)2 * n - 1;
    RT_CUDA(cudaMalloc(&tmp, nn * sizeof(RefNode44)));
    RT_CUDA(cudaMemset(tmp, 0, nn * sizeof(RefNode44)));
    const int g = (n > 1 ? n - 1 + 255 : 256) / 256;
    int* karras = nullptr;
    if (!s->built_topdown && n > 1) { karras = s->node_karras; karras_index_kernel<<<g, 256>>>(s->nodes, n, root, karras); }
    export_nodes_kernel<<<g, 256>>>(s->nodes, n, root, karras, s->bounds + 10, tmp);
    cudaError_t e = cudaMemcpy(nodes44, tmp, nn * sizeof(RefNode44), cudaMemcpyDeviceToHost);
    cudaFree(tmp);
    RT_CUDA(e);
  }
  return RT_OK;
}
