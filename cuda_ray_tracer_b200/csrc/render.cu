// render.cu — wavefront renderer: generate / traverse / shade / resolve, and rt_render / rt_trace.
//
// Replaces render(), render_kernel and render_kernel_warp_aa (src/draw.cu:94-239) and the recursive
// shading below shootPrimaryRay (src/draw.cu:260-568).  The reference runs one thread per pixel (or
// per sample) through a recursive call graph with a 64 KiB per-thread stack (main.cu:64).  Here every
// sample is a small state machine with an explicit continuation stack in HBM: each wavefront
// iteration traces exactly one ray per live sample (traverse.cuh, persistent warps over a compacted
// queue) and then advances the sample to its next hitNearest call (shade_kernel), appending the new
// ray to the next queue.  The state machine replays the reference's depth-first evaluation order, so
// the per-sample XORWOW stream is consumed in the reference's order:
//
//   shade(hit H of ray R, weight W)            = shootPrimaryRay / the "if hit" bodies, draw.cu:269-282,
//     diffuse(H)   * W(1-shin)(1-trans)          417-429, 507-518, 553-565
//     reflect(R,H) * W shin            -> F_REFLECT frame
//     refract(R,H) * W(1-shin)trans    -> F_REFRACT frame (for a reflection hit: of the OUTER ray/hit, draw.cu:424)
//     gi(H)        * W(1-shin)(1-trans) color -> F_GI frame (primary and GI hits only)
//
// Colours are accumulated with the path weight pushed down ("throughput") instead of being returned up
// the recursion; the products are the reference's, associated differently (image tolerance 1/255).
#include <algorithm>
#include <chrono>
#include <cstring>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "rng.cuh"
#include "rng_tables.h"
#include "rt_math.cuh"
#include "traverse.cuh"

namespace {

// Exact division of a 32-bit unsigned number by a divisor that is the same for a whole launch (Granlund & Montgomery's
// round-up method: q = (t + ((n - t) >> 1)) >> (L - 1), t = umulhi(m, n), L = ceil(log2 d), m = floor(2^32 (2^L - d) / d) + 1;
// exact for every n < 2^32 — checked against n / d over random and boundary values for d up to 2^32 - 1).  An integer
// division by a kernel parameter costs ~20 instructions; generate_kernel did seven per sample, one of them on 64 bits.
struct FastDiv { uint32_t d, m, sh; };
static FastDiv make_fastdiv(uint32_t d) {
  FastDiv f; f.d = d ? d : 1u;
  uint32_t L = 0; while ((1ull << L) < f.d) L++;
  f.m = (uint32_t)((((1ull << 32) * ((1ull << L) - f.d)) / f.d) + 1ull);
  f.sh = L;
  return f;
}
__device__ __forceinline__ uint32_t fdiv(const uint32_t n, const FastDiv& f) {
  if (f.sh == 0u) return n;  // d == 1 (launch-uniform)
  const uint32_t t = __umulhi(f.m, n);
  return (t + ((n - t) >> 1)) >> (f.sh - 1u);
}

// ---- per-render parameters (by value to every kernel) ------------------------------------------
struct RenderParams {
  RtCamera cam;
  int W, H;
  int mode;         // 0: aa==0 (exact pixel centre), 1: aa==1 (one jittered sample), 2: aa>=2 (one state per sample)
  int spp;          // samples per pixel actually traced
  int covered;      // pixels [0,covered) are rendered
  int tile_w, tile_h, tiles_x, num_ranks, rank, slots_per_tile, sub_w, sub_h;
  int slot0, npix, S;   // this pass: pixel slots [slot0, slot0+npix), S = npix * spp samples
  int depth;        // continuation stack depth (frames per sample)
  int ref_order /* traversal lane, RT_LANE_* */, rough_zyx, need_rng;
  int full_shade;   // mirrors, glass or GI present: the continuation stack can hold pending calls (shade_kernel<true>)
  int rng_live;     // something after generate draws random numbers (rough materials, gi): the state is stored
  int num_lights;
  FastDiv fd_spt, fd_tiles_x, fd_sub, fd_bpr, fd_sub_w, fd_spp, fd_tile_w, fd_tile_h, fd_ranks, fd_sub_h, fd_W;  // divisors of slot_to_pixel / pixel_to_slot (set_fastdivs)
};
static void set_fastdivs(RenderParams& P) {
  P.fd_spt = make_fastdiv((uint32_t)P.slots_per_tile); P.fd_tiles_x = make_fastdiv((uint32_t)P.tiles_x); P.fd_sub = make_fastdiv((uint32_t)(P.sub_w * P.sub_h));
  P.fd_bpr = make_fastdiv((uint32_t)(P.tile_w / P.sub_w)); P.fd_sub_w = make_fastdiv((uint32_t)P.sub_w); P.fd_spp = make_fastdiv((uint32_t)P.spp);
  P.fd_tile_w = make_fastdiv((uint32_t)P.tile_w); P.fd_tile_h = make_fastdiv((uint32_t)P.tile_h); P.fd_ranks = make_fastdiv((uint32_t)P.num_ranks);
  P.fd_sub_h = make_fastdiv((uint32_t)P.sub_h); P.fd_W = make_fastdiv((uint32_t)P.W);
}

struct RenderBuffers {
  uint4 *rng0, *rng1;     // XORWOW v0..v3 | v4, d, box-muller flag, cached normal
  float4* accum;          // linear rgba per sample
  uint32_t* ctl;          // stack pointer per sample
  float4* frames;         // [(level*4 + q) * S + sample]
  float4 *qa, *qb;        // ray queues, 2 float4 per ray
  float2* hits;           // per queue slot
  unsigned* fb_list;      // exact lane: rays to re-trace in reference order
  uint32_t* counters;     // [0] fetch ticket (front), [1] out-queue length (front), [2] overflow flag, [3] fallback count,
                          // [4] fallback fetch ticket, [5] out-queue length (back), [6] fetch ticket (back)
  unsigned long long* stats;  // [0] nodes, [1] leaf tests, [2] shaded hits
  const uint32_t* rng_mats;   // power matrices P^m
  const uint32_t* rng_nib;    // the same matrices as 4-bit lookup tables (RT_RNG_NIB_WORDS each)
  const uint32_t* rng_byte;   // words 0, 1, 4 of the low-seed product as 8-bit lookup tables (RT_RNG_BYTE_STRIDE each)
  const uint32_t* rng_base;   // base vectors for mode 0/1
};

enum : unsigned {
  T_W_PRIMARY = 1, T_W_GI, T_W_REFLECT, T_W_REFRACT1, T_W_REFRACT2,  // a ray is in flight; resumed with its hit
  T_DIFFUSE,                                                           // shadow ray of light `li` in flight
  T_REFLECT, T_REFRACT, T_GI                                           // pending calls
};

__device__ __forceinline__ unsigned hdr_pack(unsigned type, int bounce, unsigned flag) {
  return type | (((unsigned)bounce & 0xFFFFu) << 8) | (flag << 24);
}
__device__ __forceinline__ int hdr_bounce(unsigned h) { return (int)(short)((h >> 8) & 0xFFFFu); }

// pixel slot -> pixel.  Slots enumerate the rank's tiles (tile t belongs to rank t % num_ranks), row-major
// inside a tile; slots that fall outside the image (edge tiles) are dead.
#ifndef RT_SUB_W
#define RT_SUB_W 8
#define RT_SUB_H 4
#endif
__device__ __forceinline__ bool slot_to_pixel(const RenderParams& P, int slot, int& x, int& y, int& pix) {
  const int lt = (int)fdiv((uint32_t)slot, P.fd_spt), r = slot - lt * P.slots_per_tile;
  const int t = lt * P.num_ranks + P.rank;
  const int ty = (int)fdiv((uint32_t)t, P.fd_tiles_x), tx = t - ty * P.tiles_x;
  // inside a tile, slots run over sub_w x sub_h pixel blocks (8 x 4 = one warp) so that the 32 rays of a warp
  // cover a compact footprint and fetch the same nodes; sub = (tile_w, 1) is plain row-major
  const int sub = P.sub_w * P.sub_h, blk = (int)fdiv((uint32_t)r, P.fd_sub), in = r - blk * sub, bpr = (int)P.fd_bpr.d;
  const int by = (int)fdiv((uint32_t)blk, P.fd_bpr), bx = blk - by * bpr, iy = (int)fdiv((uint32_t)in, P.fd_sub_w);
  x = tx * P.tile_w + bx * P.sub_w + (in - iy * P.sub_w);
  y = ty * P.tile_h + by * P.sub_h + iy;
  pix = y * P.W + x;
  return x < P.W && y < P.H && pix < P.covered;
}
__device__ __forceinline__ bool pixel_to_slot(const RenderParams& P, int pix, int& slot) {
  const int y = (int)fdiv((uint32_t)pix, P.fd_W), x = pix - y * P.W;
  const int tx = (int)fdiv((uint32_t)x, P.fd_tile_w), ty = (int)fdiv((uint32_t)y, P.fd_tile_h);
  const int t = ty * P.tiles_x + tx;
  const int tq = (int)fdiv((uint32_t)t, P.fd_ranks);
  if (t - tq * P.num_ranks != P.rank) return false;
  const int lx = x - tx * P.tile_w, ly = y - ty * P.tile_h, bx = (int)fdiv((uint32_t)lx, P.fd_sub_w), by = (int)fdiv((uint32_t)ly, P.fd_sub_h);
  slot = tq * P.slots_per_tile + (by * (int)P.fd_bpr.d + bx) * (P.sub_w * P.sub_h)
         + (ly - by * P.sub_h) * P.sub_w + (lx - bx * P.sub_w);
  return true;
}

__device__ __forceinline__ void rng_load(const RenderBuffers& B, int i, Xorwow& r) {
  const uint4 a = B.rng0[i], b = B.rng1[i];
  r.v0 = a.x; r.v1 = a.y; r.v2 = a.z; r.v3 = a.w; r.v4 = b.x; r.d = b.y; r.bm_flag = b.z; r.bm_extra = __uint_as_float(b.w);
}
__device__ __forceinline__ void rng_store(const RenderBuffers& B, int i, const Xorwow& r) {
  B.rng0[i] = make_uint4(r.v0, r.v1, r.v2, r.v3);
  B.rng1[i] = make_uint4(r.v4, r.d, r.bm_flag, __float_as_uint(r.bm_extra));
}

// warp-aggregated append of one ray per participating lane (called with the whole warp converged).
// The queue has two ends: sun shadow rays (any hit, no distance limit) grow down from slot cap-1, everything else
// grows up from slot 0, so that each end gets the traversal kernel specialised for it (traverse.cuh).
__device__ __forceinline__ void queue_append(float4* __restrict__ q, uint32_t* count_front, uint32_t* count_back, unsigned cap, bool has,
                                             V3 o, V3 d, float tlimit, unsigned tag) {
  const bool back = has && (tag & RT_RAY_ANYHIT) && tlimit == INFINITY;
  const unsigned mf = __ballot_sync(0xffffffffu, has && !back), mb = __ballot_sync(0xffffffffu, back);
  if ((mf | mb) == 0u) return;
  const unsigned lane = threadIdx.x & 31, below = (1u << lane) - 1u;
  unsigned bf = 0, bb = 0;
  if (mf) {
    if (lane == (unsigned)(__ffs(mf) - 1)) bf = atomicAdd(count_front, (unsigned)__popc(mf));
    bf = __shfl_sync(0xffffffffu, bf, __ffs(mf) - 1);
  }
  if (mb) {
    if (lane == (unsigned)(__ffs(mb) - 1)) bb = atomicAdd(count_back, (unsigned)__popc(mb));
    bb = __shfl_sync(0xffffffffu, bb, __ffs(mb) - 1);
  }
  if (has) {
    const size_t r = back ? (size_t)cap - 1 - (bb + __popc(mb & below)) : (size_t)bf + __popc(mf & below);
    q[2 * r] = make_float4(o.x, o.y, o.z, tlimit);
    q[2 * r + 1] = make_float4(d.x, d.y, d.z, __uint_as_float(tag));
  }
}

// The same with ONE atomic per block and queue end: a single counter serialises in L2 at roughly a nanosecond per
// atomic, which at one atomic per warp was the whole run time of generate_kernel (56 % of its stall samples).
// Every thread of the block must call it; s_tmp holds 2 * WARPS + 2 words.
template <int WARPS>
__device__ __forceinline__ void queue_append_block(float4* __restrict__ q, uint32_t* count_front, uint32_t* count_back, unsigned cap, bool has,
                                                   V3 o, V3 d, float tlimit, unsigned tag, unsigned* s_tmp) {
  const bool back = has && (tag & RT_RAY_ANYHIT) && tlimit == INFINITY;
  const unsigned mf = __ballot_sync(0xffffffffu, has && !back), mb = __ballot_sync(0xffffffffu, back);
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5, below = (1u << lane) - 1u;
  if (lane == 0) { s_tmp[warp] = (unsigned)__popc(mf); s_tmp[WARPS + warp] = (unsigned)__popc(mb); }
  __syncthreads();
  if (threadIdx.x == 0 || threadIdx.x == 32) {
    const unsigned* c = s_tmp + (threadIdx.x ? WARPS : 0);
    unsigned tot = 0;
    for (int w = 0; w < WARPS; w++) tot += c[w];
    s_tmp[2 * WARPS + (threadIdx.x ? 1 : 0)] = tot ? atomicAdd(threadIdx.x ? count_back : count_front, tot) : 0u;
  }
  __syncthreads();
  if (has) {
    const unsigned* c = s_tmp + (back ? WARPS : 0);
    unsigned pre = s_tmp[2 * WARPS + (back ? 1 : 0)];
    for (unsigned w = 0; w < warp; w++) pre += c[w];
    const size_t r = back ? (size_t)cap - 1 - (pre + __popc(mb & below)) : (size_t)pre + __popc(mf & below);
    q[2 * r] = make_float4(o.x, o.y, o.z, tlimit);
    q[2 * r + 1] = make_float4(d.x, d.y, d.z, __uint_as_float(tag));
  }
  __syncthreads();  // s_tmp is reused by the caller's next round
}

// 4-bit lookup tables of the power matrices (rng.cuh: xw_matvec_nib), built once per table upload
__global__ void rng_nibble_kernel(const uint32_t* __restrict__ mats, int count, uint32_t* __restrict__ nib) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= count * 640) return;
  const int m = g / 640, e = g - m * 640, p = e >> 4, v = e & 15;
  const uint32_t* M = mats + (size_t)m * RT_RNG_MATRIX_WORDS;
  uint32_t r[5] = {0u, 0u, 0u, 0u, 0u};
  for (int b = 0; b < 4; b++)
    if ((v >> b) & 1)
      for (int k = 0; k < 5; k++) r[k] ^= M[8 * (4 * p + b) + k];
  for (int k = 0; k < 5; k++) nib[(size_t)m * RT_RNG_NIB_STRIDE + e * 5 + k] = r[k];
  if (e == 0) {  // M * (0, 0, v2, v3, 0) for the seed-independent words of xw_seed (seed < 2^32), rng.cuh
    uint32_t v[5], d; xw_seed(0ull, v, d);
    uint32_t in[5] = {0u, 0u, v[2], v[3], 0u}, c[5];
    xw_matvec(M, in, c);
    for (int k = 0; k < 8; k++) nib[(size_t)m * RT_RNG_NIB_STRIDE + RT_RNG_NIB_WORDS + k] = k < 5 ? c[k] : 0u;
  }
}

// 8-bit tables out of the 4-bit ones (rng.cuh: xw_matvec_byte_lowseed), built once per table upload
__global__ void rng_byte_kernel(const uint32_t* __restrict__ nib, int count, uint32_t* __restrict__ byt) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= count * 12 * 256) return;
  const int m = g / (12 * 256), e = g - m * (12 * 256), p = e >> 8, v = e & 255;
  const int word = p >> 2, b = p & 3, i = word == 2 ? 4 : word;   // byte b of state word i (0, 1 or 4)
  const uint32_t* T = nib + (size_t)m * RT_RNG_NIB_STRIDE;
  const uint32_t* lo = T + ((i * 8 + 2 * b) * 16 + (v & 15)) * 5;
  const uint32_t* hi = T + ((i * 8 + 2 * b + 1) * 16 + (v >> 4)) * 5;
  uint32_t* o = byt + (size_t)m * RT_RNG_BYTE_STRIDE + (size_t)e * 3;
  o[0] = lo[0] ^ hi[0]; o[1] = lo[1] ^ hi[1]; o[2] = lo[4] ^ hi[4];
  if (e == 0) { uint32_t* c = byt + (size_t)m * RT_RNG_BYTE_STRIDE + RT_RNG_BYTE_WORDS; c[0] = T[RT_RNG_NIB_WORDS]; c[1] = T[RT_RNG_NIB_WORDS + 1]; c[2] = T[RT_RNG_NIB_WORDS + 4]; c[3] = 0u; }
}

// ---- RNG seeding -------------------------------------------------------------------------------
// mode 0/1: curand_init(1234, pixel, 0) (draw.cu:105).  Thread g handles pixel h*256 + m with m uniform
// across the warp, so the 160 matrix rows are broadcast loads (rng.cuh).
__global__ void __launch_bounds__(256) rng_init_pixels_kernel(const RenderParams P, const RenderBuffers B, int hn) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = g / hn, h = g - m * hn;
  if (m >= 256) return;
  const int pix = h * 256 + m;
  if (pix >= P.covered) return;
  int slot;
  if (!pixel_to_slot(P, pix, slot)) return;
  if (slot < P.slot0 || slot >= P.slot0 + P.npix) return;
  uint32_t in[5], out[5];
  const uint32_t* bv = B.rng_base + 8 * (size_t)h;
#pragma unroll
  for (int k = 0; k < 5; k++) in[k] = __ldg(bv + k);
  if (m) xw_matvec(B.rng_mats + (size_t)m * RT_RNG_MATRIX_WORDS, in, out);
  else { for (int k = 0; k < 5; k++) out[k] = in[k]; }
  uint32_t v[5], d; xw_seed(1234ull, v, d);
  Xorwow r; r.v0 = out[0]; r.v1 = out[1]; r.v2 = out[2]; r.v3 = out[3]; r.v4 = out[4]; r.d = d; r.bm_flag = 0u; r.bm_extra = 0.f;
  rng_store(B, slot - P.slot0, r);
}

// ---- generate: RNG state, pixel jitter, primary ray (Ray::Ray(x,y,state,cfg), src/struct.cu:16-62) ---
#define RT_GEN_CHUNKS 8   // a block generates 8 x 256 consecutive samples: the skip-ahead table (36 KB as 8-bit, 12.8 KB as 4-bit lookups) is staged once
__global__ void __launch_bounds__(256) generate_kernel(const RenderParams P, const RenderBuffers B, float4* __restrict__ qout, int32_t* __restrict__ hit_id_out) {
  // one skip-ahead table per block: the 8-bit one (36 KB) when only the jitter draws read the state, else the 4-bit one (12.8 KB)
  __shared__ __align__(16) uint32_t s_nib[RT_RNG_BYTE_STRIDE > RT_RNG_NIB_STRIDE ? RT_RNG_BYTE_STRIDE : RT_RNG_NIB_STRIDE];
  const bool all_words = P.rng_live || P.cam.dof_focus != 0.0f;
  const long long base = (long long)blockIdx.x * (256 * RT_GEN_CHUNKS);
  // sample-major order (i = sample * npix + pixel slot): a block almost always lies inside one sample index,
  // whose skip-ahead matrix all its threads share
  const int s0 = (int)(base / P.npix);  // one 64-bit division per BLOCK (uniform), none per sample
  if (P.mode == 2 && s0 > 0) {
    const uint4* src = reinterpret_cast<const uint4*>(all_words ? B.rng_nib + (size_t)s0 * RT_RNG_NIB_STRIDE : B.rng_byte + (size_t)s0 * RT_RNG_BYTE_STRIDE);
    const int n4 = (all_words ? RT_RNG_NIB_STRIDE : RT_RNG_BYTE_STRIDE) / 4;
    for (int k = threadIdx.x; k < n4; k += 256) reinterpret_cast<uint4*>(s_nib)[k] = __ldg(src + k);
    __syncthreads();
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) B.counters[1] = P.cam.bounces != 0 ? (unsigned)P.S : 0u;  // primary rays sit at their sample's slot
  for (int chunk = 0; chunk < RT_GEN_CHUNKS; chunk++) {
  const long long g_ll = base + chunk * 256 + threadIdx.x;  // thread order: sample-major (g = sample * npix + pixel slot)
  if (base + chunk * 256 >= P.S) break;  // block-uniform
  const bool in_range = g_ll < P.S;
  bool live = in_range;
  int s = 0, lp = 0, x = 0, y = 0, pix = 0;
  if (live) {
    s = s0; long long rem = g_ll - (long long)s0 * P.npix;
    while (rem >= P.npix) { rem -= P.npix; s++; }  // a block of 1024 samples crosses a sample-index boundary at most a few times (npix >= 32)
    lp = (int)rem;
    live = slot_to_pixel(P, P.slot0 + lp, x, y, pix);
  }
  // SAMPLE ID (state index, queue slot of the primary ray, ray tag): pixel-major, i = pixel slot * spp + sample, so
  // that the 32 consecutive rays a warp of the traversal kernel fetches are the samples of ONE or two pixels — the
  // reference's own mapping (one warp = one pixel's 32 samples, draw.cu:135-213).  Such rays walk the same nodes:
  // their fetches fall into the same 128-byte lines (one L1TEX wavefront instead of 32) and they reach their leaves
  // together.  Shadow rays inherit the order through the shade kernel's appends.
  const int i = lp * P.spp + s;
  bool has = false; V3 eye = mk3(0, 0, 0), dir = mk3(0, 0, 0);
  // a sample that emits a primary ray is initialised by the first shade visit (shade_kernel<.., FIRST = true>)
  if (in_range && !(live && P.cam.bounces != 0)) {
    B.accum[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (P.full_shade) B.ctl[i] = 0u;
  }
  if (live) {
    Xorwow rng; rng.v0 = rng.v1 = rng.v2 = rng.v3 = rng.v4 = rng.d = 0u; rng.bm_flag = 0u; rng.bm_extra = 0.f;
    bool rng_dirty = false;
    float fx = (float)x, fy = (float)y;
    if (P.mode == 2) {  // curand_init(1234 + pixel_idx, lane_id, 0), draw.cu:162
      uint32_t v[5], d, out[5];
      xw_seed(1234ull + (unsigned long long)pix, v, d);
      if (s == 0) { for (int k = 0; k < 5; k++) out[k] = v[k]; }
      else if (s == s0) {  // 1234 + pix < 2^32
        if (all_words) xw_matvec_nib_lowseed<true>(s_nib, v, out);
        else xw_matvec_byte_lowseed(s_nib, v, out);  // only the two jitter draws will ever read this state: words 0, 1, 4
      }
      else xw_matvec(B.rng_mats + (size_t)s * RT_RNG_MATRIX_WORDS, v, out);  // block straddling two sample indices
      rng.v0 = out[0]; rng.v1 = out[1]; rng.v2 = out[2]; rng.v3 = out[3]; rng.v4 = out[4]; rng.d = d;
      const float jx = xw_rand(-0.5f, 0.5f, rng);  // draw.cu:165-169
      const float jy = xw_rand(-0.5f, 0.5f, rng);
      fx = (float)x + jx; fy = (float)y + jy;
      rng_dirty = true;
    } else if (P.need_rng) {
      rng_load(B, i, rng);
      if (P.mode == 1) {  // draw.cu:118-123
        fx = x + xw_rand(-0.5f, 0.5f, rng);
        fy = y + xw_rand(-0.5f, 0.5f, rng);
        rng_dirty = true;
      }
    }
    // struct.cu:16-62
    const RtCamera& C = P.cam;
    const V3 fwd = mk3(C.forward[0], C.forward[1], C.forward[2]), right = mk3(C.right[0], C.right[1], C.right[2]),
             up = mk3(C.up[0], C.up[1], C.up[2]), ceye = mk3(C.eye[0], C.eye[1], C.eye[2]);
    const float max_dim = fmax((float)C.width, (float)C.height);
    float sx = (2.0f * fx - C.width) / max_dim;
    float sy = (C.height - 2.0f * fy) / max_dim;
    eye = ceye;
    if (C.fisheye) {
      dir = (sx * right + sy * up) + sqrt(1.0f - pow(sx, 2) - pow(sy, 2)) * fwd;
    } else if (C.panorama) {
      sx = (float)fx / (float)C.width;
      sy = (float)fy / (float)C.height;
      const float theta = (sx - 0.5f) * 2.0f * 3.14159265358979323846f;
      const float phi = (sy - 0.5f) * 3.14159265358979323846f;
      dir = cos(phi) * (cos(theta) * fwd + sin(theta) * right) - sin(phi) * up;
      dir = normalize3(dir);
    } else if (C.dof_focus != 0.0f) {
      const float theta = xw_rand(0.0f, 2.0f * 3.14159265358979323846f, rng);
      const float r = xw_rand(0.0f, C.dof_lens, rng);
      rng_dirty = true;
      const float lx = r * cos(theta);
      const float ly = r * sin(theta);
      eye = eye + lx * up + ly * right;
      const V3 old_dir = fwd + sx * right + sy * up;
      dir = (ceye + normalize3(old_dir) * C.dof_focus - eye) / C.dof_focus;
    } else {
      dir = fwd + sx * right + sy * up;
    }
    dir = normalize3(dir);
    if (rng_dirty && P.rng_live) rng_store(B, i, rng);
    if (C.bounces != 0) {  // hitNearest: a ray with bounce == 0 hits nothing (draw.cu:294)
      has = true;
    } else if (hit_id_out && s == 0) hit_id_out[pix] = -1;
  }
  // no compaction: the ray of sample i goes to queue slot i (a sample without a ray leaves a record the traversal and
  // shade kernels skip); the live rays are counted for the accounting, one atomic per block
  if (in_range && P.cam.bounces != 0) {
    qout[2 * (size_t)i] = make_float4(eye.x, eye.y, eye.z, INFINITY);
    qout[2 * (size_t)i + 1] = make_float4(dir.x, dir.y, dir.z, __uint_as_float(has ? (unsigned)i : RT_RAY_DEAD));
  }
  const int nlive = __syncthreads_count(has);
  if (nlive && threadIdx.x == 0) atomicAdd(B.counters + 7, (unsigned)nlive);
  }
}

// ---- shade -------------------------------------------------------------------------------------
struct SurfHit { V3 point, normal, color, shin, trans; float ior, rough; int file_id; };

__device__ __forceinline__ void load_material(const float4* __restrict__ m, SurfHit& h) {
  const float4 a = __ldg(m), b = __ldg(m + 1), c = __ldg(m + 2);
  h.color = mk3(a.x, a.y, a.z); h.shin = mk3(a.w, b.x, b.y); h.trans = mk3(b.z, b.w, c.x); h.ior = c.y; h.rough = c.z;
}
// rebuild the ObjectInfo the reference's intersection routines return (point, facing normal, material)
__device__ __forceinline__ void reconstruct_hit(const RtDeviceScene& sc, int id, float t, V3 o, V3 d, SurfHit& h) {
  h.point = t * d + o;
  if (id >= 0) {
    const unsigned k = (unsigned)id & RT_LEAF_MASK;
    const float4 a = __ldg(sc.sorted_tv + RT_TV_STRIDE * (size_t)k);
    if ((unsigned)id & RT_SPHERE_BIT) {  // struct.cu:76-108
      const V3 c = mk3(a.x, a.y, a.z);
      const V3 cr0 = c - o;
      const bool inside = dot(cr0, cr0) < a.w * a.w;
      h.normal = normalize3(inside ? (c - h.point) : (h.point - c));
    } else {  // struct.cu:124,160
      const float4 b = __ldg(sc.sorted_tv + RT_TV_STRIDE * (size_t)k + 1);
      const V3 nor = mk3(a.w, b.x, b.y);
      h.normal = (dot(d, nor) < 0.0f) ? nor : -nor;
    }
    h.file_id = (int)__ldg(sc.perm + k);
    load_material(sc.prim_mat + 3 * (size_t)h.file_id, h);
  } else {  // plane -2-p, draw.cu:597-611
    const int p = -2 - id;
    const float* pl = reinterpret_cast<const float*>(sc.planes + p);
    const V3 nor = mk3(__ldg(pl), __ldg(pl + 1), __ldg(pl + 2));
    h.normal = (dot(nor, d) < 0.0f) ? nor : -nor;
    h.color = mk3(__ldg(pl + 6), __ldg(pl + 7), __ldg(pl + 8)); h.shin = mk3(__ldg(pl + 9), __ldg(pl + 10), __ldg(pl + 11));
    h.trans = mk3(__ldg(pl + 12), __ldg(pl + 13), __ldg(pl + 14)); h.ior = __ldg(pl + 15); h.rough = __ldg(pl + 16);
    h.file_id = id;
  }
}

// setExpose, helper.cu:40-45: the subtraction is done in double
__device__ __forceinline__ float set_expose(float c, float expose) { return expose == float(INFINITY) ? c : 1.0 - exp(-expose * c); }

// vec3(standerdD, standerdD, standerdD), draw.cu:335-337,395-397: the three draws are unsequenced
// constructor arguments; zyx selects right-to-left evaluation
__device__ __forceinline__ V3 rough_offset(float rough, Xorwow& rng, int zyx) {
  float a, b, c;
  if (zyx) { c = xw_normal(rng) * rough; b = xw_normal(rng) * rough; a = xw_normal(rng) * rough; }
  else { a = xw_normal(rng) * rough; b = xw_normal(rng) * rough; c = xw_normal(rng) * rough; }
  return mk3(a, b, c);
}

// The kernel waits on three dependent random loads per ray (hit -> leaf record -> material), so what it needs is
// many warps in flight: 64 registers (8 CTAs of 128 threads per SM; 118 registers and 17 % occupancy without the
// bound), and a FULL = false instantiation for scenes without mirrors, glass or GI, where the continuation stack
// never holds a pending call and half of the state machine is dead code.
#ifndef RT_SHADE_MIN_CTAS
#define RT_SHADE_MIN_CTAS 8
#endif
// FIRST = the iteration right after generate_kernel: every queue entry is a primary ray whose frame is the constant
// (T_W_PRIMARY, bounces, weight 1) and whose accumulator is zero, so neither is stored by generate nor loaded here.
template <bool FULL, bool FIRST>
__global__ void __launch_bounds__(128, RT_SHADE_MIN_CTAS) shade_kernel(const RenderParams P, const RtDeviceScene sc, const RenderBuffers B,
                                                   const float4* __restrict__ qin, const unsigned n_front_in, const unsigned n_back_in,
                                                   float4* __restrict__ qout, int32_t* __restrict__ hit_id_out,
                                                   const unsigned* __restrict__ counts_dev) {
  // counts_dev != nullptr: a pass queued without host round trips — the ray counts are read here, the grid is an upper
  // bound (or smaller: the blocks stride over the queue), B.counters is the NEXT iteration's counter block
  const unsigned n_front = counts_dev ? counts_dev[1] : n_front_in, n_back = counts_dev ? counts_dev[5] : n_back_in;
  __shared__ unsigned s_app[2 * 4 + 2];
  for (unsigned g0 = blockIdx.x * blockDim.x; g0 < n_front + n_back; g0 += gridDim.x * blockDim.x) {  // block-uniform
  const unsigned g = g0 + threadIdx.x;
  bool emit = false, shaded = false; V3 eo = mk3(0, 0, 0), ed = mk3(0, 0, 0); float elimit = INFINITY; unsigned etag = 0u;
  const unsigned r = g < n_front ? g : (unsigned)P.S - n_back + (g - n_front);  // queue slot: front end, then back end
  float4 q0 = make_float4(0.f, 0.f, 0.f, 0.f), q1 = make_float4(0.f, 0.f, 0.f, __uint_as_float(RT_RAY_DEAD));
  if (g < n_front + n_back) { q0 = __ldcs(qin + 2 * (size_t)r); q1 = __ldcs(qin + 2 * (size_t)r + 1); }
  // RT_RAY_DEAD: a sample without a primary ray (generate_kernel leaves such a record at the sample's slot)
  if (__float_as_uint(q1.w) != RT_RAY_DEAD) {
    const float2 hv = __ldcs(B.hits + r);
    const V3 o = mk3(q0.x, q0.y, q0.z), d = mk3(q1.x, q1.y, q1.z);
    const int i = (int)(__float_as_uint(q1.w) & ~RT_RAY_ANYHIT);
    const float ht = hv.x; const int hid = __float_as_int(hv.y);
    const size_t S = (size_t)P.S;
    float4* const F = B.frames + i;  // frame slot (L,q) is F[(L*4+q)*S]
    int sp = (FULL && !FIRST) ? (int)B.ctl[i] : 1;  // without mirrors, glass or GI exactly one frame is live while a ray is in flight
    Xorwow rng; bool rng_loaded = false, rng_dirty = false;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f); bool acc_loaded = FIRST, acc_dirty = FIRST;  // FIRST: this visit writes the initial value
    bool overflow = false;
    const int nsuns = sc.num_suns, nlights = sc.num_suns + sc.num_bulbs;
    const float expose = P.cam.expose;

#define FR(L, q) F[(size_t)((L) * 4 + (q)) * S]
#define NEED_RNG() do { if (!rng_loaded) { rng_load(B, i, rng); rng_loaded = true; } rng_dirty = true; } while (0)
#define NEED_ACC() do { if (!acc_loaded) { acc = B.accum[i]; acc_loaded = true; } acc_dirty = true; } while (0)

    // state of the diffuse light loop (draw.cu:342-374)
    bool in_diffuse = false; int li = 0;
    V3 df_point = mk3(0, 0, 0), df_origin = mk3(0, 0, 0), df_nrm = mk3(0, 0, 0), df_color = mk3(0, 0, 0), df_w = mk3(0, 0, 0);
    bool df_frame_valid = false;  // frame slots q1..q3 already hold this diffuse state

    // ---- 1. resume the frame whose ray has just been traced ----
    sp--;
    const float4 f0 = FIRST ? make_float4(__uint_as_float(hdr_pack(T_W_PRIMARY, P.cam.bounces, 0u)), 1.f, 1.f, 1.f) : FR(sp, 0);
    const unsigned hdr = __float_as_uint(f0.x);
    const unsigned type = hdr & 0xFFu;
    const int b = hdr_bounce(hdr);  // bounce of the traced ray
    const V3 Wt = mk3(f0.y, f0.z, f0.w);

    if (!FIRST && type == T_DIFFUSE) {
      const float4 f1 = FR(sp, 1), f2 = FR(sp, 2), f3 = FR(sp, 3);
      li = (int)(hdr >> 8);
      df_point = mk3(f0.y, f0.z, f0.w); df_origin = mk3(f1.x, f1.y, f1.z); df_nrm = mk3(f1.w, f2.x, f2.y);
      df_color = mk3(f2.z, f2.w, f3.x); df_w = mk3(f3.y, f3.z, f3.w);
      in_diffuse = true; df_frame_valid = true;
      // shadow test result, draw.cu:350-353 (sun: any hit) / :365-368 (bulb: hit closer than the bulb)
      V3 ldir; float lt = 0.f; V3 lcol;
      bool occluded;
      if (li < nsuns) {
        const LightSun& L = sc.suns[li];
        ldir = mk3(L.dir[0], L.dir[1], L.dir[2]); lcol = mk3(L.color[0], L.color[1], L.color[2]);
        occluded = hid != -1;
      } else {
        const LightBulb& L = sc.bulbs[li - nsuns];
        ldir = mk3(L.point[0], L.point[1], L.point[2]) - df_point; lcol = mk3(L.color[0], L.color[1], L.color[2]);
        lt = length3(ldir);
        occluded = hid != -1 && ht < lt;
      }
      if (!occluded) {
        const float lambert = fmax(dot(df_nrm, normalize3(ldir)), 0.0f);
        const float cr = df_color.x * (lcol.x * lambert), cg = df_color.y * (lcol.y * lambert), cb = df_color.z * (lcol.z * lambert);
        float er = set_expose(cr, expose), eg = set_expose(cg, expose), eb = set_expose(cb, expose);
        if (li >= nsuns) {  // getColorBulb, draw.cu:649-659
          const float fall = 1.0f / pow(lt, 2);
          er = er * fall; eg = eg * fall; eb = eb * fall;
        }
        NEED_ACC();
        acc.x += df_w.x * er; acc.y += df_w.y * eg; acc.z += df_w.z * eb;
      }
      li++;
    } else if (FULL && !FIRST && type == T_W_REFRACT1) {
      // refractionLight after the inside ray, draw.cu:484-495 (a miss leaves the default ObjectInfo)
      V3 normal = mk3(0.f, 0.f, 0.f), ip = mk3(0.f, 0.f, 0.f); float ior = 1.458f;
      if (hid != -1) {
        SurfHit oh; reconstruct_hit(sc, hid, ht, o, d, oh);
        normal = normalize3(oh.normal); ior = oh.ior; ip = oh.point;
      }
      const float k = 1.0f - ior * ior * (1.0f - pow(dot(normal, d), 2));
      const V3 rdir = ior * d - (ior * (dot(normal, d)) + sqrt(k)) * normal;
      const int nb = b - 1;
      if (nb != 0) {
        FR(sp, 0) = make_float4(__uint_as_float(hdr_pack(T_W_REFRACT2, nb, 0u)), Wt.x, Wt.y, Wt.z); sp++;
        emit = true; eo = ip - normal * 0.0001f; ed = normalize3(rdir); elimit = INFINITY; etag = (unsigned)i;
      }
    } else {
      // T_W_PRIMARY / T_W_GI / T_W_REFLECT / T_W_REFRACT2: the "if hit" body
      const unsigned flag = hdr >> 24;
      if (type == T_W_PRIMARY && hit_id_out) {
        const int lp = (int)fdiv((uint32_t)i, P.fd_spp);  // sample ids are pixel-major (generate_kernel)
        if (i - lp * P.spp == 0) {
          int x, y, pix; slot_to_pixel(P, P.slot0 + lp, x, y, pix);
          int out_id = hid;
          if (hid >= 0) out_id = (int)__ldg(sc.perm + ((unsigned)hid & RT_LEAF_MASK));
          hit_id_out[pix] = out_id;
        }
      }
      if (hid == -1) {
        if (FULL && type == T_W_REFLECT && flag) sp--;  // the provisional refract(outer) frame is never evaluated
      } else {
        SurfHit h; reconstruct_hit(sc, hid, ht, o, d, h);
        if (type == T_W_PRIMARY) { NEED_ACC(); acc.w = 1.0f; }  // color.a = 1, draw.cu:281
        const V3 one = mk3(1.0f, 1.0f, 1.0f);
        const V3 ms = one - h.shin;
        const V3 w_diff = mul3(Wt, mul3(ms, one - h.trans));
        const V3 w_refl = mul3(Wt, h.shin);
        const V3 w_refr = mul3(Wt, mul3(ms, h.trans));
        const bool shiny = !rgb_is_zero(h.shin), transp = !rgb_is_zero(h.trans);
        // frames are pushed in reverse evaluation order: gi, refract, reflect, then diffuse runs now
        if (FULL && (type == T_W_PRIMARY || type == T_W_GI) && P.cam.gi != 0) {
          const int g = type == T_W_PRIMARY ? P.cam.gi : b;
          if (sp < P.depth) {
            const V3 wg = mul3(w_diff, h.color);
            FR(sp, 0) = make_float4(__uint_as_float(hdr_pack(T_GI, g, 0u)), wg.x, wg.y, wg.z);
            FR(sp, 1) = make_float4(h.point.x, h.point.y, h.point.z, h.normal.x);
            FR(sp, 2) = make_float4(h.normal.y, h.normal.z, 0.f, 0.f);
            sp++;
          } else overflow = true;
        }
        if (FULL && type == T_W_REFLECT) {
          if (flag) { const float4 t0 = FR(sp - 1, 0); FR(sp - 1, 0) = make_float4(t0.x, w_refr.x, w_refr.y, w_refr.z); }
        } else if (FULL && transp && b > 0) {
          if (sp < P.depth) {
            FR(sp, 0) = make_float4(__uint_as_float(hdr_pack(T_REFRACT, b, 0u)), w_refr.x, w_refr.y, w_refr.z);
            FR(sp, 1) = make_float4(d.x, d.y, d.z, h.point.x);
            FR(sp, 2) = make_float4(h.point.y, h.point.z, h.normal.x, h.normal.y);
            FR(sp, 3) = make_float4(h.normal.z, h.ior, 0.f, 0.f);
            sp++;
          } else overflow = true;
        }
        if (FULL && shiny && b > 0) {
          if (sp < P.depth) {
            FR(sp, 0) = make_float4(__uint_as_float(hdr_pack(T_REFLECT, b, transp ? 1u : 0u)), w_refl.x, w_refl.y, w_refl.z);
            FR(sp, 1) = make_float4(d.x, d.y, d.z, h.point.x);
            FR(sp, 2) = make_float4(h.point.y, h.point.z, h.normal.x, h.normal.y);
            FR(sp, 3) = make_float4(h.normal.z, h.ior, h.rough, 0.f);
            sp++;
          } else overflow = true;
        }
        // diffuseLight, draw.cu:329-377
        shaded = true;
        V3 nrm = h.normal;
        if (h.rough > 0.0f) { NEED_RNG(); nrm = nrm + rough_offset(h.rough, rng, P.rough_zyx); }
        df_nrm = normalize3(nrm);
        df_point = h.point; df_origin = h.point + h.normal * 0.001f; df_color = h.color; df_w = w_diff;
        in_diffuse = true; li = 0; df_frame_valid = false;
      }
    }

    // ---- 2. run until a ray is emitted or the stack is empty ----
    while (!emit) {
      if (in_diffuse) {
        if (li < nlights) {
          V3 ldir; float lim = INFINITY;
          if (li < nsuns) { const LightSun& L = sc.suns[li]; ldir = mk3(L.dir[0], L.dir[1], L.dir[2]); }
          else {
            const LightBulb& L = sc.bulbs[li - nsuns];
            ldir = mk3(L.point[0], L.point[1], L.point[2]) - df_point;
            lim = length3(ldir);
          }
          if (sp < P.depth) {
            FR(sp, 0) = make_float4(__uint_as_float(T_DIFFUSE | ((unsigned)li << 8)), df_point.x, df_point.y, df_point.z);
            if (!df_frame_valid) {
              FR(sp, 1) = make_float4(df_origin.x, df_origin.y, df_origin.z, df_nrm.x);
              FR(sp, 2) = make_float4(df_nrm.y, df_nrm.z, df_color.x, df_color.y);
              FR(sp, 3) = make_float4(df_color.z, df_w.x, df_w.y, df_w.z);
            }
            sp++;
            emit = true; eo = df_origin; ed = normalize3(ldir);  // Ray(eye, dir, 1) normalises, object.cuh:69
            elimit = lim; etag = (unsigned)i | RT_RAY_ANYHIT;
            break;
          }
          overflow = true; li = nlights;
        }
        in_diffuse = false;
        continue;
      }
      if (!FULL || sp == 0) break;  // sample finished (without mirrors, glass or GI nothing is ever pending)
      sp--;
      const float4 g0 = FR(sp, 0);
      const unsigned gh = __float_as_uint(g0.x);
      const unsigned gt = gh & 0xFFu;
      const int gb = hdr_bounce(gh);
      const V3 W = mk3(g0.y, g0.z, g0.w);
      if (gt == T_REFLECT) {  // reflectionLight, draw.cu:389-437
        const float4 g1 = FR(sp, 1), g2 = FR(sp, 2), g3 = FR(sp, 3);
        const V3 rd = mk3(g1.x, g1.y, g1.z), hp = mk3(g1.w, g2.x, g2.y), hn = mk3(g2.z, g2.w, g3.x);
        const float hior = g3.y, hrough = g3.z;
        const bool transp = (gh >> 24) != 0u;
        V3 normal = hn;
        if (hrough > 0.0f) { NEED_RNG(); normal = normal + rough_offset(hrough, rng, P.rough_zyx); }
        normal = normalize3(normal);
        const V3 rdir = rd - 2.0f * (dot(normal, rd)) * normal;
        const int nb = gb - 1;
        if (nb == 0) continue;  // second ray cannot hit: colour (0,0,0)
        unsigned flag = 0u;
        if (transp) {  // refractionLight(ray, obj) of the OUTER pair (draw.cu:424), weight patched at resume
          if (sp + 1 < P.depth) {
            FR(sp, 0) = make_float4(__uint_as_float(hdr_pack(T_REFRACT, gb, 0u)), 0.f, 0.f, 0.f);
            FR(sp, 1) = make_float4(rd.x, rd.y, rd.z, hp.x);
            FR(sp, 2) = make_float4(hp.y, hp.z, hn.x, hn.y);
            FR(sp, 3) = make_float4(hn.z, hior, 0.f, 0.f);
            sp++; flag = 1u;
          } else overflow = true;
        }
        if (sp < P.depth) {
          FR(sp, 0) = make_float4(__uint_as_float(hdr_pack(T_W_REFLECT, nb, flag)), W.x, W.y, W.z); sp++;
          emit = true; eo = hp + hn * 0.001f; ed = normalize3(rdir); elimit = INFINITY; etag = (unsigned)i;
        } else { overflow = true; if (flag) sp--; }
      } else if (gt == T_REFRACT) {  // refractionLight, draw.cu:456-482
        const float4 g1 = FR(sp, 1), g2 = FR(sp, 2), g3 = FR(sp, 3);
        const V3 dir = mk3(g1.x, g1.y, g1.z), ip = mk3(g1.w, g2.x, g2.y);
        const V3 normal = normalize3(mk3(g2.z, g2.w, g3.x));
        const float ior = 1.0f / g3.y;
        int bounce = gb;
        const float k = 1.0f - pow(ior, 2) * (1.0f - pow(dot(normal, dir), 2));
        if (k < 0) {  // total internal reflection
          const V3 rdir = dir - 2.0f * (dot(normal, dir)) * normal;
          --bounce;
          if (bounce == 0) continue;
          if (sp < P.depth) {
            FR(sp, 0) = make_float4(__uint_as_float(hdr_pack(T_W_REFRACT2, bounce, 0u)), W.x, W.y, W.z); sp++;
            emit = true; eo = ip + normal * 0.001f; ed = normalize3(rdir); elimit = INFINITY; etag = (unsigned)i;
          } else overflow = true;
        } else {
          const V3 rdir = ior * dir - (ior * (dot(normal, dir)) + sqrt(k)) * normal;
          if (sp < P.depth) {
            FR(sp, 0) = make_float4(__uint_as_float(hdr_pack(T_W_REFRACT1, bounce, 0u)), W.x, W.y, W.z); sp++;
            emit = true; eo = ip - normal * 0.0001f; ed = normalize3(rdir); elimit = INFINITY; etag = (unsigned)i;
          } else overflow = true;
        }
      } else if (gt == T_GI) {  // globalIllumination, draw.cu:540-551
        const float4 g1 = FR(sp, 1), g2 = FR(sp, 2);
        const V3 ip = mk3(g1.x, g1.y, g1.z), normal = mk3(g1.w, g2.x, g2.y);
        NEED_RNG();
        // spherePoint, helper.cu:91-101
        const float z = 2.0f * xw_rand(0.0f, 1.0f, rng) - 1.0f;
        const float theta = 2.0f * 3.14159265f * xw_rand(0.0f, 1.0f, rng);
        const float rr = sqrt(1.0f - z * z);
        const V3 spt = mk3(rr * cos(theta), rr * sin(theta), z);
        const V3 gdir = normalize3(normal + spt);
        const int nb = gb - 1;
        if (nb == 0) continue;
        if (sp < P.depth) {
          FR(sp, 0) = make_float4(__uint_as_float(hdr_pack(T_W_GI, nb, 0u)), W.x, W.y, W.z); sp++;
          emit = true; eo = ip + normal * 0.001f; ed = normalize3(gdir); elimit = INFINITY; etag = (unsigned)i;
        } else overflow = true;
      }
    }
    if (FULL) B.ctl[i] = (unsigned)sp;
    if (rng_dirty) rng_store(B, i, rng);
    if (acc_dirty) B.accum[i] = acc;
    if (overflow) B.counters[2] = 1u;
#undef FR
#undef NEED_RNG
#undef NEED_ACC
  }
  queue_append_block<4>(qout, B.counters + 1, B.counters + 5, (unsigned)P.S, emit, eo, ed, elimit, etag, s_app);
  const int nshaded = __syncthreads_count(shaded);  // one atomic per block: a single hot counter serialises in L2
  if (nshaded && threadIdx.x == 0) atomicAdd(B.stats + 2, (unsigned long long)nshaded);
  }
}

// ---- shade, inline-sun passes ---------------------------------------------------------------------------------------
// Scenes without mirrors, glass, GI or rough surfaces, lit by one sun: a sample is its primary ray and, if that hits, one
// shadow ray, which traverse_near_kernel traces inline (hits2 / `occl`).  One visit per sample then does what
// shade_kernel<false, true> and shade_kernel<false, false> do in two, with the same expressions in the same order
// (shootPrimaryRay draw.cu:269-282, diffuseLight :329-353, getColorSun :627-647): no continuation frame, no second
// queue.  The number of shadow rays (= hits) is added to counters[5], where the accounting expects the next iteration's rays.
__global__ void __launch_bounds__(128, RT_SHADE_MIN_CTAS) shade_inline_kernel(const RenderParams P, const RtDeviceScene sc, const RenderBuffers B,
                                                                             const float4* __restrict__ qin, const unsigned* __restrict__ occl,
                                                                             int32_t* __restrict__ hit_id_out, const unsigned* __restrict__ counts_dev) {
  const unsigned n = counts_dev[1];
  for (unsigned g0 = blockIdx.x * blockDim.x; g0 < n; g0 += gridDim.x * blockDim.x) {  // block-uniform
    const unsigned r = g0 + threadIdx.x;
    bool shaded = false;
    float4 q1 = make_float4(0.f, 0.f, 0.f, __uint_as_float(RT_RAY_DEAD));
    if (r < n) q1 = __ldcs(qin + 2 * (size_t)r + 1);
    if (__float_as_uint(q1.w) != RT_RAY_DEAD) {
      const float4 q0 = __ldcs(qin + 2 * (size_t)r);
      const float2 hv = __ldcs(B.hits + r);
      const V3 o = mk3(q0.x, q0.y, q0.z), d = mk3(q1.x, q1.y, q1.z);
      const int i = (int)__float_as_uint(q1.w);
      const float ht = hv.x; const int hid = __float_as_int(hv.y);
      if (hit_id_out) {
        const int lp = (int)fdiv((uint32_t)i, P.fd_spp);
        if (i - lp * P.spp == 0) {
          int x, y, pix; slot_to_pixel(P, P.slot0 + lp, x, y, pix);
          int out_id = hid;
          if (hid >= 0) out_id = (int)__ldg(sc.perm + ((unsigned)hid & RT_LEAF_MASK));
          hit_id_out[pix] = out_id;
        }
      }
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      if (hid != -1) {
        SurfHit h; reconstruct_hit(sc, hid, ht, o, d, h);
        acc.w = 1.0f;  // color.a = 1, draw.cu:281
        shaded = true;
        if (__ldcs(occl + r) == 0u) {
          const V3 one = mk3(1.0f, 1.0f, 1.0f);
          const V3 Wt = mk3(1.f, 1.f, 1.f);
          const V3 ms = one - h.shin;
          const V3 df_w = mul3(Wt, mul3(ms, one - h.trans));
          const V3 df_nrm = normalize3(h.normal);
          const LightSun& L = sc.suns[0];
          const V3 ldir = mk3(L.dir[0], L.dir[1], L.dir[2]), lcol = mk3(L.color[0], L.color[1], L.color[2]);
          const float lambert = fmax(dot(df_nrm, normalize3(ldir)), 0.0f);
          const float cr = h.color.x * (lcol.x * lambert), cg = h.color.y * (lcol.y * lambert), cb = h.color.z * (lcol.z * lambert);
          const float er = set_expose(cr, P.cam.expose), eg = set_expose(cg, P.cam.expose), eb = set_expose(cb, P.cam.expose);
          acc.x += df_w.x * er; acc.y += df_w.y * eg; acc.z += df_w.z * eb;
        }
      }
      B.accum[i] = acc;
    }
    const int nshaded = __syncthreads_count(shaded);  // one atomic per block: a single hot counter serialises in L2
    if (nshaded && threadIdx.x == 0) { atomicAdd(B.stats + 2, (unsigned long long)nshaded); atomicAdd(B.counters + 5, (unsigned)nshaded); }
  }
}

// ---- resolve: sample sum in the butterfly order of render_kernel_warp_aa, sRGB, 8-bit store -----
__device__ __forceinline__ float rgb_to_srgb(float l) {  // RGBtosRGB, helper.cu:12-27
  float sol;
  if (l < 0.0031308f) sol = 12.92f * l;
  else sol = 1.055f * pow(l, 1 / 2.4f) - 0.055f;
  sol = fmin(1.0f, fmax(0.0f, sol));
  return sol;
}
__device__ __forceinline__ unsigned char to_u8_round(float f) {  // floatToUcharWarp, draw.cu:9-11
  return (unsigned char)(fminf(fmaxf(f, 0.0f), 1.0f) * 255.0f + 0.5f);
}
__device__ __forceinline__ float4 add4(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }

// v[i] += v[i ^ m] for m = NP2/2 .. 1 leaves in v[0] the halving-tree sum: lane 0 of the shuffle butterfly
// (draw.cu:181-189).  Samples >= spp are zero.
template <int NP2>
__device__ __forceinline__ float4 tree_sum(const float4* __restrict__ a, size_t stride, int spp) {
  float4 v[NP2 / 2 > 0 ? NP2 / 2 : 1];
  if (NP2 == 1) return a[0];
#pragma unroll
  for (int k = 0; k < NP2 / 2; k++) {
    const float4 lo = k < spp ? a[(size_t)k * stride] : make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 hi = k + NP2 / 2 < spp ? a[(size_t)(k + NP2 / 2) * stride] : make_float4(0.f, 0.f, 0.f, 0.f);
    v[k] = add4(lo, hi);
  }
#pragma unroll
  for (int half = NP2 / 4; half >= 1; half >>= 1) {
#pragma unroll
    for (int k = 0; k < half; k++) v[k] = add4(v[k], v[k + half]);
  }
  return v[0];
}

__global__ void __launch_bounds__(128) resolve_kernel(const RenderParams P, const RenderBuffers B, uint8_t* __restrict__ rgba,
                                                     uint8_t* __restrict__ written) {
  const int lp = blockIdx.x * blockDim.x + threadIdx.x;
  if (lp >= P.npix) return;
  int x, y, pix;
  if (!slot_to_pixel(P, P.slot0 + lp, x, y, pix)) return;
  const float4* a = B.accum + (size_t)lp * P.spp;  // the pixel's samples are consecutive (sample id = pixel slot * spp + sample)
  uchar4 out;
  if (P.mode != 2) {  // render_kernel store, draw.cu:129-132 (float -> unsigned char truncates)
    const float4 c = a[0];
    out.x = (unsigned char)(rgb_to_srgb(c.x) * 255); out.y = (unsigned char)(rgb_to_srgb(c.y) * 255);
    out.z = (unsigned char)(rgb_to_srgb(c.z) * 255); out.w = (unsigned char)(c.w * 255);
  } else {
    int np2 = 1; while (np2 < P.spp) np2 <<= 1;
    float4 v;
    const size_t st = 1;
    if (np2 <= 2) v = tree_sum<2>(a, st, P.spp);
    else if (np2 == 4) v = tree_sum<4>(a, st, P.spp);
    else if (np2 == 8) v = tree_sum<8>(a, st, P.spp);
    else if (np2 == 16) v = tree_sum<16>(a, st, P.spp);
    else if (np2 == 32) v = tree_sum<32>(a, st, P.spp);
    else if (np2 == 64) v = tree_sum<64>(a, st, P.spp);
    else {  // in-place halving in global memory (the accumulators are dead after this kernel)
      float4* w = B.accum + (size_t)lp * P.spp;
      for (int half = np2 / 2; half >= 1; half >>= 1)
        for (int k = 0; k < half; k++) {
          const float4 lo = k < P.spp ? w[(size_t)k * st] : make_float4(0.f, 0.f, 0.f, 0.f);
          const float4 hi = k + half < P.spp ? w[(size_t)(k + half) * st] : make_float4(0.f, 0.f, 0.f, 0.f);
          if (k < P.spp) w[(size_t)k * st] = add4(lo, hi);
        }
      v = w[0];
    }
    const float inv = 1.0f / (float)P.spp;  // draw.cu:195 (1.0f/32.0f)
    out.x = to_u8_round(rgb_to_srgb(v.x * inv)); out.y = to_u8_round(rgb_to_srgb(v.y * inv));
    out.z = to_u8_round(rgb_to_srgb(v.z * inv)); out.w = to_u8_round(v.w * inv);
  }
  reinterpret_cast<uchar4*>(rgba)[pix] = out;
  if (written) written[pix] = 1;
}

// ---- rt_trace helpers ----------------------------------------------------------------------------
__global__ void pack_rays_kernel(const float* __restrict__ org, const float* __restrict__ dir, int n, float4* __restrict__ q) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  q[2 * (size_t)i] = make_float4(org[3 * (size_t)i], org[3 * (size_t)i + 1], org[3 * (size_t)i + 2], INFINITY);
  q[2 * (size_t)i + 1] = make_float4(dir[3 * (size_t)i], dir[3 * (size_t)i + 1], dir[3 * (size_t)i + 2], __uint_as_float((unsigned)i));
}
__global__ void unpack_hits_kernel(const float2* __restrict__ hits, const uint32_t* __restrict__ perm, int n, int32_t* __restrict__ ids,
                                   float* __restrict__ ts) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float2 h = hits[i];
  int id = __float_as_int(h.y);
  if (id >= 0) id = (int)perm[(unsigned)id & RT_LEAF_MASK];
  ids[i] = id;
  if (ts) ts[i] = id == -1 ? INFINITY : h.x;
}

// ---- host side -----------------------------------------------------------------------------------
struct RngTables {  // device copies of the skip-ahead matrices, one set per device
  uint32_t* mats = nullptr; int mats_count = 0;
  uint32_t* nib = nullptr;
  uint32_t* byt = nullptr;
  uint32_t* base = nullptr; int base_count = 0;
};

struct RenderState {
  char* pool[2] = {nullptr, nullptr}; size_t pool_bytes[2] = {0, 0};  // one per pass in flight
  size_t dflt_pool = 0;  // default RT_POOL_BYTES of this device
  uint32_t* counters = nullptr; unsigned long long* stats = nullptr;  // counters: 16 words per pass context
  uint32_t* h_count = nullptr;  // pinned, 16 words per pass context
  RngTables rng;
  cudaEvent_t e0 = nullptr, e1 = nullptr, t0[2] = {nullptr, nullptr}, t1[2] = {nullptr, nullptr}, ev_cnt[2] = {nullptr, nullptr};
  cudaStream_t pass_stream = nullptr; cudaEvent_t ev_fork2 = nullptr, ev_join2 = nullptr;  // second pass in flight
  uint32_t* h_blocks = nullptr; size_t h_blocks_words = 0;  // pinned: counter blocks of every queued pass of a frame
  std::vector<cudaEvent_t> ev_pool;                         // traverse timing of queued passes
  cudaStream_t head_stream = nullptr; cudaEvent_t ev_join = nullptr;  // RT_RENDER_OVERLAP_BUILD
  int trav_grid = 0;
  uint8_t* d_rgba = nullptr; int32_t* d_ids = nullptr; uint8_t* d_written = nullptr; size_t img_px = 0;  // rt_render_host staging
  float4* trace_q = nullptr; float2* trace_hits = nullptr; unsigned* trace_fb = nullptr; size_t trace_cap = 0;
};

RenderState* get_state(rt_scene* s) {
  if (!s->rstate) s->rstate = new RenderState();
  return (RenderState*)s->rstate;
}

int ensure_common(rt_scene* s, RenderState* st) {
  if (!st->counters) {
    RT_CUDA(cudaMalloc(&st->counters, 2 * 256 * 4));  // 2 pass contexts x 16 blocks x 16 words
    RT_CUDA(cudaMalloc(&st->stats, 128));
    RT_CUDA(cudaMallocHost(&st->h_count, 128));
    RT_CUDA(cudaEventCreate(&st->e0)); RT_CUDA(cudaEventCreate(&st->e1));
    for (int k = 0; k < 2; k++) {
      RT_CUDA(cudaEventCreate(&st->t0[k])); RT_CUDA(cudaEventCreate(&st->t1[k]));
      RT_CUDA(cudaEventCreateWithFlags(&st->ev_cnt[k], cudaEventDisableTiming));
    }
    int per_sm = 0, sms = 0;
    RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, traverse_near_kernel<true, false, false, false>, RT_TRAV_THREADS, 0));
    RT_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device));
    if (per_sm < 1) per_sm = 1;
    if (const char* e = getenv("RT_TRAV_CTAS_PER_SM")) { const int v = atoi(e); if (v >= 1 && v < per_sm) per_sm = v; }  // experiments
    st->trav_grid = sms * per_sm;  // persistent grid: every SM full, 148 x resident CTAs on B200
  }
  return RT_OK;
}

RtDeviceScene device_scene(const rt_scene* s) {
  RtDeviceScene d;
  d.n = s->n; d.root = (const int*)(s->bounds + 8); d.nodes = s->nodes; d.wnodes = (s->built_wide && !getenv("RT_NO_WIDE")) ? s->wnodes : nullptr; d.sorted_tv = s->sorted_tv; d.perm = s->sorted_perm; d.prim_mat = s->prim_mat;
  d.suns = s->suns; d.num_suns = s->cfg.num_suns; d.bulbs = s->bulbs; d.num_bulbs = s->cfg.num_bulbs;
  d.planes = s->planes; d.num_planes = s->cfg.num_planes;
  return d;
}

// BASELINE.json's north star asks for the node array "pinned in L2 with an access-policy window".  Opt-in
// (RT_L2_PERSIST_NODES=1): part of L2 is set aside for persisting lines and the stream's window covers the node array
// (the wide one when built) with hit ratio 1.  Off by default: the traverse launches already run at a 92-98 % L2 hit
// rate with 1 % DRAM throughput, and the set-aside takes L2 from the 96 MB of leaf records and the ray stream.
void set_node_window(const rt_scene* s, cudaStream_t st, bool on) {
  cudaStreamAttrValue v; memset(&v, 0, sizeof(v));
  if (on) {
    int dev = 0, max_win = 0, max_persist = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, dev);
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, dev);
    const bool wide = s->built_wide && !getenv("RT_NO_WIDE");
    const size_t bytes = (size_t)(s->n > 1 ? s->n - 1 : 1) * (wide ? 128 : 64);
    const size_t win = std::min<size_t>(bytes, (size_t)std::max(max_win, 0));
    cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, std::min<size_t>(win, (size_t)std::max(max_persist, 0)));
    v.accessPolicyWindow.base_ptr = (void*)(wide ? s->wnodes : s->nodes);
    v.accessPolicyWindow.num_bytes = win;
    v.accessPolicyWindow.hitRatio = max_persist > 0 ? std::min(1.0f, (float)max_persist / (float)std::max<size_t>(win, 1)) : 0.f;
    v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
  }
  if (cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v) != cudaSuccess) cudaGetLastError();  // a hint, never an error
}

int lane_of(unsigned flags) {
  return (flags & RT_TRAVERSE_REF_ORDER) ? RT_LANE_REF : (flags & RT_TRAVERSE_NEAR_FIRST) ? RT_LANE_NEAR : RT_LANE_EXACT;
}
// one traversal of a two-ended ray queue (slots [0, nf) and [cap - nb, cap)): the lane's kernel on each end, then
// (exact lane) the reference-order pass over the rays whose certificate failed.  Returns the number of launches.
// ctr: [0] fetch ticket front, [3] fallback count, [4] fallback fetch ticket, [6] fetch ticket back.
template <bool EXACT, bool WIDE>
void launch_near_w(bool count, bool sun, int grid, cudaStream_t st, const RtDeviceScene& sc, const float4* rays, unsigned first, unsigned n,
                   float2* hits, unsigned* fetch, unsigned long long* stats, unsigned* fb_list, unsigned* fb_count, const unsigned* n_dev, unsigned cap_back,
                   unsigned* hits2) {
  if (sun) {
    if (count) traverse_near_kernel<EXACT, true, true, WIDE><<<grid, RT_TRAV_THREADS, 0, st>>>(sc, rays, first, n, hits, fetch, stats, fb_list, fb_count, n_dev, cap_back, nullptr);
    else traverse_near_kernel<EXACT, false, true, WIDE><<<grid, RT_TRAV_THREADS, 0, st>>>(sc, rays, first, n, hits, fetch, stats, fb_list, fb_count, n_dev, cap_back, nullptr);
  } else {
    if (count) traverse_near_kernel<EXACT, true, false, WIDE><<<grid, RT_TRAV_THREADS, 0, st>>>(sc, rays, first, n, hits, fetch, stats, fb_list, fb_count, n_dev, cap_back, hits2);
    else traverse_near_kernel<EXACT, false, false, WIDE><<<grid, RT_TRAV_THREADS, 0, st>>>(sc, rays, first, n, hits, fetch, stats, fb_list, fb_count, n_dev, cap_back, hits2);
  }
}
// the packet kernel's persistent grid: every SM full
template <bool EXACT>
int packet_grid(bool sun) {
  static int g[2] = {0, 0};
  int& w = g[sun ? 1 : 0];
  if (w == 0) {
    int per_sm = 0, sms = 148, dev = 0;
    if (sun) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, traverse_packet_kernel<EXACT, false, true>, RT_PKT_THREADS, 0);
    else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, traverse_packet_kernel<EXACT, false, false>, RT_PKT_THREADS, 0);
    cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (per_sm < 1) per_sm = 1;
    if (const char* e = getenv("RT_PKT_CTAS_PER_SM")) { const int v = atoi(e); if (v >= 1 && v < per_sm) per_sm = v; }  // experiments
    w = sms * per_sm;
  }
  return w;
}
template <bool EXACT>
void launch_packet(bool count, bool sun, long long n_hint, cudaStream_t st, const RtDeviceScene& sc, const float4* rays, unsigned first, unsigned n,
                   float2* hits, unsigned* fetch, unsigned long long* stats, unsigned* fb_list, unsigned* fb_count, const unsigned* n_dev, unsigned cap_back) {
  int grid = packet_grid<EXACT>(sun);
  if (n_hint >= 0) grid = (int)std::max<long long>(1, std::min<long long>(grid, (n_hint + RT_PKT_THREADS - 1) / RT_PKT_THREADS));
  if (sun) {
    if (count) traverse_packet_kernel<EXACT, true, true><<<grid, RT_PKT_THREADS, 0, st>>>(sc, rays, first, n, hits, fetch, stats, fb_list, fb_count, n_dev, cap_back);
    else traverse_packet_kernel<EXACT, false, true><<<grid, RT_PKT_THREADS, 0, st>>>(sc, rays, first, n, hits, fetch, stats, fb_list, fb_count, n_dev, cap_back);
  } else {
    if (count) traverse_packet_kernel<EXACT, true, false><<<grid, RT_PKT_THREADS, 0, st>>>(sc, rays, first, n, hits, fetch, stats, fb_list, fb_count, n_dev, cap_back);
    else traverse_packet_kernel<EXACT, false, false><<<grid, RT_PKT_THREADS, 0, st>>>(sc, rays, first, n, hits, fetch, stats, fb_list, fb_count, n_dev, cap_back);
  }
}
template <bool EXACT>
void launch_near(bool count, bool sun, int grid, cudaStream_t st, const RtDeviceScene& sc, const float4* rays, unsigned first, unsigned n,
                 float2* hits, unsigned* fetch, unsigned long long* stats, unsigned* fb_list, unsigned* fb_count, const unsigned* n_dev, unsigned cap_back,
                 unsigned* hits2) {
  if (sc.wnodes && sc.n > 1) {
    static int wide_per_sm[2] = {0, 0};  // the wide kernels hold more registers: fewer resident CTAs, hence a smaller persistent grid
    int& w = wide_per_sm[sun ? 1 : 0];
    if (w == 0) {
      if (sun) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&w, traverse_near_kernel<EXACT, false, true, true>, RT_TRAV_THREADS, 0);
      else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&w, traverse_near_kernel<EXACT, false, false, true>, RT_TRAV_THREADS, 0);
      if (w < 1) w = 1;
    }
    int sms = 148; { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
    grid = std::min(grid, sms * w);
    launch_near_w<EXACT, true>(count, sun, grid, st, sc, rays, first, n, hits, fetch, stats, fb_list, fb_count, n_dev, cap_back, hits2);
  }
  else launch_near_w<EXACT, false>(count, sun, grid, st, sc, rays, first, n, hits, fetch, stats, fb_list, fb_count, n_dev, cap_back, hits2);
}
// counts_dev == nullptr: nf / nb are the ray counts (the host read them back).  counts_dev != nullptr (a pass queued without host
// round trips): the kernels read the counts from counts_dev[1] (front) and counts_dev[5] (back); nf / nb only say which
// ends can hold rays at all (0 = skip the launch), and the grids are the full persistent grid.
// packet: the rays of this queue are coherent in queue order (primary rays and the shadow rays of their hits): one warp
// walks the tree once for 32 consecutive rays (traverse_packet_kernel) instead of 32 independent lanes
int launch_traverse(int lane, bool count, int max_grid, cudaStream_t st, const RtDeviceScene& sc, const float4* rays, unsigned nf, unsigned nb,
                    unsigned cap, float2* hits, unsigned* ctr, unsigned long long* stats, unsigned* fb_list, const unsigned* counts_dev = nullptr,
                    bool packet = false, unsigned* hits2 = nullptr) {  // hits2: inline sun shadow rays (near-first lanes, front end only)
  int launches = 0;
  for (int end = 0; end < 2; end++) {
    const unsigned n = end ? nb : nf, first = end ? cap - nb : 0u;
    if (n == 0) continue;
    unsigned* fetch = ctr + (end ? 6 : 0);
    const unsigned* n_dev = counts_dev ? counts_dev + (end ? 5 : 1) : nullptr;
    const unsigned cap_back = (counts_dev && end) ? cap : 0u;
    const int grid = counts_dev ? max_grid : (int)std::min<long long>(max_grid, ((long long)n + RT_TRAV_THREADS - 1) / RT_TRAV_THREADS);
    if (lane == RT_LANE_REF) {
      if (count) traverse_ref_kernel<true><<<grid, RT_TRAV_THREADS, 0, st>>>(sc, rays, first, n, hits, fetch, stats, nullptr, nullptr, n_dev, cap_back, nullptr);
      else traverse_ref_kernel<false><<<grid, RT_TRAV_THREADS, 0, st>>>(sc, rays, first, n, hits, fetch, stats, nullptr, nullptr, n_dev, cap_back, nullptr);
    } else if (packet && sc.n > 0) {
      const long long n_hint = counts_dev ? -1ll : (long long)n;
      if (lane == RT_LANE_NEAR) launch_packet<false>(count, end == 1, n_hint, st, sc, rays, first, n, hits, fetch, stats, fb_list, ctr + 3, n_dev, cap_back);
      else launch_packet<true>(count, end == 1, n_hint, st, sc, rays, first, n, hits, fetch, stats, fb_list, ctr + 3, n_dev, cap_back);
    } else if (lane == RT_LANE_NEAR) launch_near<false>(count, end == 1, grid, st, sc, rays, first, n, hits, fetch, stats, fb_list, ctr + 3, n_dev, cap_back, end == 0 ? hits2 : nullptr);
    else launch_near<true>(count, end == 1, grid, st, sc, rays, first, n, hits, fetch, stats, fb_list, ctr + 3, n_dev, cap_back, end == 0 ? hits2 : nullptr);
    launches++;
  }
  if (lane == RT_LANE_EXACT && nf + nb > 0) {
    const int g2 = max_grid < 148 ? max_grid : 148;
    if (count) traverse_ref_kernel<true><<<g2, RT_TRAV_THREADS, 0, st>>>(sc, rays, 0u, 0u, hits, ctr + 4, stats, fb_list, ctr + 3, nullptr, 0u, hits2);
    else traverse_ref_kernel<false><<<g2, RT_TRAV_THREADS, 0, st>>>(sc, rays, 0u, 0u, hits, ctr + 4, stats, fb_list, ctr + 3, nullptr, 0u, hits2);
    launches++;
  }
  return launches;
}

int ensure_rng_tables(RenderState* st, int mats_needed, int base_needed) {
  if (mats_needed > st->rng.mats_count) {
    std::vector<uint32_t> h;
    rt_xorwow_power_matrices(mats_needed, h);
    if (st->rng.mats) cudaFree(st->rng.mats);
    RT_CUDA(cudaMalloc(&st->rng.mats, h.size() * 4));
    RT_CUDA(cudaMemcpy(st->rng.mats, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    if (st->rng.nib) cudaFree(st->rng.nib);
    RT_CUDA(cudaMalloc(&st->rng.nib, (size_t)mats_needed * RT_RNG_NIB_STRIDE * 4));
    rng_nibble_kernel<<<(mats_needed * 640 + 255) / 256, 256>>>(st->rng.mats, mats_needed, st->rng.nib);
    if (st->rng.byt) cudaFree(st->rng.byt);
    RT_CUDA(cudaMalloc(&st->rng.byt, (size_t)mats_needed * RT_RNG_BYTE_STRIDE * 4));
    rng_byte_kernel<<<(mats_needed * 12 * 256 + 255) / 256, 256>>>(st->rng.nib, mats_needed, st->rng.byt);
    RT_CUDA(cudaGetLastError());
    RT_CUDA(cudaDeviceSynchronize());
    st->rng.mats_count = mats_needed;
  }
  if (base_needed > st->rng.base_count) {
    std::vector<uint32_t> h;
    rt_xorwow_base_vectors(1234ull, 256, base_needed, h);
    if (st->rng.base) cudaFree(st->rng.base);
    RT_CUDA(cudaMalloc(&st->rng.base, h.size() * 4));
    RT_CUDA(cudaMemcpy(st->rng.base, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    st->rng.base_count = base_needed;
  }
  return RT_OK;
}

size_t env_size(const char* name, size_t dflt) {
  const char* v = getenv(name);
  if (!v || !*v) return dflt;
  return (size_t)strtoull(v, nullptr, 10);
}

}  // namespace

int rt_render_free_state(rt_scene* s) {
  RenderState* st = (RenderState*)s->rstate;
  if (!st) return RT_OK;
  cudaFree(st->pool[0]); cudaFree(st->pool[1]); cudaFree(st->counters); cudaFree(st->stats); cudaFreeHost(st->h_count);
  cudaFree(st->rng.mats); cudaFree(st->rng.nib); cudaFree(st->rng.byt); cudaFree(st->rng.base); cudaFree(st->d_rgba); cudaFree(st->d_ids); cudaFree(st->d_written);
  cudaFree(st->trace_q); cudaFree(st->trace_hits); cudaFree(st->trace_fb);
  if (st->e0) {
    cudaEventDestroy(st->e0); cudaEventDestroy(st->e1);
    for (int k = 0; k < 2; k++) { cudaEventDestroy(st->t0[k]); cudaEventDestroy(st->t1[k]); cudaEventDestroy(st->ev_cnt[k]); }
  }
  if (st->h_blocks) cudaFreeHost(st->h_blocks);
  for (cudaEvent_t e : st->ev_pool) cudaEventDestroy(e);
  if (st->pass_stream) { cudaStreamDestroy(st->pass_stream); cudaEventDestroy(st->ev_fork2); cudaEventDestroy(st->ev_join2); }
  if (st->head_stream) { cudaStreamDestroy(st->head_stream); cudaEventDestroy(st->ev_join); }
  delete st;
  s->rstate = nullptr;
  return RT_OK;
}

extern "C" int rt_render(rt_scene* s, int w, int h, int aa, unsigned flags, const rt_tile_set* tiles, uint8_t* d_rgba, int32_t* d_hit_id,
                         uint8_t* d_written, rt_counters* counters, void* stream) {
  if (!s || !d_rgba) { rt_set_error("rt_render: null argument"); return RT_ERR_INVALID; }
  if (w <= 0 || h <= 0 || (long long)w * h > 0x7FFFFFFFll / 64) { rt_set_error("rt_render: bad image size"); return RT_ERR_INVALID; }
  if (s->n > 0 && !s->built) { rt_set_error("rt_render: rt_build_lbvh has not run"); return RT_ERR_NOT_BUILT; }
  if (tiles && (tiles->tile_w <= 0 || tiles->tile_h <= 0 || tiles->num_ranks <= 0 || tiles->rank < 0 || tiles->rank >= tiles->num_ranks)) {
    rt_set_error("rt_render: bad tile set"); return RT_ERR_INVALID;
  }
  RT_CUDA(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  RenderState* rs = get_state(s);
  int rc = ensure_common(s, rs); if (rc) return rc;

  RenderParams P; memset(&P, 0, sizeof(P));
  RtCamera& C = P.cam;
  C.width = w; C.height = h; C.bounces = s->cfg.bounces; C.gi = s->cfg.gi; C.dof_focus = s->cfg.dof_focus; C.dof_lens = s->cfg.dof_lens;
  C.expose = s->cfg.expose; C.fisheye = s->cfg.fisheye; C.panorama = s->cfg.panorama;
  for (int k = 0; k < 3; k++) { C.forward[k] = s->cfg.forward[k]; C.right[k] = s->cfg.right[k]; C.up[k] = s->cfg.up[k]; C.eye[k] = s->cfg.eye[k]; }
  P.W = w; P.H = h;
  const long long npx = (long long)w * h;
  if (aa <= 1) { P.mode = aa <= 0 ? 0 : 1; P.spp = 1; P.covered = (int)npx; }
  else {
    P.mode = 2;
    if (flags & RT_SPP_TRUE) { P.spp = aa; P.covered = (int)npx; }
    else {  // render(), draw.cu:226-238: ceil(P*aa/128) blocks of 4 warps, one warp = one pixel, 32 samples
      P.spp = 32;
      const long long grid = (npx * aa + 127) / 128;
      P.covered = (int)std::min<long long>(npx, grid * 4);
    }
  }
  if (tiles) { P.tile_w = std::min(tiles->tile_w, w); P.tile_h = std::min(tiles->tile_h, h); P.num_ranks = tiles->num_ranks; P.rank = tiles->rank; }
  else { P.tile_w = w; P.tile_h = h; P.num_ranks = 1; P.rank = 0; }
  P.tiles_x = (w + P.tile_w - 1) / P.tile_w;
  const int tiles_y = (h + P.tile_h - 1) / P.tile_h;
  const long long ntiles = (long long)P.tiles_x * tiles_y;
  const long long owned = ntiles > P.rank ? (ntiles - P.rank + P.num_ranks - 1) / P.num_ranks : 0;
  P.slots_per_tile = P.tile_w * P.tile_h;
  if (P.tile_w % RT_SUB_W == 0 && P.tile_h % RT_SUB_H == 0) { P.sub_w = RT_SUB_W; P.sub_h = RT_SUB_H; }
  else { P.sub_w = P.tile_w; P.sub_h = 1; }
  long long nslots = owned * P.slots_per_tile;
  if (!tiles && P.covered < npx) { P.sub_w = P.tile_w; P.sub_h = 1; nslots = P.covered; }  // F3's partial image, one tile: slot == pixel
  set_fastdivs(P);  // every divisor of the slot <-> pixel maps is final here
  P.ref_order = lane_of(flags);
  P.rough_zyx = (flags & RT_ROUGH_ZYX) ? 1 : 0;
  const bool count = (flags & RT_COUNTERS) != 0;
  const bool time_trav = (flags & (RT_COUNTERS | RT_TIME_TRAVERSE)) != 0;
  // packet traversal (opt-in, RT_TRAVERSE_PACKET): 0 never, 1 coherent queues (primary rays; every ray of a scene without
  // mirrors, glass or GI), 2 every launch.  Measured on the 4K soup frame (profiles/packet_kernel_r02.json): only 17 of 32
  // lanes want the node their warp fetches, 3300 Mrays/s against 4159 with independent lanes over the same pixel-major queue.
  int pkt_mode = (flags & RT_TRAVERSE_PACKET) ? 2 : 0;
  if (const char* e = getenv("RT_PACKET")) pkt_mode = atoi(e);  // experiments
  const bool both = s->any_shiny && s->any_trans;
  const int gi = s->cfg.gi != 0 ? 1 : 0;
  // Frames per sample.  Without mirrors, glass or GI the only frame a sample ever holds is the diffuse state of the hit
  // whose shadow rays are in flight (pushed at level 0 by the visit that shades the primary hit): one level.  Otherwise
  // the deepest stack is reached along a chain of reflections: every reflect call in progress keeps at most its own
  // W_REFLECT frame plus the provisional refract(outer) frame below it (draw.cu:424) while the hit it found pushes its
  // own gi / refract / reflect calls, i.e. two frames per nested bounce when a primitive is both shiny and transparent,
  // bounded by `bounces` (each nested call runs with bounce - 1, draw.cu:404,474) or |gi| for the GI chain; one pending
  // call otherwise.  + 2 for the frames of the innermost hit (pending call being expanded + diffuse state), + 1 with GI.
  // tests/test_render_gpu.py::test_deep_recursion_fits_the_continuation_stack drives the worst case (every primitive shiny
  // and transparent, bounces 16, ~90 000 rays per pixel) against the oracle's own recursion.
  const bool full_shade = s->any_shiny || s->any_trans || s->cfg.gi != 0;
  P.depth = !full_shade ? 1 : 2 + gi + (both ? 2 * std::max(std::max(s->cfg.bounces, std::abs(s->cfg.gi)), 1) + 2 : ((s->any_shiny || s->any_trans) ? 2 : 0));
  if (P.depth > 200) { rt_set_error("rt_render: bounces too large for the continuation stack"); return RT_ERR_INVALID; }
  P.rng_live = (s->any_rough || s->cfg.gi != 0) ? 1 : 0;
  P.full_shade = (s->any_shiny || s->any_trans || s->cfg.gi != 0) ? 1 : 0;
  P.need_rng = (P.mode != 0) || s->cfg.dof_focus != 0.0f || s->any_rough || s->cfg.gi != 0;
  if (s->cfg.fisheye || s->cfg.panorama) { if (P.mode == 0 && !s->any_rough && s->cfg.gi == 0) P.need_rng = 0; }
  P.num_lights = s->cfg.num_suns + s->cfg.num_bulbs;

  // RNG tables
  if (P.mode == 2) { rc = ensure_rng_tables(rs, std::max(32, P.spp), 0); if (rc) return rc; }
  else if (P.need_rng) { rc = ensure_rng_tables(rs, 256, (int)((npx + 255) / 256)); if (rc) return rc; }

  // pass size: all per-sample state of a pass lives in one pool (RT_POOL_BYTES each)
  const size_t per_sample = 32 + 16 + 4 + (size_t)P.depth * 64 + 2 * 32 + 8 + 4;
  // 16 GB per pool (two pools): the 4K x 16 spp frame is two passes = 4 traverse launches instead of 12 with 6 GB pools,
  // and every launch costs ~0.46 ms whatever its size (measured 47.6 -> 46.6 ms per frame).  A B200 has 180 GB; on a
  // smaller or busy device the default shrinks to a quarter of the free memory.
  if (rs->dflt_pool == 0) {  // decided once per scene (cudaMemGetInfo is a driver round trip, not something for every frame)
    rs->dflt_pool = (size_t)16 << 30;
    size_t fr = 0, tot = 0;
    if (cudaMemGetInfo(&fr, &tot) == cudaSuccess) rs->dflt_pool = std::min(rs->dflt_pool, std::max<size_t>(fr / 4, (size_t)64 << 20));
  }
  const size_t budget = env_size("RT_POOL_BYTES", rs->dflt_pool);
  const long long total_samples = nslots * P.spp;
  long long max_s = (long long)(budget / per_sample);
  long long npix_pass = std::max<long long>(1, max_s / P.spp);
  if (npix_pass >= nslots) npix_pass = nslots;
  else npix_pass = std::max<long long>(32, npix_pass & ~31ll);
  // TWO PASSES IN FLIGHT.  Every traverse launch pays ~0.46 ms whatever its size (scripts/launch_fixed_cost.py): the
  // kernel cannot end before its longest rays have, a ray is a chain of dependent node fetches, and the longest of a
  // launch's rays takes ~30x the mean number of steps.  Passes are independent (disjoint pixels), so two of them run on
  // two streams with a pool each and the kernels of one fill what the other leaves idle: 47.9 -> 46.6 ms on the 4K
  // frame.  (Less than the 4 x 0.46 ms the drains add up to: every resident CTA holds one of the long rays, so the SMs
  // free up late.)  Same pixels: a pass is the same computation on whichever stream it runs
  // (tests/test_render_gpu.py::test_pass_schedules_give_the_same_pixels, test_multi_pass_equals_single_pass).
  const bool serial = (flags & RT_RENDER_SERIAL_PASSES) != 0 || getenv("RT_SERIAL_PASSES") != nullptr;
  // A frame that fits one pass stays one pass: splitting it in two hides one drain and pays two more launches
  // (measured: 24.2 ms as one pass, 25.0 ms as two half passes in flight on a 1/2 slice of the 4K frame).
  long long npass0 = nslots > 0 ? (nslots + npix_pass - 1) / npix_pass : 0;
  if (const char* e = getenv("RT_MIN_PASSES")) { const long long v = atoll(e); if (npass0 > 0 && v > npass0 && total_samples >= (1ll << 21)) npass0 = v; }  // experiments
  const int nctx = (!serial && npass0 >= 2) ? 2 : 1;
  if (nslots > 0) {  // balanced passes (an even number of them when two are in flight)
    const long long npass = nctx == 2 ? ((npass0 + 1) & ~1ll) : npass0;
    npix_pass = std::min<long long>(nslots, std::max<long long>(32, ((nslots + npass - 1) / npass + 31) & ~31ll));
  } else npix_pass = 0;
  const size_t Smax = (size_t)npix_pass * P.spp;
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  const size_t need = al(Smax * 16) * 2 + al(Smax * 16) + al(Smax * 4) + al(Smax * 64 * P.depth) + al(Smax * 32) * 2 + al(Smax * 8) + al(Smax * 4) + 4096;
  const RtDeviceScene sc = device_scene(s);
  struct PassCtx {
    RenderBuffers B; RenderParams P; cudaStream_t st = nullptr; uint32_t* counters = nullptr; uint32_t* h_count = nullptr;
    cudaEvent_t ev_cnt = nullptr, t0 = nullptr, t1 = nullptr;
    bool active = false, after_gen = false; unsigned n_front = 0, n_back = 0; float4 *qin = nullptr, *qout = nullptr; int iter_in_pass = 0;
  } ctx[2];
  if (nctx == 2 && !rs->pass_stream) {
    RT_CUDA(cudaStreamCreateWithFlags(&rs->pass_stream, cudaStreamNonBlocking));
    RT_CUDA(cudaEventCreateWithFlags(&rs->ev_fork2, cudaEventDisableTiming));
    RT_CUDA(cudaEventCreateWithFlags(&rs->ev_join2, cudaEventDisableTiming));
  }
  for (int k = 0; k < nctx; k++) {
    if (need > rs->pool_bytes[k]) {
      if (rs->pool[k]) cudaFree(rs->pool[k]);
      rs->pool[k] = nullptr; rs->pool_bytes[k] = 0;
      cudaError_t e = cudaMalloc((void**)&rs->pool[k], need);
      if (e != cudaSuccess) { rt_set_error(std::string("rt_render: state pool cudaMalloc: ") + cudaGetErrorString(e)); return RT_ERR_NOMEM; }
      rs->pool_bytes[k] = need;
    }
    PassCtx& c = ctx[k];
    RenderBuffers& B = c.B; memset(&B, 0, sizeof(B));
    char* p = rs->pool[k];
    B.rng0 = (uint4*)p; p += al(Smax * 16); B.rng1 = (uint4*)p; p += al(Smax * 16);
    B.accum = (float4*)p; p += al(Smax * 16); B.ctl = (uint32_t*)p; p += al(Smax * 4);
    B.frames = (float4*)p; p += al(Smax * 64 * P.depth);
    B.qa = (float4*)p; p += al(Smax * 32); B.qb = (float4*)p; p += al(Smax * 32);
    B.hits = (float2*)p; p += al(Smax * 8);
    B.fb_list = (unsigned*)p; p += al(Smax * 4);
    c.counters = rs->counters + 256 * k; c.h_count = rs->h_count + 16 * k;  // 16 counter blocks of 16 words per context
    B.counters = c.counters; B.stats = rs->stats; B.rng_mats = rs->rng.mats; B.rng_nib = rs->rng.nib; B.rng_byte = rs->rng.byt; B.rng_base = rs->rng.base;
    c.st = k == 0 ? st : rs->pass_stream;
    c.ev_cnt = rs->ev_cnt[k]; c.t0 = rs->t0[k]; c.t1 = rs->t1[k];
  }

  // The head of the frame (clears, RNG seeding and primary ray generation of the first pass) reads nothing the build
  // writes.  With RT_RENDER_OVERLAP_BUILD it runs on its own stream from the point in `st` before this frame's
  // update/build calls (s->ev_fork) and joins `st` before the first traversal.
  const bool overlap = (flags & RT_RENDER_OVERLAP_BUILD) && s->fork_armed && s->fork_stream == st && !getenv("RT_NO_OVERLAP");
  s->fork_armed = false;
  cudaStream_t hs = st;
  if (overlap) {
    if (!rs->head_stream) {
      RT_CUDA(cudaStreamCreateWithFlags(&rs->head_stream, cudaStreamNonBlocking));
      RT_CUDA(cudaEventCreateWithFlags(&rs->ev_join, cudaEventDisableTiming));
    }
    hs = rs->head_stream;
    RT_CUDA(cudaStreamWaitEvent(hs, s->ev_fork, 0));
  }
  const bool l2win = getenv("RT_L2_PERSIST_NODES") != nullptr && s->n > 1;
  if (l2win) for (int k = 0; k < nctx; k++) set_node_window(s, ctx[k].st, true);
  RT_CUDA(cudaEventRecord(rs->e0, st));
  RT_CUDA(cudaMemsetAsync(d_rgba, 0, (size_t)npx * 4, hs));
  if (d_written) RT_CUDA(cudaMemsetAsync(d_written, 0, (size_t)npx, hs));
  if (d_hit_id) RT_CUDA(cudaMemsetAsync(d_hit_id, 0xFF, (size_t)npx * 4, hs));
  RT_CUDA(cudaMemsetAsync(rs->stats, 0, 128, hs));
  RT_CUDA(cudaMemsetAsync(rs->counters, 0, 2 * 256 * 4, hs));
  if (nctx == 2) {  // the second pass stream starts after the caller's work on `st` (the build) and after the clears
    RT_CUDA(cudaEventRecord(rs->ev_fork2, st));
    RT_CUDA(cudaStreamWaitEvent(rs->pass_stream, rs->ev_fork2, 0));
    if (hs != st) {
      RT_CUDA(cudaEventRecord(rs->ev_join2, hs));
      RT_CUDA(cudaStreamWaitEvent(rs->pass_stream, rs->ev_join2, 0));
    }
  }
  uint64_t rays = 0, samples = 0; uint32_t iters = 0, passes = 0, launches = 0; float trav_ms = 0.f; uint64_t fb_rays = 0; uint32_t trav_launches = 0;
  long long next_slot0 = 0;
  bool head_joined = (hs == st);
  int err = RT_OK;
#define RT_PASS_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { rt_set_error(std::string("rt_render: ") + cudaGetErrorString(e_)); return RT_ERR_CUDA; } } while (0)
  // queue the head of the next pass (RNG seeding, primary rays) on context c and ask for its ray count
  auto start_pass = [&](PassCtx& c) -> int {
    c.P = P;
    c.P.slot0 = (int)next_slot0; c.P.npix = (int)std::min<long long>(npix_pass, nslots - next_slot0); c.P.S = c.P.npix * c.P.spp;
    next_slot0 += npix_pass;
    passes++;
    samples += (uint64_t)c.P.S;
    // the very first pass of the frame may run its head on the build-overlap stream
    cudaStream_t gs = (!head_joined && c.st == st) ? hs : c.st;
    RT_PASS_CUDA(cudaMemsetAsync(c.counters, 0, 32, gs));
    if (c.P.mode != 2 && c.P.need_rng) {
      const int hn = (int)(((npx + 255) / 256 + 31) & ~31ll);
      rng_init_pixels_kernel<<<(unsigned)(((long long)hn * 256 + 255) / 256), 256, 0, gs>>>(c.P, c.B, hn);
      launches++;
    }
    generate_kernel<<<(unsigned)((c.P.S + 256 * RT_GEN_CHUNKS - 1) / (256 * RT_GEN_CHUNKS)), 256, 0, gs>>>(c.P, c.B, c.B.qa, d_hit_id);
    launches++;
    RT_PASS_CUDA(cudaGetLastError());
    if (gs != c.st) {  // join: everything below is ordered after both the build and the head
      RT_PASS_CUDA(cudaEventRecord(rs->ev_join, gs));
      RT_PASS_CUDA(cudaStreamWaitEvent(c.st, rs->ev_join, 0));
      head_joined = true;
    }
    RT_PASS_CUDA(cudaMemcpyAsync(c.h_count, c.counters + 1, 28, cudaMemcpyDeviceToHost, c.st));  // [1] slots to scan, [7] rays among them
    RT_PASS_CUDA(cudaEventRecord(c.ev_cnt, c.st));
    c.active = true; c.after_gen = true; c.qin = c.B.qa; c.qout = c.B.qb; c.iter_in_pass = 0;
    return RT_OK;
  };
  // the ray counts context c asked for are in: queue its next wavefront iteration, or finish the pass and start another
  auto advance = [&](PassCtx& c) -> int {
    RT_PASS_CUDA(cudaEventSynchronize(c.ev_cnt));
    uint64_t nrays_iter = 0;
    if (c.after_gen) { c.n_front = c.h_count[0]; c.n_back = 0; c.after_gen = false; nrays_iter = c.h_count[6]; }  // primary rays: front end, slot = sample id
    else {
      fb_rays += c.h_count[2];
      if (time_trav) { float ms = 0.f; RT_PASS_CUDA(cudaEventElapsedTime(&ms, c.t0, c.t1)); trav_ms += ms; }
      c.n_front = c.h_count[0]; c.n_back = c.h_count[4];
      if (c.h_count[1]) { rt_set_error("rt_render: continuation stack overflow (more nested bounces than the stack was sized for)"); return RT_ERR_INVALID; }
      std::swap(c.qin, c.qout);
      nrays_iter = (uint64_t)c.n_front + c.n_back;
    }
    const unsigned n_front = c.n_front, n_back = c.n_back;
    if (n_front + n_back > 0) {
      iters++; rays += nrays_iter;
      RT_PASS_CUDA(cudaMemsetAsync(c.counters, 0, 28, c.st));
      if (time_trav) RT_PASS_CUDA(cudaEventRecord(c.t0, c.st));
      const bool pkt = pkt_mode == 2 || (pkt_mode == 1 && (c.iter_in_pass == 0 || !P.full_shade));
      const int tl = launch_traverse(P.ref_order, count, rs->trav_grid, c.st, sc, (const float4*)c.qin, n_front, n_back, (unsigned)c.P.S, c.B.hits,
                                     c.counters, rs->stats, c.B.fb_list, nullptr, pkt);
      launches += tl; trav_launches += tl - (P.ref_order == RT_LANE_EXACT ? 1 : 0);  // the (near-empty) fallback pass is not a traversal launch of its own
      if (time_trav) RT_PASS_CUDA(cudaEventRecord(c.t1, c.st));
      {
        const unsigned sg = (n_front + n_back + 127) / 128;
        const bool first = c.iter_in_pass == 0;
        if (P.full_shade) { if (first) shade_kernel<true, true><<<sg, 128, 0, c.st>>>(c.P, sc, c.B, c.qin, n_front, n_back, c.qout, d_hit_id, nullptr);
                            else shade_kernel<true, false><<<sg, 128, 0, c.st>>>(c.P, sc, c.B, c.qin, n_front, n_back, c.qout, d_hit_id, nullptr); }
        else { if (first) shade_kernel<false, true><<<sg, 128, 0, c.st>>>(c.P, sc, c.B, c.qin, n_front, n_back, c.qout, d_hit_id, nullptr);
               else shade_kernel<false, false><<<sg, 128, 0, c.st>>>(c.P, sc, c.B, c.qin, n_front, n_back, c.qout, d_hit_id, nullptr); }
        c.iter_in_pass++;
      }
      launches++;
      RT_PASS_CUDA(cudaGetLastError());
      RT_PASS_CUDA(cudaMemcpyAsync(c.h_count, c.counters + 1, 20, cudaMemcpyDeviceToHost, c.st));
      RT_PASS_CUDA(cudaEventRecord(c.ev_cnt, c.st));
      return RT_OK;
    }
    resolve_kernel<<<(c.P.npix + 127) / 128, 128, 0, c.st>>>(c.P, c.B, d_rgba, d_written);
    launches++;
    RT_PASS_CUDA(cudaGetLastError());
    c.active = false;
    if (next_slot0 < nslots) return start_pass(c);
    return RT_OK;
  };
  // QUEUED PASSES.  Without mirrors, glass or GI a sample's life is: primary ray, then one shadow ray per light, so a
  // pass is exactly 1 + num_lights wavefront iterations and nothing about its launches depends on the ray counts
  // except sizes.  The kernels then read their ray counts from device memory (one counter block per iteration:
  // generate fills block 0, iteration k traverses the counts of block k and its shade kernel appends to block k+1),
  // the whole frame is queued without a single host round trip, and the counts come back in one copy per pass for
  // the accounting.  No bubbles between kernels, and nothing for a busy host to delay.  Scenes whose continuation
  // stack can hold pending calls keep the host-driven loop above (their iteration count depends on the data).
  const int nl = P.num_lights;
  const bool queued = !P.full_shade && nl + 2 <= 16 && getenv("RT_HOST_DRIVEN") == nullptr;
  const long long npasses_total = npix_pass > 0 ? (nslots + npix_pass - 1) / npix_pass : 0;
  // (a lambda: RT_CUDA returns from IT, so a failure inside lands in the common error path below, which waits for the
  // kernels already queued on the pass and head streams before the call returns — as the host-driven loop does)
  // inline sun shadow rays (traverse.cuh): one traverse launch and one shade visit per pass
  // (experiment build only, -DRT_INLINE_SUN=1: measured slower than the two specialised launches, traverse.cuh)
  const bool inline_sun = RT_INLINE_SUN && queued && nl == 1 && s->cfg.num_suns == 1 && !s->any_rough && P.ref_order != RT_LANE_REF && pkt_mode == 0 &&
                          getenv("RT_NO_INLINE_SUN") == nullptr;
  auto run_queued = [&]() -> int {
    const size_t words = (size_t)npasses_total * 256;
    if (words > rs->h_blocks_words) {
      if (rs->h_blocks) cudaFreeHost(rs->h_blocks);
      rs->h_blocks = nullptr; rs->h_blocks_words = 0;
      RT_CUDA(cudaMallocHost(&rs->h_blocks, words * 4));
      rs->h_blocks_words = words;
    }
    const size_t nev = time_trav ? (size_t)npasses_total * (nl + 1) * 2 : 0;
    while (rs->ev_pool.size() < nev) { cudaEvent_t e; RT_CUDA(cudaEventCreate(&e)); rs->ev_pool.push_back(e); }
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device);
    for (long long pi = 0; pi < npasses_total; pi++) {
      PassCtx& c = ctx[pi % nctx];
      c.P = P;
      c.P.slot0 = (int)next_slot0; c.P.npix = (int)std::min<long long>(npix_pass, nslots - next_slot0); c.P.S = c.P.npix * c.P.spp;
      next_slot0 += npix_pass;
      passes++; samples += (uint64_t)c.P.S;
      cudaStream_t gs = (!head_joined && c.st == st) ? hs : c.st;
      RT_CUDA(cudaMemsetAsync(c.counters, 0, (size_t)(nl + 2) * 64, gs));
      if (c.P.mode != 2 && c.P.need_rng) {
        const int hn = (int)(((npx + 255) / 256 + 31) & ~31ll);
        rng_init_pixels_kernel<<<(unsigned)(((long long)hn * 256 + 255) / 256), 256, 0, gs>>>(c.P, c.B, hn);
        launches++;
      }
      generate_kernel<<<(unsigned)((c.P.S + 256 * RT_GEN_CHUNKS - 1) / (256 * RT_GEN_CHUNKS)), 256, 0, gs>>>(c.P, c.B, c.B.qa, d_hit_id);  // counts -> block 0
      launches++;
      RT_CUDA(cudaGetLastError());
      if (gs != c.st) {
        RT_CUDA(cudaEventRecord(rs->ev_join, gs));
        RT_CUDA(cudaStreamWaitEvent(c.st, rs->ev_join, 0));
        head_joined = true;
      }
      float4 *qin = c.B.qa, *qout = c.B.qb;
      const unsigned sgrid = (unsigned)std::min<long long>(((long long)c.P.S + 127) / 128, (long long)sms * RT_SHADE_MIN_CTAS * 4);
      if (inline_sun) {
        uint32_t* blk = c.counters;
        unsigned* occl = (unsigned*)c.B.qb;  // the second queue is not used by these passes: one word per slot
        if (time_trav) RT_CUDA(cudaEventRecord(rs->ev_pool[2 * (pi * (nl + 1))], c.st));
        const int tl = launch_traverse(P.ref_order, count, rs->trav_grid, c.st, sc, (const float4*)qin, 1u, 0u, (unsigned)c.P.S, c.B.hits,
                                       blk, rs->stats, c.B.fb_list, blk, false, occl);
        launches += tl; trav_launches += tl - (P.ref_order == RT_LANE_EXACT ? 1 : 0);
        if (time_trav) {
          RT_CUDA(cudaEventRecord(rs->ev_pool[2 * (pi * (nl + 1)) + 1], c.st));
          RT_CUDA(cudaEventRecord(rs->ev_pool[2 * (pi * (nl + 1) + 1)], c.st)); RT_CUDA(cudaEventRecord(rs->ev_pool[2 * (pi * (nl + 1) + 1) + 1], c.st));  // no second iteration
        }
        RenderBuffers Bn = c.B; Bn.counters = blk + 16;
        shade_inline_kernel<<<sgrid, 128, 0, c.st>>>(c.P, sc, Bn, qin, occl, d_hit_id, blk);
        launches++;
      } else
      for (int it = 0; it <= nl; it++) {
        uint32_t* blk = c.counters + 16 * it;
        // which queue ends can hold rays: primary rays go to the front; shadow rays to a sun go to the back, to a bulb to the front
        const unsigned may_front = it == 0 ? 1u : (s->cfg.num_bulbs > 0 ? 1u : 0u), may_back = it == 0 ? 0u : (s->cfg.num_suns > 0 ? 1u : 0u);
        if (time_trav) RT_CUDA(cudaEventRecord(rs->ev_pool[2 * (pi * (nl + 1) + it)], c.st));
        const int tl = launch_traverse(P.ref_order, count, rs->trav_grid, c.st, sc, (const float4*)qin, may_front, may_back, (unsigned)c.P.S, c.B.hits,
                                       blk, rs->stats, c.B.fb_list, blk, pkt_mode != 0);
        launches += tl; trav_launches += tl - (P.ref_order == RT_LANE_EXACT ? 1 : 0);
        if (time_trav) RT_CUDA(cudaEventRecord(rs->ev_pool[2 * (pi * (nl + 1) + it) + 1], c.st));
        RenderBuffers Bn = c.B; Bn.counters = blk + 16;  // this iteration's shade appends the next iteration's rays
        if (it == 0) shade_kernel<false, true><<<sgrid, 128, 0, c.st>>>(c.P, sc, Bn, qin, 0u, 0u, qout, d_hit_id, blk);
        else shade_kernel<false, false><<<sgrid, 128, 0, c.st>>>(c.P, sc, Bn, qin, 0u, 0u, qout, d_hit_id, blk);
        launches++;
        std::swap(qin, qout);
      }
      resolve_kernel<<<(c.P.npix + 127) / 128, 128, 0, c.st>>>(c.P, c.B, d_rgba, d_written);
      launches++;
      RT_CUDA(cudaGetLastError());
      RT_CUDA(cudaMemcpyAsync(rs->h_blocks + 256 * pi, c.counters, (size_t)(nl + 2) * 64, cudaMemcpyDeviceToHost, c.st));
    }
    return RT_OK;
  };
  if (queued && npasses_total > 0) err = run_queued();
  else {
    for (int k = 0; k < nctx && err == RT_OK && next_slot0 < nslots; k++) err = start_pass(ctx[k]);
    while (err == RT_OK && (ctx[0].active || ctx[1].active))
      for (int k = 0; k < nctx && err == RT_OK; k++)
        if (ctx[k].active) err = advance(ctx[k]);
  }
#undef RT_PASS_CUDA
  if (err != RT_OK) {  // nothing of this frame stays in flight, and the caller's stream gets its default L2 policy back
    const std::string keep = rt_last_error();
    cudaDeviceSynchronize();
    if (l2win) for (int k = 0; k < nctx; k++) set_node_window(s, ctx[k].st, false);
    rt_set_error(keep);
    return err;
  }
  if (!head_joined) {  // nothing to render: the clears still belong to this call
    RT_CUDA(cudaEventRecord(rs->ev_join, hs));
    RT_CUDA(cudaStreamWaitEvent(st, rs->ev_join, 0));
  }
  if (nctx == 2) {  // the caller's stream continues after both pass streams
    RT_CUDA(cudaEventRecord(rs->ev_join2, rs->pass_stream));
    RT_CUDA(cudaStreamWaitEvent(st, rs->ev_join2, 0));
  }
  RT_CUDA(cudaEventRecord(rs->e1, st));
  RT_CUDA(cudaEventSynchronize(rs->e1));
  if (l2win) for (int k = 0; k < nctx; k++) set_node_window(s, ctx[k].st, false);  // the caller's stream gets its default policy back
  if (queued) {  // the accounting of the queued passes, from their counter blocks
    bool overflow = false;
    for (long long pi = 0; pi < npasses_total; pi++)
      for (int it = 0; it <= nl; it++) {
        const uint32_t* blk = rs->h_blocks + 256 * pi + 16 * it;
        const uint64_t nr = it == 0 ? (uint64_t)blk[7] : (uint64_t)blk[1] + blk[5];  // block 0: [1] = slots scanned, [7] = primary rays among them
        if (nr) iters++;
        rays += nr; fb_rays += blk[3];
        overflow = overflow || blk[16 + 2] != 0u;
        if (time_trav) { float ms = 0.f; RT_CUDA(cudaEventElapsedTime(&ms, rs->ev_pool[2 * (pi * (nl + 1) + it)], rs->ev_pool[2 * (pi * (nl + 1) + it) + 1])); trav_ms += ms; }
      }
    if (overflow) { rt_set_error("rt_render: continuation stack overflow (more nested bounces than the stack was sized for)"); return RT_ERR_INVALID; }
  }
  if (counters) {
    memset(counters, 0, sizeof(*counters));
    counters->rays = rays; counters->samples = samples; counters->wavefront_iters = iters; counters->passes = passes;
    counters->kernel_launches = launches; counters->traverse_ms = trav_ms; counters->traverse_launches = trav_launches;
    counters->fallback_rays = fb_rays;
    RT_CUDA(cudaEventElapsedTime(&counters->render_ms, rs->e0, rs->e1));
    unsigned long long hs[10];
    RT_CUDA(cudaMemcpy(hs, rs->stats, 80, cudaMemcpyDeviceToHost));
    counters->nodes_visited = hs[0]; counters->leaf_tests = hs[1]; counters->shaded_hits = hs[2]; counters->packet_steps = hs[9];
#ifdef RT_LANE_STATS
    if (count) {
      unsigned long long q[9]; RT_CUDA(cudaMemcpy(q, rs->stats, 72, cudaMemcpyDeviceToHost));
      fprintf(stderr, "[lane stats] node-phase iters %llu avg lanes %.2f (parked at leaf %.2f) | leaf-phase iters %llu avg lanes %.2f | idle lanes/iter %.2f\n",
              q[3], (double)q[5] / q[3], (double)q[7] / q[3], q[4], (double)q[6] / q[4], (double)q[8] / (q[3] + q[4]));
    }
#endif
  }
  (void)total_samples;
  return RT_OK;
}

// the device image rt_render_host renders into (the reference's cudaMalloc of d_image, main.cu:56-58), allocated ahead of
// the first frame so that the frame's wall time holds no allocation
extern "C" int rt_render_reserve(rt_scene* s, int w, int h) {
  if (!s || w <= 0 || h <= 0) { rt_set_error("rt_render_reserve: bad argument"); return RT_ERR_INVALID; }
  RT_CUDA(cudaSetDevice(s->device));
  RenderState* rs = get_state(s);
  const size_t npx = (size_t)w * h;
  if (npx > rs->img_px) {
    cudaFree(rs->d_rgba); cudaFree(rs->d_ids); cudaFree(rs->d_written);
    rs->d_rgba = nullptr; rs->d_ids = nullptr; rs->d_written = nullptr; rs->img_px = 0;
    RT_CUDA(cudaMalloc(&rs->d_rgba, npx * 4)); RT_CUDA(cudaMalloc(&rs->d_ids, npx * 4)); RT_CUDA(cudaMalloc(&rs->d_written, npx));
    rs->img_px = npx;
  }
  return RT_OK;
}

extern "C" int rt_render_host(rt_scene* s, int w, int h, int aa, unsigned flags, const rt_tile_set* tiles, uint8_t* h_rgba, int32_t* h_hit_id,
                              uint8_t* h_written, rt_counters* counters) {
  if (!s || !h_rgba) { rt_set_error("rt_render_host: null argument"); return RT_ERR_INVALID; }
  if (w <= 0 || h <= 0) { rt_set_error("rt_render_host: bad image size"); return RT_ERR_INVALID; }
  int rc = rt_render_reserve(s, w, h);
  if (rc) return rc;
  RenderState* rs = get_state(s);
  const size_t npx = (size_t)w * h;
  rc = rt_render(s, w, h, aa, flags, tiles, rs->d_rgba, h_hit_id ? rs->d_ids : nullptr, h_written ? rs->d_written : nullptr, counters, nullptr);
  if (rc) return rc;
  const auto t_d2h = std::chrono::steady_clock::now();
  RT_CUDA(cudaMemcpy(h_rgba, rs->d_rgba, npx * 4, cudaMemcpyDeviceToHost));  // main.cu:77
  if (counters) counters->d2h_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t_d2h).count();
  if (h_hit_id) RT_CUDA(cudaMemcpy(h_hit_id, rs->d_ids, npx * 4, cudaMemcpyDeviceToHost));
  if (h_written) RT_CUDA(cudaMemcpy(h_written, rs->d_written, npx, cudaMemcpyDeviceToHost));
  return RT_OK;
}

extern "C" int rt_trace(rt_scene* s, int n, const float* d_org, const float* d_dir, unsigned flags, int32_t* d_hit_id, float* d_hit_t,
                        rt_counters* counters, void* stream) {
  if (!s || !d_org || !d_dir || !d_hit_id || n < 0) { rt_set_error("rt_trace: bad argument"); return RT_ERR_INVALID; }
  if (s->n > 0 && !s->built) { rt_set_error("rt_trace: rt_build_lbvh has not run"); return RT_ERR_NOT_BUILT; }
  RT_CUDA(cudaSetDevice(s->device));
  cudaStream_t st = (cudaStream_t)stream;
  s->fork_armed = false;
  RenderState* rs = get_state(s);
  int rc = ensure_common(s, rs); if (rc) return rc;
  if (n == 0) return RT_OK;
  if ((size_t)n > rs->trace_cap) {
    cudaFree(rs->trace_q); cudaFree(rs->trace_hits); cudaFree(rs->trace_fb); rs->trace_q = nullptr; rs->trace_hits = nullptr; rs->trace_fb = nullptr; rs->trace_cap = 0;
    RT_CUDA(cudaMalloc(&rs->trace_q, (size_t)n * 32)); RT_CUDA(cudaMalloc(&rs->trace_hits, (size_t)n * 8)); RT_CUDA(cudaMalloc(&rs->trace_fb, (size_t)n * 4));
    rs->trace_cap = (size_t)n;
  }
  const bool count = (flags & RT_COUNTERS) != 0;
  const RtDeviceScene sc = device_scene(s);
  RT_CUDA(cudaMemsetAsync(rs->counters, 0, 28, st));
  RT_CUDA(cudaMemsetAsync(rs->stats, 0, 128, st));
  pack_rays_kernel<<<(n + 255) / 256, 256, 0, st>>>(d_org, d_dir, n, rs->trace_q);
  RT_CUDA(cudaEventRecord(rs->e0, st));
  launch_traverse(lane_of(flags), count, rs->trav_grid, st, sc, (const float4*)rs->trace_q, (unsigned)n, 0u, (unsigned)n, rs->trace_hits, rs->counters,
                  rs->stats, rs->trace_fb, nullptr, (flags & RT_TRAVERSE_PACKET) != 0);  // arbitrary rays: per-ray lanes unless the caller says they are coherent
  RT_CUDA(cudaEventRecord(rs->e1, st));
  unpack_hits_kernel<<<(n + 255) / 256, 256, 0, st>>>(rs->trace_hits, s->sorted_perm, n, d_hit_id, d_hit_t);
  RT_CUDA(cudaGetLastError());
  if (counters) {
    RT_CUDA(cudaStreamSynchronize(st));
    memset(counters, 0, sizeof(*counters));
    counters->rays = (uint64_t)n; counters->kernel_launches = 3; counters->traverse_launches = 1;
    RT_CUDA(cudaEventElapsedTime(&counters->traverse_ms, rs->e0, rs->e1));
    counters->render_ms = counters->traverse_ms;
    unsigned long long hs[10];
    RT_CUDA(cudaMemcpy(hs, rs->stats, 80, cudaMemcpyDeviceToHost));
    counters->nodes_visited = hs[0]; counters->leaf_tests = hs[1]; counters->packet_steps = hs[9];
    unsigned fb = 0;
    RT_CUDA(cudaMemcpy(&fb, rs->counters + 3, 4, cudaMemcpyDeviceToHost));
    counters->fallback_rays = fb;
  }
  return RT_OK;
}

// ---- rt_trace_static: rt_trace for callers that must not allocate ---------------------------------------------------
// No counterpart in the reference.  Some callers cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  rt_trace keeps its scratch — the
// packed ray queue (32 B per ray), the hit records (8 B), the exact lane's fallback list (4 B), fetch tickets — in
// buffers it cudaMallocs on demand.  This variant keeps the same scratch in ONE file-scope device array of fixed size,
// so the call allocates nothing: no cudaMalloc / cudaFree / cudaMallocHost, no events, no library routine that allocates.
//   * limit: RT_TRACE_STATIC_MAX_RAYS rays per call (131072 => 5.5 MiB); more is rejected with RT_ERR_INVALID.  Every GPU
//     that loads libb200rt.so holds the array until the process exits.
//   * the array is shared by every call on the device and keeps the last call's contents: the tickets / fallback count
//     are cleared at the start of each call (cudaMemsetAsync on the caller's stream), everything else is overwritten
//     before it is read.  Calls that OVERLAP on the GPU (two streams, two host threads) race on it: serialise them.
//   * in host code `g_trace_scratch` is a host-side stand-in, not a device address: the address the kernels get is
//     fetched once per device, with that device current, by cudaGetSymbolAddress.
#define RT_TRACE_STATIC_MAX_RAYS 131072
#define RT_TRACE_STATIC_HEADER 512   /* 64 counter words + 32 stats words */
static __device__ __align__(256) unsigned char g_trace_scratch[RT_TRACE_STATIC_HEADER + (size_t)RT_TRACE_STATIC_MAX_RAYS * (32 + 8 + 4)];

namespace {
struct StaticTraceDev { unsigned char* scratch = nullptr; int grid = 0; };
std::mutex g_static_trace_mu;
StaticTraceDev g_static_trace[64];

int static_trace_setup(int device, StaticTraceDev& out) {  // the caller made `device` current
  if (device < 0 || device >= 64) { rt_set_error("rt_trace_static: device index out of range"); return RT_ERR_INVALID; }
  std::lock_guard<std::mutex> lk(g_static_trace_mu);
  StaticTraceDev& d = g_static_trace[device];
  if (!d.scratch) {
    void* p = nullptr;
    RT_CUDA(cudaGetSymbolAddress(&p, g_trace_scratch));
    int per_sm = 0, sms = 0;
    RT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, traverse_near_kernel<true, false, false, false>, RT_TRAV_THREADS, 0));
    RT_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    d.grid = sms * (per_sm < 1 ? 1 : per_sm);
    d.scratch = (unsigned char*)p;
  }
  out = d;
  return RT_OK;
}
}  // namespace

extern "C" int rt_trace_static_max_rays(void) { return RT_TRACE_STATIC_MAX_RAYS; }

extern "C" int rt_trace_static(rt_scene* s, int n, const float* d_org, const float* d_dir, unsigned flags, int32_t* d_hit_id, float* d_hit_t,
                               void* stream) {
  if (!s || !d_org || !d_dir || !d_hit_id || n < 0) { rt_set_error("rt_trace_static: bad argument"); return RT_ERR_INVALID; }
  if (n > RT_TRACE_STATIC_MAX_RAYS) {
    rt_set_error("rt_trace_static: " + std::to_string(n) + " rays exceed the static scratch (RT_TRACE_STATIC_MAX_RAYS = " +
                 std::to_string(RT_TRACE_STATIC_MAX_RAYS) + "); split the batch or use rt_trace");
    return RT_ERR_INVALID;
  }
  if (s->n > 0 && !s->built) { rt_set_error("rt_trace_static: rt_build_lbvh has not run"); return RT_ERR_NOT_BUILT; }
  RT_CUDA(cudaSetDevice(s->device));
  StaticTraceDev dev;
  const int rc = static_trace_setup(s->device, dev);
  if (rc) return rc;
  if (n == 0) return RT_OK;
  cudaStream_t st = (cudaStream_t)stream;
  uint32_t* ctr = (uint32_t*)dev.scratch;
  unsigned long long* stats = (unsigned long long*)(dev.scratch + 256);
  float4* q = (float4*)(dev.scratch + RT_TRACE_STATIC_HEADER);
  float2* hits = (float2*)(dev.scratch + RT_TRACE_STATIC_HEADER + (size_t)RT_TRACE_STATIC_MAX_RAYS * 32);
  unsigned* fb = (unsigned*)(dev.scratch + RT_TRACE_STATIC_HEADER + (size_t)RT_TRACE_STATIC_MAX_RAYS * 40);
  const RtDeviceScene sc = device_scene(s);
  RT_CUDA(cudaMemsetAsync(dev.scratch, 0, RT_TRACE_STATIC_HEADER, st));  // tickets, fallback count, stats: must start at zero
  pack_rays_kernel<<<(n + 255) / 256, 256, 0, st>>>(d_org, d_dir, n, q);
  launch_traverse(lane_of(flags), false, dev.grid, st, sc, (const float4*)q, (unsigned)n, 0u, (unsigned)n, hits, ctr, stats, fb, nullptr,
                  (flags & RT_TRAVERSE_PACKET) != 0);
  unpack_hits_kernel<<<(n + 255) / 256, 256, 0, st>>>(hits, s->sorted_perm, n, d_hit_id, d_hit_t);
  RT_CUDA(cudaGetLastError());
  return RT_OK;
}
