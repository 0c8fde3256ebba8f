// build_sort.cu — scene bounds, Morton codes (K1) and a hand-written onesweep radix sort (K2).
//
// Replaces generate_morton_codes_kernel (src/lbvh_utils.cu:32-75) and the thrust::sort_by_key call
// (src/lbvh_utils.cu:110-115).  The sort is a least-significant-digit radix sort with 8-bit digits in
// which every pass is ONE kernel ("onesweep", Adinets & Merrill 2022): tiles are handed out by an
// atomic ticket, each tile ranks its keys with warp match operations, publishes its per-digit counts
// and obtains its global offsets by decoupled look-back over the preceding tiles.  All digit
// histograms are produced by the Morton kernel itself, so the keys are read exactly once per pass.
// Stable by construction (tile order = memory order, in-tile order = memory order), which the build
// depends on: with the reference's shipped bounds every code is 0 and the order must stay the file
// order.  The payload is the 4-byte file-order primitive index.
#include "build.cuh"

namespace {

constexpr int SORT_THREADS = 256;
// keys per thread: 16 (4096-key tiles).  A pass over 1M keys is only 245 tiles on 148 SMs and bound by the latency of one
// tile's chain (ncu: 20 % of the warp slots, 14 % issue), so 2048-key tiles (RT_SORT_SMALL_N = the key count up to which
// they are used) were measured: 0.336 ms per build at 1M triangles against 0.328 with 4096-key tiles, 0.490 against 0.469
// with 63-bit keys (profiles/build_ab_r02.txt) — the longer look-back chain costs what the shorter tile saves.  Off.
#ifndef RT_SORT_SMALL_N
#define RT_SORT_SMALL_N 0
#endif
#ifndef RT_SORT_TMA
#define RT_SORT_TMA 0   // experiment: cp.async.bulk staging of the tile (see onesweep_kernel)
#endif
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr uint32_t LB_AGG = 1u << 30, LB_INC = 2u << 30, LB_VAL = (1u << 30) - 1;
#ifndef RT_HIST_HYBRID
#define RT_HIST_HYBRID 1  // 0: every digit through __match_any_sync (round 1), for A/B
#endif
#ifndef RT_RANK_BATCH
#define RT_RANK_BATCH 4  // 1: one match, one update at a time (round 1); measured equal
#endif
constexpr int RANK_BATCH = RT_RANK_BATCH;
#ifndef RT_LB_WINDOW
#define RT_LB_WINDOW 8
#endif
constexpr int LB_WINDOW = RT_LB_WINDOW;

// ---- state init: one launch instead of four memsets -------------------------------------------
__global__ void init_build_state_kernel(uint32_t* hist, int hist_words, uint32_t* lookback, size_t lb_words, uint32_t* tickets,
                                        uint32_t* bounds_enc, float3 mn, float3 mx) {
  rt_pdl_trigger();
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  for (size_t k = i; k < (size_t)hist_words; k += stride) hist[k] = 0;
  for (size_t k = i; k < lb_words; k += stride) lookback[k] = 0;
  if (i < 64) tickets[i] = 0;
  if (i == 0) {
    bounds_enc[0] = rt_float_order(mn.x); bounds_enc[1] = rt_float_order(mn.y); bounds_enc[2] = rt_float_order(mn.z);
    bounds_enc[3] = rt_float_order(mx.x); bounds_enc[4] = rt_float_order(mx.y); bounds_enc[5] = rt_float_order(mx.z);
  }
}

// ---- scene bounds: fminf/fmaxf over the primitive AABBs (parse.cpp:147-155,193-200) -----------
__global__ void __launch_bounds__(256) bounds_kernel(const float4* __restrict__ bv, int n, uint32_t* bounds_enc) {
  rt_pdl_trigger(); rt_pdl_wait();
  float mnx = INFINITY, mny = INFINITY, mnz = INFINITY, mxx = -INFINITY, mxy = -INFINITY, mxz = -INFINITY;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float4 a = __ldg(bv + RT_PRIM_STRIDE * (size_t)i), b = __ldg(bv + RT_PRIM_STRIDE * (size_t)i + 1), c = __ldg(bv + RT_PRIM_STRIDE * (size_t)i + 2);
    RtBox bx = rt_prim_box(a, b, c);
    mnx = fminf(mnx, bx.xmin); mny = fminf(mny, bx.ymin); mnz = fminf(mnz, bx.zmin);
    mxx = fmaxf(mxx, bx.xmax); mxy = fmaxf(mxy, bx.ymax); mxz = fmaxf(mxz, bx.zmax);
  }
  for (int o = 16; o > 0; o >>= 1) {
    mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o)); mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o));
    mnz = fminf(mnz, __shfl_xor_sync(0xffffffffu, mnz, o)); mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
    mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o)); mxz = fmaxf(mxz, __shfl_xor_sync(0xffffffffu, mxz, o));
  }
  __shared__ float red[6][8];
  int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { red[0][w] = mnx; red[1][w] = mny; red[2][w] = mnz; red[3][w] = mxx; red[4][w] = mxy; red[5][w] = mxz; }
  __syncthreads();
  if (threadIdx.x < 6) {
    float v = red[threadIdx.x][0];
    for (int k = 1; k < 8; k++) v = threadIdx.x < 3 ? fminf(v, red[threadIdx.x][k]) : fmaxf(v, red[threadIdx.x][k]);
    if (threadIdx.x < 3) atomicMin(bounds_enc + threadIdx.x, rt_float_order(v));
    else atomicMax(bounds_enc + threadIdx.x, rt_float_order(v));
  }
}

// ---- K1: Morton codes -------------------------------------------------------------------------
__device__ inline uint32_t expand_bits10(uint32_t v) {  // lbvh_utils.cu:10-17
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
__device__ inline uint64_t expand_bits21(uint64_t v) {
  v &= 0x1FFFFFull;
  v = (v | (v << 32)) & 0x1F00000000FFFFull;
  v = (v | (v << 16)) & 0x1F0000FF0000FFull;
  v = (v | (v << 8)) & 0x100F00F00F00F00Full;
  v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
  v = (v | (v << 2)) & 0x1249249249249249ull;
  return v;
}
__device__ inline uint32_t quantize(float c, float mn, float range, float scale) {  // lbvh_utils.cu:24-30
  if (range <= 1e-6f) return 0u;
  float nrm = (c - mn) / range;
  nrm = fmaxf(0.0f, fminf(1.0f, nrm));
  return __float2uint_rz(nrm * scale);
}

// lanes of the (full) warp whose 8-bit digit equals mine.  MATCH.ANY resolves one distinct value after the other (about 30
// of them among 32 random digits) and sat on 30 % of the pass's stall samples; eight ballots, one per bit, are independent
// of each other and pipeline.  RT_RANK_BALLOT=0: the match instruction (round 1), for A/B.
#ifndef RT_RANK_BALLOT
#define RT_RANK_BALLOT 1
#endif
__device__ __forceinline__ unsigned rt_match_digit8(unsigned d) {
#if RT_RANK_BALLOT
  unsigned peers = 0xffffffffu;
#pragma unroll
  for (int b = 0; b < 8; b++) {
    const bool bit = (d >> b) & 1u;
    const unsigned m = __ballot_sync(0xffffffffu, bit);
    peers &= bit ? m : ~m;
  }
  return peers;
#else
  return __match_any_sync(0xffffffffu, d);
#endif
}

template <typename KeyT, int NPASS>
__global__ void __launch_bounds__(256) morton_kernel(const float4* __restrict__ bv, int n, const uint32_t* __restrict__ bounds_enc,
                                                    KeyT* __restrict__ keys, uint32_t* __restrict__ perm, uint32_t* __restrict__ hist,
                                                    int* __restrict__ node_flags) {
  __shared__ uint32_t s_hist[NPASS][256];
  rt_pdl_trigger();
  for (int k = threadIdx.x; k < NPASS * 256; k += blockDim.x) (&s_hist[0][0])[k] = 0;
  rt_pdl_wait();  // the bounds (and the zeroed histograms) come from the kernels before
  __syncthreads();
  const float mnx = rt_float_unorder(bounds_enc[0]), mny = rt_float_unorder(bounds_enc[1]), mnz = rt_float_unorder(bounds_enc[2]);
  const float rx = rt_float_unorder(bounds_enc[3]) - mnx, ry = rt_float_unorder(bounds_enc[4]) - mny,
              rz = rt_float_unorder(bounds_enc[5]) - mnz;  // lbvh_utils.cu:66-68
  const unsigned lane = threadIdx.x & 31;
  const int n_round = (n + 31) & ~31;  // whole warps stay together for the match operations
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += gridDim.x * blockDim.x) {
    const bool valid = i < n;
    KeyT code = 0;
    if (valid) {
      float4 a = __ldg(bv + RT_PRIM_STRIDE * (size_t)i), b = __ldg(bv + RT_PRIM_STRIDE * (size_t)i + 1), c = __ldg(bv + RT_PRIM_STRIDE * (size_t)i + 2);
      float cx, cy, cz;
      if (c.w == RT_TAG_SPHERE) { cx = a.x; cy = a.y; cz = a.z; }  // lbvh_utils.cu:48-51
      else {  // (p0 + p1 + p2) / 3.0f, lbvh_utils.cu:55-58
        cx = ((a.x + a.w) + b.z) / 3.0f; cy = ((a.y + b.x) + b.w) / 3.0f; cz = ((a.z + b.y) + c.x) / 3.0f;
      }
      if (sizeof(KeyT) == 4) {
        uint32_t qx = quantize(cx, mnx, rx, 1023.0f), qy = quantize(cy, mny, ry, 1023.0f), qz = quantize(cz, mnz, rz, 1023.0f);
        code = (KeyT)(expand_bits10(qx) | (expand_bits10(qy) << 1) | (expand_bits10(qz) << 2));  // lbvh_utils.cu:19-22
      } else {
        uint64_t qx = quantize(cx, mnx, rx, 2097151.0f), qy = quantize(cy, mny, ry, 2097151.0f), qz = quantize(cz, mnz, rz, 2097151.0f);
        code = (KeyT)(expand_bits21(qx) | (expand_bits21(qy) << 1) | (expand_bits21(qz) << 2));
      }
      keys[i] = code; perm[i] = (uint32_t)i; node_flags[i] = -1;
    }
    const unsigned act = __ballot_sync(0xffffffffu, valid);
    if (act == 0u) continue;
    const int first = __ffs(act) - 1;
#pragma unroll
    for (int p = 0; p < NPASS; p++) {
      unsigned d = valid ? (unsigned)((code >> (8 * p)) & 0xFF) : 256u + lane;  // invalid lanes match nobody
      // Primitives in no particular order (the soup) have unrelated digits: 32 plain shared atomics, hardly a bank conflict,
      // instead of one __match_any_sync round per digit (ncu r02: short_scoreboard 25 per issue, 37 of the build's 300 us).
      // Digits that many lanes share - a mesh in file order, or every code equal as in the shipped reference - would
      // serialise those atomics, so a warp whose first lane's digit is shared by more than 8 lanes aggregates first.
      const unsigned d0 = __shfl_sync(0xffffffffu, d, first);  // every lane takes part: not inside the && below
      const unsigned same = RT_HIST_HYBRID ? __ballot_sync(0xffffffffu, valid && d == d0) : 0xffffffffu;
      if (__popc(same) > 8) {
        const unsigned peers = rt_match_digit8(d) & act;  // (an invalid lane's low 8 bits may equal a valid digit: & act)
        if (valid && lane == (unsigned)(__ffs(peers) - 1)) atomicAdd(&s_hist[p][d], (uint32_t)__popc(peers));
      } else if (valid) atomicAdd(&s_hist[p][d], 1u);
    }
  }
  __syncthreads();
  for (int k = threadIdx.x; k < NPASS * 256; k += blockDim.x) {
    uint32_t v = (&s_hist[0][0])[k];
    if (v) atomicAdd(hist + k, v);
  }
}

// ---- K2: one onesweep pass --------------------------------------------------------------------
__device__ inline uint32_t ld_volatile(const uint32_t* p) { return *(const volatile uint32_t*)p; }
__device__ inline void st_volatile(uint32_t* p, uint32_t v) { *(volatile uint32_t*)p = v; }

template <typename KeyT, int SORT_ITEMS>
__global__ void __launch_bounds__(SORT_THREADS) onesweep_kernel(const KeyT* __restrict__ kin, KeyT* __restrict__ kout,
                                                               const uint32_t* __restrict__ vin, uint32_t* __restrict__ vout, int n,
                                                               int shift, const uint32_t* __restrict__ hist, uint32_t* lookback,
                                                               uint32_t* ticket) {
  __shared__ uint32_t s_warp[SORT_WARPS][256];  // per-warp digit counts, then exclusive warp offsets
  __shared__ uint32_t s_gbase[256];             // global position of local slot 0 of each digit run
  __shared__ uint32_t s_lbase[256];             // local exclusive digit offsets
  __shared__ uint32_t s_scan[2][SORT_WARPS];
  __shared__ uint32_t s_tile;
  constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;
  extern __shared__ __align__(16) unsigned char s_raw[];
  KeyT* s_keys = reinterpret_cast<KeyT*>(s_raw);
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_raw);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  rt_pdl_trigger();
  for (int k = tid; k < SORT_WARPS * 256; k += SORT_THREADS) (&s_warp[0][0])[k] = 0;
  rt_pdl_wait();  // the previous pass (or the Morton kernel) must be complete before a ticket is taken
  if (tid == 0) s_tile = atomicAdd(ticket, 1u);
  // a digit every key shares makes the pass the identity: copy the tile and leave
  const int trivial = __syncthreads_or(hist[tid] == (uint32_t)n);
  const uint32_t tile = s_tile;
  const size_t base = (size_t)tile * SORT_TILE;
  if (trivial) {
#pragma unroll
    for (int j = 0; j < SORT_ITEMS; j++) {
      size_t idx = base + (size_t)j * SORT_THREADS + tid;
      if (idx < (size_t)n) { kout[idx] = kin[idx]; vout[idx] = vin[idx]; }
    }
    return;
  }
  const int tile_valid = (int)min((size_t)SORT_TILE, (size_t)n - base);

  // warp-striped load: warp w owns keys [base + w*512, +512), item j of lane l is offset j*32 + l
  KeyT key[SORT_ITEMS];
  uint32_t val[SORT_ITEMS], pos[SORT_ITEMS];
  const size_t wbase = base + (size_t)warp * (32 * SORT_ITEMS);
#if RT_SORT_TMA
  // EXPERIMENT (BASELINE.json: "TMA bulk staging on these streaming passes"), off by default: a full tile's keys and
  // values come in with two cp.async.bulk copies (UBLKCP) that one thread issues and an mbarrier signals; the lanes
  // then pick their items from shared memory (SASS: UBLKCP.S.G + SYNCS).  Measured against the plain coalesced loads
  // below (profiles/build_ab_r02.txt): 0.336 vs 0.336 ms per build at 1M triangles, 2.528 vs 2.494 ms at 10M — the pass is
  // bound by the ranking chain and the look-back, not by how the 32 KB tile arrives.
  if (tile_valid == SORT_TILE && SORT_TILE * (sizeof(KeyT) + 4) <= 40 * 1024) {  // staging must fit beside the 9 KB of static shared memory
    __shared__ __align__(8) unsigned long long s_bar;
    KeyT* st_k = reinterpret_cast<KeyT*>(s_raw);
    uint32_t* st_v = reinterpret_cast<uint32_t*>(s_raw + SORT_TILE * sizeof(KeyT));
    const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar);
    if (tid == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (tid == 0) {
      const unsigned bk = SORT_TILE * (unsigned)sizeof(KeyT), bvb = SORT_TILE * 4u;
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bk + bvb) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"((unsigned)__cvta_generic_to_shared(st_k)),
                   "l"(kin + base), "r"(bk), "r"(bar) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"((unsigned)__cvta_generic_to_shared(st_v)),
                   "l"(vin + base), "r"(bvb), "r"(bar) : "memory");
    }
    unsigned done = 0;
    while (!done)
      asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(bar) : "memory");
#pragma unroll
    for (int j = 0; j < SORT_ITEMS; j++) {
      const int li = warp * (32 * SORT_ITEMS) + j * 32 + lane;
      key[j] = st_k[li]; val[j] = st_v[li];
    }
    __syncthreads();  // the staging area is the scatter area below
  } else
#endif
#pragma unroll
  for (int j = 0; j < SORT_ITEMS; j++) {
    size_t idx = wbase + (size_t)j * 32 + lane;
    bool valid = idx < (size_t)n;
    key[j] = valid ? kin[idx] : ~(KeyT)0;
    val[j] = valid ? vin[idx] : 0u;
  }
  // rank inside the warp, item by item (memory order), with match_any
  // The matches do not depend on the counters, only the counter updates depend on each other: RANK_BATCH matches go out
  // back to back before their updates (ncu r02: 30 % of the pass's stall samples sat on the first use of one match's
  // result, 16 of them in a row).  MEASURED: batches of 1, 4 and 8 build in the same 0.266 ms at 1M (profiles/
  // build_sort_r02.txt) - MATCH.ANY's own rate bounds the chain, not the order it is issued in.  Kept at 4 (tested), no gain.
  const unsigned lt = (1u << lane) - 1u;
#pragma unroll
  for (int j0 = 0; j0 < SORT_ITEMS; j0 += RANK_BATCH) {
    unsigned dg[RANK_BATCH], peers[RANK_BATCH];
#pragma unroll
    for (int b = 0; b < RANK_BATCH; b++) {
      dg[b] = (unsigned)((key[j0 + b] >> shift) & 0xFF);
      peers[b] = rt_match_digit8(dg[b]);
    }
#pragma unroll
    for (int b = 0; b < RANK_BATCH; b++) {
      const int leader = __ffs(peers[b]) - 1;
      uint32_t prev = 0;
      if (lane == leader) { prev = s_warp[warp][dg[b]]; s_warp[warp][dg[b]] = prev + (uint32_t)__popc(peers[b]); }
      prev = __shfl_sync(0xffffffffu, prev, leader);
      pos[j0 + b] = prev + (uint32_t)__popc(peers[b] & lt);
      __syncwarp();
    }
  }
  __syncthreads();

  // thread d owns digit d: warp offsets, tile count, look-back
  {
    const int d = tid;
    uint32_t cnt = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; w++) { uint32_t c = s_warp[w][d]; s_warp[w][d] = cnt; cnt += c; }
    st_volatile(lookback + (size_t)tile * 256 + d, (tile == 0 ? LB_INC : LB_AGG) | cnt);
    // two block-wide exclusive scans over digits: tile counts (local bases) and the global histogram
    const uint32_t h = hist[d];
    uint32_t a = cnt, b = h;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t ta = __shfl_up_sync(0xffffffffu, a, o), tb = __shfl_up_sync(0xffffffffu, b, o);
      if (lane >= o) { a += ta; b += tb; }
    }
    if (lane == 31) { s_scan[0][warp] = a; s_scan[1][warp] = b; }
    __syncthreads();
    uint32_t oa = 0, ob = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; w++) if (w < warp) { oa += s_scan[0][w]; ob += s_scan[1][w]; }
    const uint32_t lbase = oa + a - cnt, gdigit = ob + b - h;
    uint32_t excl = 0;
    if (tile > 0) {
      // LB_WINDOW predecessors are read at once and then consumed in order: all tiles of a 1M sort are resident together,
      // publish their aggregates at about the same time and then walk back until they meet an inclusive prefix; one
      // dependent L2 read per predecessor made that walk the longest chain of the pass
      int t = (int)tile - 1;
      bool done = false;
      while (!done) {
        uint32_t v[LB_WINDOW];
#pragma unroll
        for (int k = 0; k < LB_WINDOW; k++) v[k] = t - k >= 0 ? ld_volatile(lookback + (size_t)(t - k) * 256 + d) : LB_INC;
#pragma unroll
        for (int k = 0; k < LB_WINDOW; k++) {
          if (done || (v[k] >> 30) == 0) break;  // not published yet: read again from here
          excl += v[k] & LB_VAL;
          t--;
          done = (v[k] & LB_INC) != 0u;
        }
      }
      st_volatile(lookback + (size_t)tile * 256 + d, LB_INC | (excl + cnt));
    }
    s_lbase[d] = lbase;
    s_gbase[d] = gdigit + excl - lbase;
  }
  __syncthreads();

  // local scatter through shared memory, then coalesced runs to global
#pragma unroll
  for (int j = 0; j < SORT_ITEMS; j++) {
    const unsigned d = (unsigned)((key[j] >> shift) & 0xFF);
    pos[j] += s_lbase[d] + s_warp[warp][d];
    s_keys[pos[j]] = key[j];
  }
  __syncthreads();
  uint32_t gpos[SORT_ITEMS];
#pragma unroll
  for (int j = 0; j < SORT_ITEMS; j++) {
    const int idx = j * SORT_THREADS + tid;
    if (idx < tile_valid) {
      const KeyT k = s_keys[idx];
      gpos[j] = s_gbase[(unsigned)((k >> shift) & 0xFF)] + (uint32_t)idx;
      kout[gpos[j]] = k;
    }
  }
  __syncthreads();
#pragma unroll
  for (int j = 0; j < SORT_ITEMS; j++) s_vals[pos[j]] = val[j];
  __syncthreads();
#pragma unroll
  for (int j = 0; j < SORT_ITEMS; j++) {
    const int idx = j * SORT_THREADS + tid;
    if (idx < tile_valid) vout[gpos[j]] = s_vals[idx];
  }
}

template <typename KeyT, int NPASS, int ITEMS>
int sort_impl(rt_scene* s, cudaStream_t st, uint32_t* launches) {
  const int n = s->n;
  constexpr int TILE = SORT_THREADS * ITEMS;
  const int tiles = (n + TILE - 1) / TILE;
  KeyT *ka = (KeyT*)s->keys_a, *kb = (KeyT*)s->keys_b;
  uint32_t *pa = s->perm_a, *pb = s->perm_b;
  int grid = (n + 255) / 256;
  if (grid > 148 * 8) grid = 148 * 8;
  RT_CUDA(rt_launch_chained(rt_pdl_for(n), morton_kernel<KeyT, NPASS>, dim3(grid), dim3(256), 0, st, s->prim_bv, n, (const uint32_t*)s->bounds, ka, pa, s->hist, s->node_flags));
  (*launches)++;
  for (int p = 0; p < NPASS; p++) {
    RT_CUDA(rt_launch_chained(rt_pdl_for(n), onesweep_kernel<KeyT, ITEMS>, dim3(tiles), dim3(SORT_THREADS), TILE * (sizeof(KeyT) + ((RT_SORT_TMA && TILE * (sizeof(KeyT) + 4) <= 40 * 1024) ? 4 : 0)), st,
                              (const KeyT*)ka, kb, (const uint32_t*)pa, pb, n, 8 * p, (const uint32_t*)(s->hist + 256 * p), s->lookback + (size_t)p * tiles * 256, s->tickets + p));
    (*launches)++;
    KeyT* tk = ka; ka = kb; kb = tk;
    uint32_t* tp = pa; pa = pb; pb = tp;
  }
  RT_CUDA(cudaGetLastError());
  if (sizeof(KeyT) == 4) { s->sorted_keys32 = (uint32_t*)ka; s->sorted_keys64 = nullptr; }
  else { s->sorted_keys64 = (uint64_t*)ka; s->sorted_keys32 = nullptr; }
  s->sorted_perm = pa;
  return RT_OK;
}

}  // namespace

int rt_build_codes_and_sort(rt_scene* s, int morton_bits, unsigned flags, cudaStream_t st, uint32_t* launches) {
  const int n = s->n;
  const int npass = morton_bits == 30 ? 4 : 8;
  const bool small = n <= RT_SORT_SMALL_N;
  const int tile_keys = SORT_THREADS * (small ? 8 : 16);
  const int tiles = (n + tile_keys - 1) / tile_keys;
  float3 mn = make_float3(s->cfg.scene_min[0], s->cfg.scene_min[1], s->cfg.scene_min[2]);
  float3 mx = make_float3(s->cfg.scene_max[0], s->cfg.scene_max[1], s->cfg.scene_max[2]);
  if (flags & RT_BOUNDS_COMPUTED) { mn = make_float3(INFINITY, INFINITY, INFINITY); mx = make_float3(-INFINITY, -INFINITY, -INFINITY); }
  const size_t lb_words = (size_t)npass * tiles * 256;
  int g0 = (int)((lb_words + 255) / 256); if (g0 < 8) g0 = 8; if (g0 > 148 * 4) g0 = 148 * 4;
  init_build_state_kernel<<<g0, 256, 0, st>>>(s->hist, 8 * 256, s->lookback, lb_words, s->tickets, (uint32_t*)s->bounds, mn, mx);
  (*launches)++;
  if (flags & RT_BOUNDS_COMPUTED) {
    int g = (n + 255) / 256; if (g > 148 * 8) g = 148 * 8;
    RT_CUDA(rt_launch_chained(rt_pdl_for(n), bounds_kernel, dim3(g), dim3(256), 0, st, (const float4*)s->prim_bv, n, (uint32_t*)s->bounds));
    (*launches)++;
  }
  RT_CUDA(cudaGetLastError());
  if (morton_bits == 30) return small ? sort_impl<uint32_t, 4, 8>(s, st, launches) : sort_impl<uint32_t, 4, 16>(s, st, launches);
  return small ? sort_impl<uint64_t, 8, 8>(s, st, launches) : sort_impl<uint64_t, 8, 16>(s, st, launches);
}
