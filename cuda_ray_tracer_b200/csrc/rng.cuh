// rng.cuh — device XORWOW, bit-compatible with the cuRAND calls the reference makes:
// curand_init (draw.cu:105,162), curand_uniform via randD (helper.cu:82-85), curand_normal via
// standerdD (helper.cu:87-89).  Arithmetic follows curand_kernel.h / curand_uniform.h:69-72 /
// curand_normal.h:70-87,313-326 (device branch: __sincosf, logf, sqrtf).  Built with -fmad=false:
// x * 2^-32 + 2^-33 is a multiply and an add, as in the reference's parity build.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define RT_RNG_NIB_WORDS (40 * 16 * 5)   // 4-bit lookup table of one skip-ahead matrix (xw_matvec_nib), words
#define RT_RNG_NIB_STRIDE (RT_RNG_NIB_WORDS + 8)  // + the constant vector of xw_matvec_nib_lowseed, padded to 32 bytes

struct Xorwow {
  uint32_t v0, v1, v2, v3, v4, d;
  uint32_t bm_flag;   // EXTRA_FLAG_NORMAL = 1 when a cached Box-Muller value is pending
  float bm_extra;
};

__device__ __forceinline__ uint32_t xw_next(Xorwow& s) {
  uint32_t t = (s.v0 ^ (s.v0 >> 2));
  s.v0 = s.v1; s.v1 = s.v2; s.v2 = s.v3; s.v3 = s.v4;
  s.v4 = (s.v4 ^ (s.v4 << 4)) ^ (t ^ (t << 1));
  s.d += 362437u;
  return s.v4 + s.d;
}
__device__ __forceinline__ float xw_uniform(Xorwow& s) {
  uint32_t x = xw_next(s);
  return x * 2.3283064e-10f + (2.3283064e-10f / 2.0f);
}
__device__ __forceinline__ float xw_normal(Xorwow& s) {
  if (s.bm_flag != 1u) {
    uint32_t x = xw_next(s), y = xw_next(s);
    float u = x * 2.3283064e-10f + (2.3283064e-10f / 2);
    float v = y * (2.3283064e-10f * 6.2831855f) + ((2.3283064e-10f * 6.2831855f) / 2);
    float sq = sqrtf(-2.0f * logf(u));
    float sn, cs;
    __sincosf(v, &sn, &cs);
    sn *= sq; cs *= sq;
    s.bm_extra = cs; s.bm_flag = 1u;
    return sn;
  }
  s.bm_flag = 0u;
  return s.bm_extra;
}
// randD(start, end), helper.cu:82-85
__device__ __forceinline__ float xw_rand(float a, float b, Xorwow& s) { float u = xw_uniform(s); return a + (b - a) * u; }

// seed scramble of _curand_init_inplace (curand_kernel.h)
__device__ __forceinline__ void xw_seed(uint64_t seed, uint32_t* v, uint32_t& d) {
  uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
  uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
  uint32_t t0 = 1099087573u * s0;
  uint32_t t1 = 2591861531u * s1;
  d = 6615241u + t1 + t0;
  v[0] = 123456789u + t0; v[1] = 362436069u ^ t0; v[2] = 521288629u + t1; v[3] = 88675123u ^ t1; v[4] = 5783321u + t0;
}

// out = M * in over GF(2); M = 160 rows of 8 words (5 used), row b = image of basis vector b.
// One subsequence skip of any length is ONE such product with a host-built power matrix
// (rng_tables.cpp); all lanes of a warp normally share M, so the row loads are broadcasts.
__device__ __forceinline__ void xw_matvec(const uint32_t* __restrict__ M, const uint32_t* in, uint32_t* out) {
  uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0, r4 = 0;
#pragma unroll
  for (int i = 0; i < 5; i++) {
    const uint32_t w = in[i];
#pragma unroll 8
    for (int j = 0; j < 32; j++) {
      const uint4 a = __ldg(reinterpret_cast<const uint4*>(M + 8 * (i * 32 + j)));
      const uint32_t b = __ldg(M + 8 * (i * 32 + j) + 4);
      const uint32_t m = 0u - ((w >> j) & 1u);
      r0 ^= a.x & m; r1 ^= a.y & m; r2 ^= a.z & m; r3 ^= a.w & m; r4 ^= b & m;
    }
  }
  out[0] = r0; out[1] = r1; out[2] = r2; out[3] = r3; out[4] = r4;
}
// Seeds below 2^32 (1234 + pixel always is) leave state words v2, v3 of xw_seed independent of the seed, so their
// share of the product is one constant vector per matrix (tab[RT_RNG_NIB_WORDS .. +5), written by the table
// kernel) and only the 24 nibbles of v0, v1, v4 are looked up.
// ALL = false: only words 0, 1 and 4 of the result, which is all that the first two draws read (xw_next consumes
// v0 and v4, then what was v1): enough when nothing but the pixel jitter is ever drawn from this state.
template <bool ALL>
__device__ __forceinline__ void xw_matvec_nib_lowseed(const uint32_t* tab, const uint32_t* in, uint32_t* out) {
  uint32_t r0 = tab[RT_RNG_NIB_WORDS], r1 = tab[RT_RNG_NIB_WORDS + 1], r2 = tab[RT_RNG_NIB_WORDS + 2], r3 = tab[RT_RNG_NIB_WORDS + 3],
           r4 = tab[RT_RNG_NIB_WORDS + 4];
#pragma unroll
  for (int ii = 0; ii < 3; ii++) {
    const int i = ii == 2 ? 4 : ii;
    const uint32_t w = in[i];
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const uint32_t* e = tab + ((i * 8 + j) * 16 + ((w >> (4 * j)) & 15u)) * 5;
      r0 ^= e[0]; r1 ^= e[1]; if (ALL) { r2 ^= e[2]; r3 ^= e[3]; } r4 ^= e[4];
    }
  }
  out[0] = r0; out[1] = r1; out[2] = r2; out[3] = r3; out[4] = r4;
}

// 8-bit tables for the same low-seed product when only words 0, 1 and 4 of the result are wanted (the pixel jitter's two
// draws): T8[p][v] = T4[2p][v & 15] ^ T4[2p + 1][v >> 4] for the 12 byte positions of v0, v1, v4, three words per entry:
// 12 x 256 x 3 words = 36 KB per matrix, 12 lookups x (3 loads + 3 xors) instead of 24 x.  Entry stride 3 words (odd), so
// 32 lanes with different byte values spread over the banks.  [RT_RNG_BYTE_WORDS .. +3) = the seed-independent share.
#define RT_RNG_BYTE_WORDS (12 * 256 * 3)
#define RT_RNG_BYTE_STRIDE (RT_RNG_BYTE_WORDS + 4)
__device__ __forceinline__ void xw_matvec_byte_lowseed(const uint32_t* tab, const uint32_t* in, uint32_t* out) {
  uint32_t r0 = tab[RT_RNG_BYTE_WORDS], r1 = tab[RT_RNG_BYTE_WORDS + 1], r4 = tab[RT_RNG_BYTE_WORDS + 2];
#pragma unroll
  for (int ii = 0; ii < 3; ii++) {
    const uint32_t w = in[ii == 2 ? 4 : ii];
#pragma unroll
    for (int b = 0; b < 4; b++) {
      const uint32_t* e = tab + ((ii * 4 + b) * 256 + ((w >> (8 * b)) & 255u)) * 3;
      r0 ^= e[0]; r1 ^= e[1]; r4 ^= e[2];
    }
  }
  out[0] = r0; out[1] = r1; out[2] = 0u; out[3] = 0u; out[4] = r4;
}

// The same product through 4-bit tables (method of the four Russians): T[p][v] = XOR of the rows 4p+b, b in v,
// 40 nibble positions x 16 values x 5 words = 12.8 KB per matrix, staged in shared memory by the caller.  An
// entry is 5 words, so the 16 entries of a position start in 16 different banks: lanes with different nibble
// values do not conflict, lanes with equal values broadcast.  40 x (5 loads + 5 xors) instead of 160 x.
__device__ __forceinline__ void xw_matvec_nib(const uint32_t* tab, const uint32_t* in, uint32_t* out) {
  uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0, r4 = 0;
#pragma unroll
  for (int i = 0; i < 5; i++) {
    const uint32_t w = in[i];
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const uint32_t* e = tab + ((i * 8 + j) * 16 + ((w >> (4 * j)) & 15u)) * 5;
      r0 ^= e[0]; r1 ^= e[1]; r2 ^= e[2]; r3 ^= e[3]; r4 ^= e[4];
    }
  }
  out[0] = r0; out[1] = r1; out[2] = r2; out[3] = r3; out[4] = r4;
}
