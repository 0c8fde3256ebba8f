// rng_tables.cpp — host side of the XORWOW seeding used by the render kernels.
//
// The reference seeds one cuRAND XORWOW state per thread: curand_init(1234, tid, 0) in
// render_kernel (src/draw.cu:105) and curand_init(1234 + pixel, lane, 0) in render_kernel_warp_aa
// (:162).  curand_init skips `subsequence * 2^67` draws ahead by multiplying the 160-bit state with
// precomputed GF(2) matrices, one full matrix-vector product per non-zero base-4 digit value of the
// subsequence (curand_kernel.h, _skipahead_sequence_inplace) — thousands of loads per thread.
// Here the powers P^m of the subsequence-step matrix P = T^(2^67) are built once on the host by
// repeated squaring of the one-draw transition matrix T, so that the device does exactly ONE
// warp-uniform matrix-vector product per sample (rng.cuh), with bit-identical results.
#include <cstring>
#include <mutex>
#include <vector>

#include "rng_tables.h"

namespace {

inline uint32_t xw_step(uint32_t* v) {  // curand(curandStateXORWOW_t*), without the Weyl term d
  uint32_t t = (v[0] ^ (v[0] >> 2));
  v[0] = v[1]; v[1] = v[2]; v[2] = v[3]; v[3] = v[4];
  v[4] = (v[4] ^ (v[4] << 4)) ^ (t ^ (t << 1));
  return v[4];
}
struct Gf2 { uint32_t img[160][5]; };
inline void apply(const Gf2& M, const uint32_t* in, uint32_t* out) {
  uint32_t r[5] = {0, 0, 0, 0, 0};
  for (int i = 0; i < 5; i++)
    for (uint32_t bits = in[i]; bits; bits &= bits - 1) {
      const uint32_t* im = M.img[i * 32 + __builtin_ctz(bits)];
      for (int k = 0; k < 5; k++) r[k] ^= im[k];
    }
  memcpy(out, r, sizeof(r));
}
inline Gf2 mul(const Gf2& A, const Gf2& B) {  // (A after B)
  Gf2 R;
  for (int b = 0; b < 160; b++) apply(A, B.img[b], R.img[b]);
  return R;
}
const Gf2& subsequence_step() {  // P = T^(2^67)
  static Gf2 P; static std::once_flag once;
  std::call_once(once, [] {
    Gf2 T;
    for (int b = 0; b < 160; b++) {
      uint32_t v[5] = {0, 0, 0, 0, 0};
      v[b / 32] = 1u << (b % 32);
      xw_step(v);
      memcpy(T.img[b], v, sizeof(v));
    }
    P = T;
    for (int i = 0; i < 67; i++) P = mul(P, P);
  });
  return P;
}
Gf2 identity() {
  Gf2 I; memset(&I, 0, sizeof(I));
  for (int b = 0; b < 160; b++) I.img[b][b / 32] = 1u << (b % 32);
  return I;
}
Gf2 power(const Gf2& base, unsigned long long e) {
  Gf2 R = identity(), B = base;
  while (e) { if (e & 1) R = mul(B, R); e >>= 1; if (e) B = mul(B, B); }
  return R;
}
void pack(const Gf2& M, uint32_t* dst) {  // device layout: 160 rows of 8 words (5 used), 32-byte aligned
  for (int b = 0; b < 160; b++) { memcpy(dst + 8 * b, M.img[b], 20); dst[8 * b + 5] = dst[8 * b + 6] = dst[8 * b + 7] = 0; }
}

}  // namespace

void rt_xorwow_seed(uint64_t seed, uint32_t* v5, uint32_t* d) {  // _curand_init_inplace, scramble part
  uint32_t s0 = ((uint32_t)seed) ^ 0xaad26b49u;
  uint32_t s1 = (uint32_t)(seed >> 32) ^ 0xf7dcefddu;
  uint32_t t0 = 1099087573u * s0;
  uint32_t t1 = 2591861531u * s1;
  *d = 6615241u + t1 + t0;
  v5[0] = 123456789u + t0; v5[1] = 362436069u ^ t0; v5[2] = 521288629u + t1; v5[3] = 88675123u ^ t1; v5[4] = 5783321u + t0;
}

// matrices P^0 .. P^(count-1), packed for the device
void rt_xorwow_power_matrices(int count, std::vector<uint32_t>& out) {
  out.assign((size_t)count * RT_RNG_MATRIX_WORDS, 0u);
  const Gf2& P = subsequence_step();
  Gf2 M = identity();
  for (int m = 0; m < count; m++) { pack(M, out.data() + (size_t)m * RT_RNG_MATRIX_WORDS); if (m + 1 < count) M = mul(P, M); }
}

// base[h] = P^(stride*h) * v(seed) for h in [0,count): the per-256-pixel starting vectors of the
// one-sample kernel (seed fixed, subsequence = pixel index)
void rt_xorwow_base_vectors(uint64_t seed, int stride, int count, std::vector<uint32_t>& out) {
  out.assign((size_t)count * 8, 0u);
  uint32_t v[5], d; rt_xorwow_seed(seed, v, &d);
  const Gf2 Q = power(subsequence_step(), (unsigned long long)stride);
  for (int h = 0; h < count; h++) { memcpy(out.data() + 8 * (size_t)h, v, 20); apply(Q, v, v); }
}

extern "C" void rt_xorwow_init(uint64_t seed, uint64_t subsequence, uint32_t* state6) {
  uint32_t v[5], d; rt_xorwow_seed(seed, v, &d);
  if (subsequence) { Gf2 M = power(subsequence_step(), subsequence); apply(M, v, v); }
  memcpy(state6, v, 20); state6[5] = d;  // d is unchanged: the skip is a multiple of 2^32 draws
}
