// Which pipe does what on sm_100a?  Issue-rate microbenchmark: N independent chains of one PTX op (or an
// interleaved pair) per thread, all SMs full; reports warp-instructions per cycle per SM sub-partition.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 2048
#define CHAINS 8

#define KERNEL(name, decl, body_a, body_b, fin)                                             \
  __global__ void name(float* out, float s, float t) {                                      \
    decl;                                                                                   \
    for (int it = 0; it < ITERS; it++) {                                                    \
      _Pragma("unroll") for (int c = 0; c < CHAINS; c++) { body_a; body_b; }                \
    }                                                                                       \
    float acc = 0.f;                                                                        \
    _Pragma("unroll") for (int c = 0; c < CHAINS; c++) { fin; }                             \
    if (acc == 12345.678f) out[0] = acc;                                                    \
  }

#define DECLF float x[CHAINS], y[CHAINS]; for (int c = 0; c < CHAINS; c++) { x[c] = s + c + threadIdx.x; y[c] = t - c; }
#define DECLU unsigned x[CHAINS], y[CHAINS]; for (int c = 0; c < CHAINS; c++) { x[c] = __float_as_uint(s) + c + threadIdx.x; y[c] = __float_as_uint(t) - c; }
#define DECL2 unsigned long long x[CHAINS], y[CHAINS]; for (int c = 0; c < CHAINS; c++) { x[c] = (unsigned long long)__float_as_uint(s + c + threadIdx.x) << 32 | __float_as_uint(s); y[c] = (unsigned long long)__float_as_uint(t) << 32 | __float_as_uint(t - c); }
#define FINF acc += x[c] + y[c]
#define FINU acc += __uint_as_float(x[c] ^ y[c])
#define FIN2 acc += __uint_as_float((unsigned)(x[c] >> 32) ^ (unsigned)x[c] ^ (unsigned)y[c])
#define NOP

KERNEL(k_fadd, DECLF, asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x[c]) : "f"(t)), NOP, FINF)
KERNEL(k_fmul, DECLF, asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(x[c]) : "f"(t)), NOP, FINF)
KERNEL(k_ffma, DECLF, asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[c]) : "f"(t), "f"(s)), NOP, FINF)
KERNEL(k_fmnmx, DECLF, asm volatile("min.f32 %0, %0, %1;" : "+f"(x[c]) : "f"(y[c])), NOP, FINF)
KERNEL(k_fmnmx3, DECLF, asm volatile("min.f32 %0, %0, %1, %2;" : "+f"(x[c]) : "f"(y[c]), "f"(t)), NOP, FINF)
KERNEL(k_lop3, DECLU, asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[c]) : "r"(y[c]), "r"(__float_as_uint(t))), NOP, FINU)
KERNEL(k_iadd, DECLU, asm volatile("add.u32 %0, %0, %1;" : "+r"(x[c]) : "r"(y[c])), NOP, FINU)
KERNEL(k_imad, DECLU, asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[c]) : "r"(y[c]), "r"(__float_as_uint(t))), NOP, FINU)
KERNEL(k_selp, DECLF, asm volatile("{.reg .pred p; setp.lt.f32 p, %0, %1; selp.f32 %0, %1, %2, p;}" : "+f"(x[c]) : "f"(y[c]), "f"(t)), NOP, FINF)
KERNEL(k_fadd2, DECL2, asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(x[c]) : "l"(y[c])), NOP, FIN2)
KERNEL(k_fmul2, DECL2, asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(x[c]) : "l"(y[c])), NOP, FIN2)
KERNEL(k_ffma2, DECL2, asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(x[c]) : "l"(y[c])), NOP, FIN2)
KERNEL(k_fsetp, DECLF, asm volatile("{.reg .pred p; setp.lt.f32 p, %0, %1; @p add.rn.f32 %0, %0, %1;}" : "+f"(x[c]) : "f"(y[c])), NOP, FINF)
KERNEL(k_fsetp2, DECLF, asm volatile("{.reg .pred p, q; setp.lt.f32 p, %0, %1; setp.gt.and.f32 q, %0, %2, p; @q add.rn.f32 %0, %0, %1;}" : "+f"(x[c]) : "f"(y[c]), "f"(t)), NOP, FINF)
KERNEL(k_isetp, DECLU, asm volatile("{.reg .pred p; setp.lt.s32 p, %0, %1; @p add.u32 %0, %0, %1;}" : "+r"(x[c]) : "r"(y[c])), NOP, FINU)
KERNEL(k_sel, DECLU, asm volatile("{.reg .pred p; setp.lt.s32 p, %1, 0; selp.b32 %0, %0, %1, p;}" : "+r"(x[c]) : "r"(y[c])), NOP, FINU)
KERNEL(k_prmt, DECLU, asm volatile("prmt.b32 %0, %0, %1, 0x7610;" : "+r"(x[c]) : "r"(y[c])), NOP, FINU)
KERNEL(k_shf, DECLU, asm volatile("shf.l.wrap.b32 %0, %0, %1, 3;" : "+r"(x[c]) : "r"(y[c])), NOP, FINU)
KERNEL(k_fsel_fma, DECLF, asm volatile("{.reg .pred p; setp.lt.f32 p, %1, 0f00000000; @p fma.rn.f32 %0, %1, 0f3F800000, 0f00000000;}" : "+f"(x[c]) : "f"(y[c])), NOP, FINF)
// pairs: do they share a pipe?
KERNEL(k_fadd_fmnmx, DECLF, asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x[c]) : "f"(t)), asm volatile("min.f32 %0, %0, %1;" : "+f"(y[c]) : "f"(s)), FINF)
KERNEL(k_fmul_fmnmx, DECLF, asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(x[c]) : "f"(t)), asm volatile("min.f32 %0, %0, %1;" : "+f"(y[c]) : "f"(s)), FINF)
KERNEL(k_fadd_fmul, DECLF, asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(x[c]) : "f"(t)), asm volatile("mul.rn.f32 %0, %0, %1;" : "+f"(y[c]) : "f"(s)), FINF)
KERNEL(k_ffma_fmnmx, DECLF, asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[c]) : "f"(t), "f"(s)), asm volatile("min.f32 %0, %0, %1;" : "+f"(y[c]) : "f"(s)), FINF)
KERNEL(k_fadd2_fmnmx, DECL2, asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(x[c]) : "l"(y[c])), asm volatile("{.reg .f32 a, b; mov.b64 {a, b}, %0; min.f32 a, a, b; mov.b64 %0, {a, b};}" : "+l"(y[c])), FIN2)
KERNEL(k_lop3_fmnmx, DECLU, asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[c]) : "r"(y[c]), "r"(__float_as_uint(t))), asm volatile("{.reg .f32 a; mov.b32 a, %0; min.f32 a, a, 1.5; mov.b32 %0, a;}" : "+r"(y[c])), FINU)
KERNEL(k_imad_fmnmx, DECLU, asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[c]) : "r"(3u), "r"(__float_as_uint(t))), asm volatile("{.reg .f32 a; mov.b32 a, %0; min.f32 a, a, 1.5; mov.b32 %0, a;}" : "+r"(y[c])), FINU)

template <typename K>
void run(const char* name, K kernel, int ops_per_iter) {
  float* out; cudaMalloc(&out, 4);
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const int blocks = sms * 2, threads = 1024;
  kernel<<<blocks, threads>>>(out, 1.0001f, 0.9999f);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  kernel<<<blocks, threads>>>(out, 1.0001f, 0.9999f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double warp_instr = (double)blocks * (threads / 32) * ITERS * CHAINS * ops_per_iter;
  const double cycles = ms * 1e-3 * khz * 1e3;
  printf("%-16s %8.3f ms  %.3f warp-instr/cycle/SMSP (nominal clock %d MHz)\n", name, ms, warp_instr / cycles / sms / 4, khz / 1000);
  cudaFree(out);
}

int main() {
  run("FADD", k_fadd, 1); run("FMUL", k_fmul, 1); run("FFMA", k_ffma, 1); run("FMNMX", k_fmnmx, 1); run("FMNMX3", k_fmnmx3, 1);
  run("LOP3", k_lop3, 1); run("IADD", k_iadd, 1); run("IMAD", k_imad, 1); run("FSETP+SEL", k_selp, 2);
  run("FSETP+@FADD", k_fsetp, 2); run("2FSETP+@FADD", k_fsetp2, 3); run("ISETP+@IADD", k_isetp, 2); run("ISETP+SEL", k_sel, 2); run("PRMT", k_prmt, 1); run("SHF", k_shf, 1);
  run("FSETP+@FFMAmov", k_fsel_fma, 2);
  run("FADD2", k_fadd2, 1); run("FMUL2", k_fmul2, 1); run("FFMA2", k_ffma2, 1);
  run("FADD+FMNMX", k_fadd_fmnmx, 2); run("FMUL+FMNMX", k_fmul_fmnmx, 2); run("FADD+FMUL", k_fadd_fmul, 2); run("FFMA+FMNMX", k_ffma_fmnmx, 2);
  run("FADD2+FMNMX", k_fadd2_fmnmx, 2); run("LOP3+FMNMX", k_lop3_fmnmx, 2); run("IMAD+FMNMX", k_imad_fmnmx, 2);
  return 0;
}
