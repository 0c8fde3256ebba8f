// L2 bandwidth of the B200 for the access shapes of the traversal kernel: an L2-resident array (64 MB, the node
// array of the 1M-triangle scene) read (a) as a coalesced stream of 32-byte loads, (b) as 64-byte records at
// random indices, one record per lane per step (what an incoherent warp of rays does).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o l2bw l2bw.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void ld256(const float4* p, float4& a, float4& b) {
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}
__global__ void stream_kernel(const float4* buf, size_t n32, int reps, float* out) {  // n32 = number of 32-byte units
  float acc = 0.f;
  for (int r = 0; r < reps; r++)
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n32; i += (size_t)gridDim.x * blockDim.x) {
      float4 a, b; ld256(buf + 2 * i, a, b); acc += a.x + b.w;
    }
  if (acc == 1.2345f) out[0] = acc;
}
__global__ void gather_kernel(const float4* buf, unsigned nrec, int steps, float* out) {  // 64-byte records, random walk
  unsigned x = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  float acc = 0.f;
  for (int s = 0; s < steps; s++) {
    x = x * 1664525u + 1013904223u;
    const unsigned rec = (unsigned)(((unsigned long long)x * nrec) >> 32);
    float4 a, b, c, d; ld256(buf + 4 * (size_t)rec, a, b); ld256(buf + 4 * (size_t)rec + 2, c, d);
    acc += a.x + b.w + c.y + d.z;
  }
  if (acc == 1.2345f) out[0] = acc;
}
// The same 64-byte records fetched by LANE PAIRS: in one instruction lane 2k reads the first 32 bytes and lane 2k+1
// the second 32 bytes of ONE record (16 distinct 128-byte lines per warp instruction instead of 32), first for the
// even lane's record, then for the odd lane's.  L1TEX charges one wavefront per distinct line per instruction
// (B300_MICROARCH.md, "L1tex wavefront queue"), so the 32 records of a warp cost 32 wavefronts instead of 64.
// XCHG = 0: each lane keeps the halves it loaded (raw fetch rate).  XCHG = 1: eight shuffles hand every lane its own
// complete record (what a drop-in replacement of the per-lane fetch costs).
template <int XCHG>
__global__ void gather_pair_kernel(const float4* buf, unsigned nrec, int steps, float* out) {
  unsigned x = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  const unsigned odd = threadIdx.x & 1u;
  float acc = 0.f;
  for (int s = 0; s < steps; s++) {
    x = x * 1664525u + 1013904223u;
    const unsigned rec = (unsigned)(((unsigned long long)x * nrec) >> 32);
    const unsigned other = __shfl_xor_sync(0xffffffffu, rec, 1);
    const unsigned rec_e = odd ? other : rec, rec_o = odd ? rec : other;
    float4 a, b, c, d;
    ld256(buf + 4 * (size_t)rec_e + 2 * odd, a, b);   // my half of the even lane's record
    ld256(buf + 4 * (size_t)rec_o + 2 * odd, c, d);   // my half of the odd lane's record
    if (XCHG) {  // even lane: keeps (a,b), needs the odd lane's (a,b); odd lane: keeps (c,d), needs the even lane's (c,d)
      float4 s0 = odd ? a : c, s1 = odd ? b : d;
      s0.x = __shfl_xor_sync(0xffffffffu, s0.x, 1); s0.y = __shfl_xor_sync(0xffffffffu, s0.y, 1);
      s0.z = __shfl_xor_sync(0xffffffffu, s0.z, 1); s0.w = __shfl_xor_sync(0xffffffffu, s0.w, 1);
      s1.x = __shfl_xor_sync(0xffffffffu, s1.x, 1); s1.y = __shfl_xor_sync(0xffffffffu, s1.y, 1);
      s1.z = __shfl_xor_sync(0xffffffffu, s1.z, 1); s1.w = __shfl_xor_sync(0xffffffffu, s1.w, 1);
      if (odd) { a = s0; b = s1; } else { c = s0; d = s1; }
    }
    acc += a.x + b.w + c.y + d.z;
  }
  if (acc == 1.2345f) out[0] = acc;
}
int main() {
  const size_t bytes = 64ull << 20;
  float4* buf; float* out; cudaMalloc(&buf, bytes); cudaMalloc(&out, 4); cudaMemset(buf, 0, bytes);
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1); float ms;
  const int reps = 40;
  stream_kernel<<<sms * 8, 256>>>(buf, bytes / 32, 2, out);
  cudaEventRecord(e0); stream_kernel<<<sms * 8, 256>>>(buf, bytes / 32, reps, out); cudaEventRecord(e1); cudaEventSynchronize(e1);
  cudaEventElapsedTime(&ms, e0, e1);
  printf("L2 stream, 64 MB resident, 32-byte loads : %.0f GB/s\n", (double)bytes * reps / (ms * 1e-3) / 1e9);
  for (int tpb = 128; tpb <= 128; tpb *= 2) {
    const int blocks = sms * 16, steps = 2000;
    gather_kernel<<<blocks, tpb>>>(buf, (unsigned)(bytes / 64), 100, out);
    cudaEventRecord(e0); gather_kernel<<<blocks, tpb>>>(buf, (unsigned)(bytes / 64), steps, out); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    printf("L2 gather, 64-byte records at random      : %.0f GB/s (%d x %d threads)\n", (double)blocks * tpb * steps * 64.0 / (ms * 1e-3) / 1e9, blocks, tpb);
  }
  for (int v = 0; v < 2; v++) {
    const int blocks = sms * 16, steps = 2000, tpb = 128;
    if (v) gather_pair_kernel<1><<<blocks, tpb>>>(buf, (unsigned)(bytes / 64), 100, out); else gather_pair_kernel<0><<<blocks, tpb>>>(buf, (unsigned)(bytes / 64), 100, out);
    cudaEventRecord(e0);
    if (v) gather_pair_kernel<1><<<blocks, tpb>>>(buf, (unsigned)(bytes / 64), steps, out); else gather_pair_kernel<0><<<blocks, tpb>>>(buf, (unsigned)(bytes / 64), steps, out);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    printf("L2 gather, 64-byte records, lane pairs%s: %.0f GB/s\n", v ? " + 8 shuffles" : " (no exchange)", (double)blocks * tpb * steps * 64.0 / (ms * 1e-3) / 1e9);
  }
  // the same gather from an array that does not fit L2 (1 GB): HBM-bound
  cudaFree(buf); const size_t big = 1ull << 30; cudaMalloc(&buf, big); cudaMemset(buf, 0, big);
  { const int blocks = sms * 16, steps = 500;
    gather_kernel<<<blocks, 128>>>(buf, (unsigned)(big / 64), 50, out);
    cudaEventRecord(e0); gather_kernel<<<blocks, 128>>>(buf, (unsigned)(big / 64), steps, out); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    printf("HBM gather, 64-byte records at random (1 GB): %.0f GB/s\n", (double)blocks * 128 * steps * 64.0 / (ms * 1e-3) / 1e9); }
  return 0;
}
