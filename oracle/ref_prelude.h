/* oracle/ref_prelude.h — force-included (-include) when the reference's device sources are compiled
 * for the HOST (oracle/_ref/libref_host.so).  TEST INFRASTRUCTURE ONLY.
 * Makes every __device__ function of src/{bvh_traversal,struct,helper,draw}.cu an ordinary host
 * symbol and lets cuRAND's device API take its host branches (curand_kernel.h precalc_*_host). */
#include <cuda_runtime.h>
#define QUALIFIERS static __forceinline__ __host__ __device__
#include <curand_kernel.h>
#undef __device__
#define __device__ __host__ __location__(device)
