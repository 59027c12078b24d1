#include <cuda_runtime.h>
#include <cuda_bf16.h>
__device__ __forceinline__ unsigned bmax2(unsigned a, unsigned b) { unsigned d; asm("max.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ unsigned bge2(unsigned a, unsigned b) { unsigned d; asm("set.ge.u32.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__global__ void k(const uint4* a, unsigned* o, int n, unsigned thr) {
  int hits = 0;
  for (int i = threadIdx.x; i < n; i += 256) {
    uint4 w = a[i];
    unsigned m = bmax2(bmax2(w.x, w.y), bmax2(w.z, w.w));
    if (bge2(m, thr)) hits++;
  }
  o[threadIdx.x] = hits;
}
