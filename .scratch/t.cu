#include <cuda_runtime.h>
__global__ void k(const unsigned* a, unsigned* o, int n) {
  unsigned m = 0;
  for (int i = threadIdx.x; i < n; i += 32) m = __vmaxu2(m, a[i]);
  unsigned c = __vcmpgeu2(m, a[0]);
  o[threadIdx.x] = m ^ c;
}
