// HBM read-bandwidth probe: how fast can 148 SMs stream a weight matrix with 1-D bulk async copies (the TMA engine) into a
// shared-memory ring, as a function of bytes in flight per SM, chunk size, CTAs per SM and the address pattern?
// Answers "what can the decode GEMM's weight stream reach" (gemm_dec streams 16 KB tiles through an 8-stage ring, 1 CTA/SM).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bw_probe bw_probe.cu && ./bw_probe
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__global__ void probe(const char* __restrict__ base, long long n_chunks, int chunk, int stages, int pattern, int panel) {
  extern __shared__ __align__(128) unsigned char smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
  unsigned char* ring = smem + 1024;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar + s)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const long long G = gridDim.x, c = blockIdx.x;
    // pattern 0: chunk i of CTA c = c + i*G (interleaved); pattern 1: CTA c owns panels of `panel` consecutive chunks,
    // panel j of CTA c = (c + j*G) (tile-major weight layout: a weight tile row-panel is contiguous)
    auto addr = [&](long long i) -> long long {
      if (pattern == 0) return c + i * G;
      const long long j = i / panel, r = i % panel;
      return (c + j * G) * panel + r;
    };
    long long mine = 0;
    if (pattern == 0) mine = (n_chunks - c + G - 1) / G;
    else { const long long panels = n_chunks / panel; mine = ((panels - c + G - 1) / G) * panel; }
    long long issued = 0;
    auto issue = [&](long long i) {
      const int s = static_cast<int>(i % stages);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar + s)), "r"(chunk) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                   ::"r"(smem_u32(ring + static_cast<size_t>(s) * chunk)), "l"(base + addr(i) * chunk), "r"(chunk),
                     "r"(smem_u32(bar + s)), "l"(0x12F0000000000000ull) : "memory");
    };
    for (; issued < mine && issued < stages; ++issued) issue(issued);
    for (long long i = 0; i < mine; ++i) {
      const int s = static_cast<int>(i % stages);
      const uint32_t parity = static_cast<uint32_t>((i / stages) & 1);
      uint32_t done = 0;
      do {
        asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.b32 %0, 1, 0, P1;\n\t}"
                     : "=r"(done) : "r"(smem_u32(bar + s)), "r"(parity) : "memory");
      } while (!done);
      if (issued < mine) { issue(issued); ++issued; }
    }
  }
}

int main() {
  const size_t total = 3ull << 30;                       // 3 GB of "weights": far larger than the 126 MB L2
  char* buf = nullptr;
  if (cudaMalloc(&buf, total) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMemset(buf, 1, total);
  const size_t slice = 201326592;                        // the gate|up matrix of the 8B shape
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  struct Cfg { int ctas_per_sm, stages, chunk, pattern, panel; };
  std::vector<Cfg> cfgs;
  for (int pattern : {1, 0})
    for (int chunk : {16384, 32768, 8192})
      for (int cps : {1, 2})
        for (int inflight_kb : {64, 128, 192, 216}) {
          const int per_cta = inflight_kb * 1024 / cps;
          const int stages = per_cta / chunk;
          if (stages < 2) continue;
          cfgs.push_back({cps, stages, chunk, pattern, 1048576 / chunk});
        }
  printf("pattern ctas/SM stages chunk inflight_KB/SM  GB/s\n");
  for (const Cfg& c : cfgs) {
    const int grid = 148 * c.ctas_per_sm;
    const size_t smem = 1024 + static_cast<size_t>(c.stages) * c.chunk;
    if (smem * c.ctas_per_sm > 227 * 1024) continue;
    const long long n_chunks = slice / c.chunk;
    const int reps = 14;                                 // 14 distinct slices = 2.8 GB: every launch streams from HBM
    for (int warm = 0; warm < 2; ++warm) {
      cudaEventRecord(e0);
      for (int r = 0; r < reps; ++r)
        probe<<<grid, 32, smem>>>(buf + static_cast<size_t>(r) * slice, n_chunks, c.chunk, c.stages, c.pattern, c.panel);
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
    }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (cudaGetLastError() != cudaSuccess) { printf("launch error\n"); return 1; }
    printf("%d %d %d %d %d  %.0f\n", c.pattern, c.ctas_per_sm, c.stages, c.chunk,
           static_cast<int>(static_cast<size_t>(c.stages) * c.chunk * c.ctas_per_sm / 1024), reps * (slice / 1e9) / (ms / 1e3));
  }
  // one long launch (no launch gaps): the whole 2.8 GB in one kernel
  for (int stages : {8, 12}) {
    const long long n_chunks = 14 * slice / 16384;
    cudaEventRecord(e0);
    probe<<<148, 32, 1024 + stages * 16384>>>(buf, n_chunks, 16384, stages, 1, 64);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    printf("single launch, %d stages x 16 KB: %.0f GB/s\n", stages, 14 * (slice / 1e9) / (ms / 1e3));
  }
  return 0;
}
