// Shared device/host helpers for libmosstts_sm100 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <utility>

#include <cuda_fp16.h>
#include "../../include/mosstts.h"

namespace mt {

// ---- error plumbing (no exceptions cross the C ABI) ------------------------------------------
void set_error(const std::string& msg);
int fail(int code, const char* fmt, ...);

#define MT_CUDA(expr)                                                                          \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess)                                                                     \
      return mt::fail(MT_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #expr,                \
                      cudaGetErrorString(_e));                                                 \
  } while (0)

#define MT_CHECK(cond, ...)                                                                    \
  do {                                                                                         \
    if (!(cond)) return mt::fail(MT_ERR_ARG, __VA_ARGS__);                                     \
  } while (0)

#define MT_TRY(expr)                                                                           \
  do {                                                                                         \
    int _s = (expr);                                                                           \
    if (_s != MT_OK) return _s;                                                                \
  } while (0)

#define MT_LAUNCH_CHECK() MT_CUDA(cudaGetLastError())

constexpr int kNumSMs = 148;  // B200

// ---- programmatic dependent launch (PDL) -----------------------------------------------------
// Every kernel of the library starts with pdl_trigger() (the next kernel on the stream may be
// scheduled once all our CTAs are resident) and calls pdl_wait() before it touches any memory an
// earlier kernel of the stream may still be writing.  Work that only reads constants (weights,
// tables) may be issued before pdl_wait(): that is how a GEMM streams the first weight tiles into
// shared memory / L2 while the previous kernel is still running.  Both instructions are no-ops for
// a kernel launched without the attribute.  MT_NO_PDL=1 in the environment disables the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();
// A/B switches between alternative (all correct) kernels are read from the environment only in a library built with
// -DMT_EXPERIMENTS (make EXPERIMENTS=1); the product build ignores the environment and has no work-skipping switch.
int exp_env(const char* name, int dflt);
// Debug timeline buffer (mt_debug_set_trace): returns the next `words` uint64 slots, or nullptr when tracing is off.
unsigned long long* trace_slot(int words);

template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(std::forward<Args>(args))...);
}

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
static inline long long cdivll(long long a, long long b) { return (a + b - 1) / b; }

// ---- bf16 helpers ----------------------------------------------------------------------------
typedef __nv_bfloat16 bf16;

__device__ __forceinline__ float bf2f(bf16 x) { return __bfloat162float(x); }
__device__ __forceinline__ bf16 f2bf(float x) { return __float2bfloat16_rn(x); }
// round a float through bf16 (emulates a PyTorch op that returns a bf16 tensor)
__device__ __forceinline__ float rbf(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

__device__ __forceinline__ float bits_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ float bits_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ uint32_t pack_f16(float lo, float hi) {
  __half2 v = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// ---- reductions ------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide sum; `scratch` must hold >= 33 floats. All threads get the result.
__device__ __forceinline__ float block_sum(float v, float* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  if (warp == 0) {
    float t = lane < nw ? scratch[lane] : 0.f;
    t = warp_sum(t);
    if (lane == 0) scratch[32] = t;
  }
  __syncthreads();
  return scratch[32];
}
__device__ __forceinline__ float block_max(float v, float* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  if (warp == 0) {
    float t = lane < nw ? scratch[lane] : -INFINITY;
    t = warp_max(t);
    if (lane == 0) scratch[32] = t;
  }
  __syncthreads();
  return scratch[32];
}

// streaming 16-byte loads that do not pollute L1
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

}  // namespace mt
