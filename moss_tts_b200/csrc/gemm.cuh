// Internal interface of the Linear-layer GEMMs (gemm_tc.cu).
#pragma once
#include <cuda_fp16.h>
#include <cuda.h>
#include "common.cuh"

namespace mt {

// A tensor map for a row-major [rows, K] matrix with a {128 B x box_rows} box and 128B swizzle.
int make_tmap_2d(CUtensorMap* out, const void* base, int dtype, long long rows, long long K, long long ld_elems,
                 int box_rows);

// Same for a tile-major weight matrix (see LinearPlan::w_tiled); box = one 16 KB tile.
int make_tmap_tiled(CUtensorMap* out, const void* base, int dtype, long long N, long long K);

struct LinearPlan {
  // y[M,N] = x[M,K] . w[N,K]^T
  int dtype_in = MT_DTYPE_BF16, dtype_out = MT_DTYPE_BF16;
  int M = 0, N = 0, K = 0;
  const void* x = nullptr; long long ldx = 0;
  const void* w = nullptr; long long ldw = 0;
  void* out = nullptr; long long ldo = 0;
  const float* bias = nullptr;
  int act = 0;
  const void* residual = nullptr; long long ldr = 0;
  const float* scale = nullptr;
  int round_bf16 = 0;
  float* workspace = nullptr; long long workspace_bytes = 0;
  // If non-null the kernel writes fp32 split-K partials [splits][M][N] to `workspace` and leaves the
  // reduction + epilogue to the consumer kernel; *out_splits receives the split count (>=1).
  int* out_splits = nullptr;
  int impl = MT_GEMM_AUTO;
  const CUtensorMap* w_tmap = nullptr;   // optional cached weight map (box rows 128)
  // w is constant while the stream's work is in flight: weight tiles may be fetched (smem + L2
  // prefetch) before the previous kernel finished (PDL).  Never set it for a w produced on-stream.
  bool w_static = false;
  // w is stored tile-major: [ceil(N/128)][K/k_elems][128][k_elems] (k_elems = 64 bf16 / 32 fp32), so
  // each TMA box is 16 KB of consecutive HBM addresses.  ldw is ignored; w_tmap must match.
  bool w_tiled = false;
  // Token-major kernels, 16-bit output, no activation: rotary embedding applied to output columns [0, rope_cols) in the
  // epilogue (interleaved pairs inside 64-wide heads; rope = [rope_T][32] (cos, sin) floats, row = token index % rope_T) --
  // the codec's in_proj hands the attention kernel rotated q and k.
  const float* rope = nullptr; int rope_T = 0, rope_cols = 0;
  // Decode orientation only: when *tile_filter_flag == 0 (device int) only output features
  // [tile_lo, tile_hi] (rounded out to 128-row weight tiles) are computed; the rest of `out` is left
  // untouched.  nullptr = always compute everything.
  const int* tile_filter_flag = nullptr;
  int tile_lo = 0, tile_hi = 0;
};

// ---------------------------------------------------------------------------------------------
// epilogue description shared by the tcgen05 kernel, the SIMT kernel and the split-K reducer
// ---------------------------------------------------------------------------------------------
struct Epi {
  void* out; long long ldo; int out_f32;   // element type of out / residual: 1 fp32, 0 bf16, 2 fp16
  const float* bias; int act;
  const void* residual; long long ldr;
  const float* scale; int round_bf16;
  int M, N;                       // logical output extent
  float* partial;                 // != nullptr: write raw fp32 partials [split][M][N]
  int direct_store;               // token-major kernels: 1 = per-lane row stores (A/B of the warp-transposed epilogue)
  const float4* rope; int rope_T, rope_cols;   // != nullptr: rotate columns [0, rope_cols) (see LinearPlan::rope)
};

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752440f)); }

__device__ __forceinline__ float epi_apply(const Epi& e, int m, int n, float v) {
  if (e.bias) v += e.bias[n];
  if (e.act == 1) v = gelu_erf(v);
  if (e.round_bf16) v = rbf(v);
  if (e.residual) {
    float r = e.out_f32 == 1 ? reinterpret_cast<const float*>(e.residual)[m * e.ldr + n]
              : e.out_f32 == 2 ? __half2float(reinterpret_cast<const __half*>(e.residual)[m * e.ldr + n])
                               : bf2f(reinterpret_cast<const bf16*>(e.residual)[m * e.ldr + n]);
    v = r + (e.scale ? e.scale[n] * v : v);
  } else if (e.scale) {
    v = e.scale[n] * v;
  }
  return v;
}
__device__ __forceinline__ void epi_store(const Epi& e, long long m, long long n, float v) {
  if (e.out_f32 == 1) reinterpret_cast<float*>(e.out)[m * e.ldo + n] = v;
  else if (e.out_f32 == 2) reinterpret_cast<__half*>(e.out)[m * e.ldo + n] = __float2half_rn(v);
  else reinterpret_cast<bf16*>(e.out)[m * e.ldo + n] = f2bf(v);
}

static inline int dtype_bytes(int dtype) { return dtype == MT_DTYPE_F32 ? 4 : 2; }
static inline int out_kind(int dtype) { return dtype == MT_DTYPE_F32 ? 1 : (dtype == MT_DTYPE_F16 ? 2 : 0); }
// tcgen05 kernel flavour of an input dtype: 1 bf16 (kind::f16), 2 tf32 (kind::tf32, fp32 storage), 3 fp16 (kind::f16)
static inline int mma_kind(int dtype) { return dtype == MT_DTYPE_BF16 ? 1 : (dtype == MT_DTYPE_F32 ? 2 : 3); }

// Persistent 128x256-tile kernel for token-major problems (M > 256): prefill and the codec (gemm_big.cu).
int linear_big(const LinearPlan& p, const Epi& ep, cudaStream_t s);
// The same problems on CTA pairs: 256x256 tiles, tcgen05.mma.cta_group::2 (gemm_pair.cu); mt_gemm_debug_set_pair(1) keeps
// the one-SM kernel for A/B measurements.
bool linear_pair_applies(const LinearPlan& p);
int linear_pair(const LinearPlan& p, const Epi& ep, cudaStream_t s);
extern int g_gemm_pair_mode;

// Persistent decode-orientation kernel (M <= 256): one CTA per SM streams (weight tile, K split) units (gemm_dec.cu).
int linear_dec(const LinearPlan& p, Epi ep, cudaStream_t s);
int dec_grid_ctas(int M, int N, int K, long long workspace_bytes);
// Deterministic reduction of [splits][M][N] fp32 partials + epilogue (gemm_tc.cu).
int splitk_reduce(const Epi& ep, int splits, cudaStream_t s);

// Launches the GEMM.  With out_splits == nullptr the result is complete in `out` on return of the
// stream work (split-K, when chosen, is reduced by a second kernel).
int linear(const LinearPlan& p, cudaStream_t s);

extern bool g_force_simt;
extern long long g_launches;
inline void count_launch(int n = 1) { g_launches += n; }

}  // namespace mt
