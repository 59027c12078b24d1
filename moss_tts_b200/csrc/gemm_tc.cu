// Linear-layer GEMM front end for sm_100a:  y[M,N] = epilogue(x[M,K] . w[N,K]^T).
// Holds what the two tcgen05 kernels share: TMA tensor maps, the deterministic split-K reducer, the SIMT
// cross-check kernel (tests only) and the dispatcher `linear()`:
//   M <= 256 (decode)         -> gemm_dec.cu  (TMEM lanes = 128 weight rows, columns = tokens)
//   M  > 256 (prefill, codec) -> gemm_big.cu  (persistent 128x256 output tiles)
#include "gemm.cuh"
#include "ptx.cuh"

#include <mutex>
#include <stdlib.h>

namespace mt {

bool g_force_simt = false;
long long g_launches = 0;

// ---------------------------------------------------------------------------------------------
// tensor maps
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int make_tmap_2d(CUtensorMap* out, const void* base, int dtype, long long rows, long long K, long long ld_elems,
                 int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return fail(MT_ERR_CUDA, "cuTensorMapEncodeTiled is unavailable (no CUDA driver?)");
  const int esz = dtype_bytes(dtype);
  if ((reinterpret_cast<uintptr_t>(base) & 15) || ((ld_elems * esz) & 15))
    return fail(MT_ERR_ARG, "TMA operand must be 16-byte aligned with a 16-byte-multiple row pitch (ld=%lld)", ld_elems);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(K), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld_elems) * esz};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(128 / esz), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, dtype == MT_DTYPE_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                         : (dtype == MT_DTYPE_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32), 2,
                  const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(MT_ERR_CUDA, "cuTensorMapEncodeTiled failed with %d", static_cast<int>(r));
  return MT_OK;
}

// Tile-major weights: [ceil(N/128)][K/k_elems][128][k_elems], i.e. every {128 rows x 128 B} operand
// tile is one contiguous 16 KB run in HBM (a TMA box then reads 16 KB of consecutive addresses instead
// of 128 runs of 128 B at a K*esz pitch).  Rows beyond N in the last tile are zero padding.
int make_tmap_tiled(CUtensorMap* out, const void* base, int dtype, long long N, long long K) {
  const int k_elems = 128 / dtype_bytes(dtype);
  const long long rows = ((N + 127) / 128) * (K / k_elems) * 128;
  return make_tmap_2d(out, base, dtype, rows, k_elems, k_elems, 128);
}

// ---------------------------------------------------------------------------------------------
// split-K reduction + epilogue (deterministic order: split 0,1,2,...)
// ---------------------------------------------------------------------------------------------
__global__ void splitk_reduce_kernel(Epi ep, int splits) {
  pdl_trigger();
  pdl_wait();
  const long long total = static_cast<long long>(ep.M) * ep.N;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float v = 0.f;
    for (int s = 0; s < splits; ++s) v += ep.partial[s * total + i];
    const int m = static_cast<int>(i / ep.N), n = static_cast<int>(i % ep.N);
    epi_store(ep, m, n, epi_apply(ep, m, n, v));
  }
}

// ---------------------------------------------------------------------------------------------
// SIMT cross-check kernel (fp32 FMA; tests only).  One thread per output, 16x16 tiles.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void gemm_simt_kernel(const T* __restrict__ x, long long ldx, const T* __restrict__ w, long long ldw, int K,
                                 Epi ep, int w_tiled) {
  constexpr int KE = 128 / static_cast<int>(sizeof(T));   // elements per 128-byte k-block
  __shared__ float sx[16][17], sw[16][17];
  pdl_trigger();
  pdl_wait();
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int m = blockIdx.y * 16 + ty, n = blockIdx.x * 16 + tx;
  float acc = 0.f;
  for (int k0 = 0; k0 < K; k0 += 16) {
    const int mk = blockIdx.y * 16 + ty, nk = blockIdx.x * 16 + ty;
    float a = 0.f, b = 0.f;
    if (mk < ep.M && k0 + tx < K) a = static_cast<float>(x[static_cast<long long>(mk) * ldx + k0 + tx]);
    if (nk < ep.N && k0 + tx < K) {
      const int k = k0 + tx;
      const long long off = w_tiled ? ((static_cast<long long>(nk >> 7) * (K / KE) + k / KE) * 128 + (nk & 127)) * KE + k % KE
                                    : static_cast<long long>(nk) * ldw + k;
      b = static_cast<float>(w[off]);
    }
    sx[ty][tx] = a;
    sw[ty][tx] = b;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 16; ++k) acc += sx[ty][k] * sw[tx][k];
    __syncthreads();
  }
  if (m < ep.M && n < ep.N) {
    if (ep.partial) ep.partial[static_cast<long long>(m) * ep.N + n] = acc;
    else epi_store(ep, m, n, epi_apply(ep, m, n, acc));
  }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
int linear(const LinearPlan& p, cudaStream_t s) {
  MT_CHECK(p.M > 0 && p.N > 0 && p.K > 0, "linear: empty problem M=%d N=%d K=%d", p.M, p.N, p.K);
  MT_CHECK(p.x && p.w, "linear: null operand");
  Epi ep{};
  ep.out = p.out; ep.ldo = p.ldo; ep.out_f32 = out_kind(p.dtype_out);
  ep.bias = p.bias; ep.act = p.act; ep.residual = p.residual; ep.ldr = p.ldr; ep.scale = p.scale;
  ep.round_bf16 = p.round_bf16; ep.M = p.M; ep.N = p.N; ep.partial = nullptr;
  ep.direct_store = (g_gemm_pair_mode & 2) ? 1 : 0;
  ep.rope = reinterpret_cast<const float4*>(p.rope); ep.rope_T = p.rope_T; ep.rope_cols = p.rope_cols;
  if (p.rope) {
    // the rotation lives in the warp-transposed epilogue of the token-major tcgen05 kernels only
    MT_CHECK(p.M > 256 && p.impl != MT_GEMM_SIMT && !g_force_simt && !ep.direct_store && p.act == 0 && !p.out_splits && !p.residual &&
                 out_kind(p.dtype_out) != 1 && p.N % 32 == 0 && p.ldo % 8 == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15) == 0 &&
                 p.rope_T > 0 && p.rope_cols % 64 == 0 && (reinterpret_cast<uintptr_t>(p.rope) & 15) == 0,
             "linear: a rotary epilogue needs the token-major kernel, a 16-bit 16-byte aligned output and N %% 32 == 0");
  }

  if (p.impl == MT_GEMM_SIMT || g_force_simt) {
    if (p.out_splits) { *p.out_splits = 1; ep.partial = p.workspace; }
    dim3 grid(cdiv(p.N, 16), cdiv(p.M, 16)), block(16, 16);
    if (p.dtype_in == MT_DTYPE_BF16)
      MT_CUDA(launch(gemm_simt_kernel<bf16>, grid, block, 0, s, reinterpret_cast<const bf16*>(p.x), p.ldx,
                     reinterpret_cast<const bf16*>(p.w), p.ldw, p.K, ep, static_cast<int>(p.w_tiled)));
    else if (p.dtype_in == MT_DTYPE_F16)
      MT_CUDA(launch(gemm_simt_kernel<__half>, grid, block, 0, s, reinterpret_cast<const __half*>(p.x), p.ldx,
                     reinterpret_cast<const __half*>(p.w), p.ldw, p.K, ep, static_cast<int>(p.w_tiled)));
    else
      MT_CUDA(launch(gemm_simt_kernel<float>, grid, block, 0, s, reinterpret_cast<const float*>(p.x), p.ldx,
                     reinterpret_cast<const float*>(p.w), p.ldw, p.K, ep, static_cast<int>(p.w_tiled)));
    count_launch();
    return MT_OK;
  }

  if (p.w_tiled) {
    const int k_elems = 128 / dtype_bytes(p.dtype_in);
    MT_CHECK(p.K % k_elems == 0, "linear: tile-major weights need K %% %d == 0 (K=%d)", k_elems, p.K);
  }
  if (p.M <= 256) return linear_dec(p, ep, s);
  return linear_pair_applies(p) ? linear_pair(p, ep, s) : linear_big(p, ep, s);
}

int splitk_reduce(const Epi& ep, int splits, cudaStream_t s) {
  const long long total = static_cast<long long>(ep.M) * ep.N;
  int blocks = static_cast<int>(cdivll(total, 256));
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  MT_CUDA(launch(splitk_reduce_kernel, dim3(blocks), dim3(256), 0, s, ep, splits));
  count_launch();
  return MT_OK;
}

}  // namespace mt
