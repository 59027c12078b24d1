// Linear-layer GEMM for sm_100a:  y[M,N] = epilogue(x[M,K] . w[N,K]^T), bf16 (kind::f16) or fp32
// storage with TF32 math (kind::tf32), fp32 accumulation in TMEM.
//
// One kernel, two orientations:
//   normal  : TMEM lanes = rows of x (tokens, tile 128),  TMEM columns = rows of w (tile BN)
//   swapped : TMEM lanes = rows of w (out features, 128), TMEM columns = rows of x (tokens, BN=pad16(M))
// The swapped orientation is the decode path (M = batch <= 256): the 128-row UMMA M dimension is
// filled by weight rows, weights stream once from HBM through a TMA ring, and split-K spreads one
// layer over all 148 SMs.  Both operands are K-major, so the same TMA box / UMMA descriptor code
// serves both roles.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread MMA
// issuer, warps 2..5 = epilogue (TMEM -> registers -> global), one TMEM lane quarter each.
#include "gemm.cuh"
#include "ptx.cuh"

#include <mutex>
#include <stdlib.h>

namespace mt {

bool g_force_simt = false;
long long g_launches = 0;
// bytes of weight tiles a decode GEMM asks L2 to prefetch before it waits for the previous kernel
// (MT_PF_MB in the environment overrides; 0 disables)
// experiment: BN=64 decode GEMMs with a 9-stage ring and one CTA per SM instead of 4 stages x 2 CTAs
static const bool g_deep = [] { const char* e = getenv("MT_DEEP"); return e && e[0] == '1'; }();
long long g_pf_bytes = [] {
  const char* e = getenv("MT_PF_MB");
  return (e ? atoll(e) : 40ll) << 20;
}();

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
// tcgen05 kernel
// ---------------------------------------------------------------------------------------------
constexpr int kTileA = 128;          // UMMA M
constexpr int kTileKBytes = 128;     // one 128B-swizzle row of K per k-block
constexpr int kGemmThreads = 192;

template <int BN, int STAGES>
struct GemmSmem {
  static constexpr int kABytes = kTileA * kTileKBytes;
  static constexpr int kBBytes = BN * kTileKBytes;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kBarOffset = STAGES * kStageBytes;
  static constexpr int kTotal = kBarOffset + (2 * STAGES + 1) * 8 + 16 + 1024 /* alignment slack */;
};

template <int BN, int STAGES, int KIND /*1 bf16, 2 tf32*/>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, Epi ep,
               int swapped, int rowsA, int rowsB, int k_blocks, int splits, uint64_t hintA, uint64_t hintB,
               int early /*0: none, 1: operand A is constant weights, 2: operand B is*/, int pf_kblocks,
               int tiled /*0: row-major operands, 1: A is tile-major, 2: B is*/) {
  using L = GemmSmem<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + L::kBarOffset);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tileA = blockIdx.x, tileB = blockIdx.y, split = blockIdx.z;
  const int kb0 = static_cast<int>((static_cast<long long>(k_blocks) * split) / splits);
  const int kb1 = static_cast<int>((static_cast<long long>(k_blocks) * (split + 1)) / splits);
  constexpr int kKElems = KIND == 1 ? 64 : 32;          // elements per 128-byte k-block
  constexpr uint32_t kTmemCols = BN < 32 ? 32 : BN;     // power of two >= 32

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int i = 0; i < STAGES; ++i) {
      ptx::mbar_init(&full[i], 1);
      ptx::mbar_init(&empty[i], 1);
    }
    ptx::mbar_init(tmem_full, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_holder, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  if (threadIdx.x == 0) pdl_trigger();   // the next kernel may take our place as CTAs retire

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      // Operand loads.  `tiled` (1: A, 2: B) marks a weight operand stored tile-major: every
      // {128 rows x 128 B} block of the matrix is one contiguous 16 KB run, addressed as rows
      // [(row_tile * k_blocks + kb) * 128, +128) of a [*, 128 B] matrix (see LinearPlan::w_tiled).
      auto load_a = [&](int st, int kb) {
        uint8_t* sa = smem + st * L::kStageBytes;
        if (tiled == 1) ptx::tma_load_2d(sa, &tmA, &full[st], 0, (tileA * k_blocks + kb) * kTileA, hintA);
        else ptx::tma_load_2d(sa, &tmA, &full[st], kb * kKElems, tileA * kTileA, hintA);
      };
      auto load_b = [&](int st, int kb) {
        uint8_t* sb = smem + st * L::kStageBytes + L::kABytes;
        if (tiled == 2) {
#pragma unroll
          for (int r = 0; r < BN / 128; ++r)
            ptx::tma_load_2d(sb + r * 128 * kTileKBytes, &tmB, &full[st], 0,
                             ((tileB * (BN / 128) + r) * k_blocks + kb) * 128, hintB);
        } else {
          ptx::tma_load_2d(sb, &tmB, &full[st], kb * kKElems, tileB * BN, hintB);
        }
      };
      // Constant weights do not depend on the previous kernel: fill the ring with their tiles (and ask
      // L2 for the following ones) before waiting for it, so HBM keeps streaming across the boundary.
      const int n = kb1 - kb0;
      const int pre = early ? (n < STAGES ? n : STAGES) : 0;
      for (int it = 0; it < pre; ++it) {
        ptx::mbar_expect_tx(&full[it], L::kStageBytes);
        if (early == 1) load_a(it, kb0 + it); else load_b(it, kb0 + it);
      }
      if (early == 1 && pf_kblocks > 0) {
        const int pf_end = pre + pf_kblocks < n ? pre + pf_kblocks : n;
        for (int it = pre; it < pf_end; ++it) {
          if (tiled == 1) ptx::tma_prefetch_2d(&tmA, 0, (tileA * k_blocks + kb0 + it) * kTileA);
          else ptx::tma_prefetch_2d(&tmA, (kb0 + it) * kKElems, tileA * kTileA);
        }
      }
      pdl_wait();
      for (int it = 0; it < pre; ++it) {
        if (early == 1) load_b(it, kb0 + it); else load_a(it, kb0 + it);
      }
      for (int kb = kb0 + pre, it = pre; kb < kb1; ++kb, ++it) {
        const int st = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        ptx::mbar_wait(&empty[st], ph ^ 1);
        ptx::mbar_expect_tx(&full[st], L::kStageBytes);
        load_a(st, kb);
        load_b(st, kb);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread) =====
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::make_idesc(KIND, kTileA, BN);
      for (int kb = kb0, it = 0; kb < kb1; ++kb, ++it) {
        const int st = it % STAGES;
        const uint32_t ph = (it / STAGES) & 1;
        ptx::mbar_wait(&full[st], ph);
        ptx::tc_fence_after();
        const uint32_t sa = ptx::smem_u32(smem + st * L::kStageBytes);
        const uint64_t adesc = ptx::make_smem_desc_sw128(sa);
        const uint64_t bdesc = ptx::make_smem_desc_sw128(sa + L::kABytes);
#pragma unroll
        for (int k = 0; k < 4; ++k) {   // 4 x 32-byte UMMA_K slices per 128-byte row
          if (KIND == 1) ptx::mma_bf16(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, (it | k) != 0);
          else ptx::mma_tf32(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, (it | k) != 0);
        }
        ptx::mma_commit(&empty[st]);    // frees the smem stage when these MMAs retire
      }
      ptx::mma_commit(tmem_full);       // accumulator complete
    }
  } else {
    // ===== epilogue: warps 2..5, TMEM lane quarter = warp % 4 =====
    const int q = warp & 3;
    const int a_row = tileA * kTileA + q * 32 + lane;
    ptx::mbar_wait(tmem_full, 0);
    ptx::tc_fence_after();
    pdl_wait();                          // residual reads / output writes are ordered after the previous kernel
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    constexpr int CH = BN < 32 ? 16 : 32;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += CH) {
      uint32_t r[CH];
      if constexpr (CH == 32) ptx::tmem_ld32(taddr + c0, r);
      else ptx::tmem_ld16(taddr + c0, r);
      ptx::tmem_ld_wait();
      const int b_row0 = tileB * BN + c0;
      if (swapped) {
        // lane <-> out feature n (contiguous in memory): each j is a coalesced warp store
        const int n = a_row;
        if (n < rowsA) {
#pragma unroll
          for (int j = 0; j < CH; ++j) {
            const int m = b_row0 + j;
            if (m < rowsB) {
              const float v = __uint_as_float(r[j]);
              if (ep.partial) ep.partial[(static_cast<long long>(split) * ep.M + m) * ep.N + n] = v;
              else epi_store(ep, m, n, epi_apply(ep, m, n, v));
            }
          }
        }
      } else {
        const int m = a_row;
        if (m < rowsA) {
          if (ep.partial) {
            float* dst = ep.partial + (static_cast<long long>(split) * ep.M + m) * ep.N + b_row0;
            if (b_row0 + CH <= rowsB && (ep.N & 3) == 0) {
#pragma unroll
              for (int j = 0; j < CH; j += 4)
                *reinterpret_cast<float4*>(dst + j) = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                                                                  __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
            } else {
              for (int j = 0; j < CH; ++j)
                if (b_row0 + j < rowsB) dst[j] = __uint_as_float(r[j]);
            }
          } else {
            float v[CH];
#pragma unroll
            for (int j = 0; j < CH; ++j)
              v[j] = (b_row0 + j < rowsB) ? epi_apply(ep, m, b_row0 + j, __uint_as_float(r[j])) : 0.f;
            const bool full_chunk = b_row0 + CH <= rowsB;
            if (ep.out_f32 == 1) {
              float* dst = reinterpret_cast<float*>(ep.out) + static_cast<long long>(m) * ep.ldo + b_row0;
              if (full_chunk && (ep.ldo & 3) == 0) {
#pragma unroll
                for (int j = 0; j < CH; j += 4)
                  *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
              } else {
                for (int j = 0; j < CH; ++j)
                  if (b_row0 + j < rowsB) dst[j] = v[j];
              }
            } else {
              bf16* dst = reinterpret_cast<bf16*>(ep.out) + static_cast<long long>(m) * ep.ldo + b_row0;
              if (full_chunk && (ep.ldo & 7) == 0) {
#pragma unroll
                for (int j = 0; j < CH; j += 8) {
                  uint4 pk;
                  pk.x = pack_bf16(v[j], v[j + 1]);
                  pk.y = pack_bf16(v[j + 2], v[j + 3]);
                  pk.z = pack_bf16(v[j + 4], v[j + 5]);
                  pk.w = pack_bf16(v[j + 6], v[j + 7]);
                  *reinterpret_cast<uint4*>(dst + j) = pk;
                }
              } else {
                for (int j = 0; j < CH; ++j)
                  if (b_row0 + j < rowsB) dst[j] = f2bf(v[j]);
              }
            }
          }
        }
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, kTmemCols);
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
template <int BN, int STAGES, int KIND>
static int launch_tc(const CUtensorMap& tmA, const CUtensorMap& tmB, const Epi& ep, int swapped, int rowsA, int rowsB,
                     int k_blocks, int splits, uint64_t hintA, uint64_t hintB, int early, long long pf_bytes,
                     int tiled, cudaStream_t s) {
  using L = GemmSmem<BN, STAGES>;
  auto kern = gemm_tc_kernel<BN, STAGES, KIND>;
  static bool attr_set = false;
  if (!attr_set) {
    MT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
    attr_set = true;
  }
  dim3 grid(cdiv(rowsA, kTileA), cdiv(rowsB, BN), splits);
  // L2 prefetch depth per CTA: spread pf_bytes of weight tiles over the CTAs of the first wave
  int pf_kblocks = 0;
  if (early) {
    const long long ctas = static_cast<long long>(grid.x) * grid.y * grid.z;
    const long long resident = static_cast<long long>(kNumSMs) * (L::kTotal > 113 * 1024 ? 1 : 2);
    const long long tile_bytes = early == 1 ? L::kABytes : L::kBBytes;
    pf_kblocks = static_cast<int>(pf_bytes / ((ctas < resident ? ctas : resident) * tile_bytes));
  }
  MT_CUDA(launch(kern, grid, dim3(kGemmThreads), L::kTotal, s, tmA, tmB, ep, swapped, rowsA, rowsB, k_blocks, splits,
                 hintA, hintB, early, pf_kblocks, tiled));
  count_launch();
  return MT_OK;
}

template <int KIND>
static int dispatch_tc(int BN, const CUtensorMap& tmA, const CUtensorMap& tmB, const Epi& ep, int swapped, int rowsA,
                       int rowsB, int k_blocks, int splits, uint64_t hA, uint64_t hB, int early, long long pf,
                       int tiled, cudaStream_t s) {
  switch (BN) {
    case 16: return launch_tc<16, 4, KIND>(tmA, tmB, ep, swapped, rowsA, rowsB, k_blocks, splits, hA, hB, early, pf, tiled, s);
    case 32: return launch_tc<32, 4, KIND>(tmA, tmB, ep, swapped, rowsA, rowsB, k_blocks, splits, hA, hB, early, pf, tiled, s);
    case 64:
      if (g_deep) return launch_tc<64, 9, KIND>(tmA, tmB, ep, swapped, rowsA, rowsB, k_blocks, splits, hA, hB, early, pf, tiled, s);
      return launch_tc<64, 4, KIND>(tmA, tmB, ep, swapped, rowsA, rowsB, k_blocks, splits, hA, hB, early, pf, tiled, s);
    case 128: return launch_tc<128, 6, KIND>(tmA, tmB, ep, swapped, rowsA, rowsB, k_blocks, splits, hA, hB, early, pf, tiled, s);
    case 256: return launch_tc<256, 4, KIND>(tmA, tmB, ep, swapped, rowsA, rowsB, k_blocks, splits, hA, hB, early, pf, tiled, s);
  }
  return fail(MT_ERR_ARG, "unsupported BN %d", BN);
}

static int pick_bn_swapped(int M) {
  if (M <= 16) return 16;
  if (M <= 32) return 32;
  if (M <= 64) return 64;
  if (M <= 128) return 128;
  return 256;
}

int linear(const LinearPlan& p, cudaStream_t s) {
  MT_CHECK(p.M > 0 && p.N > 0 && p.K > 0, "linear: empty problem M=%d N=%d K=%d", p.M, p.N, p.K);
  MT_CHECK(p.x && p.w, "linear: null operand");
  Epi ep{};
  ep.out = p.out; ep.ldo = p.ldo; ep.out_f32 = out_kind(p.dtype_out);
  ep.bias = p.bias; ep.act = p.act; ep.residual = p.residual; ep.ldr = p.ldr; ep.scale = p.scale;
  ep.round_bf16 = p.round_bf16; ep.M = p.M; ep.N = p.N; ep.partial = nullptr;
  ep.gu_counters = nullptr; ep.gu_act = nullptr;

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

  const int esz = dtype_bytes(p.dtype_in);
  const int k_elems = 128 / esz;
  const int k_blocks = cdiv(p.K, k_elems);
  const bool swapped = p.M <= 256;
  static const bool old_big = [] { const char* e = getenv("MT_OLD_BIG"); return e && e[0] == '1'; }();
  if (!swapped && !old_big) return linear_big(p, ep, s);
  static const bool old_dec = [] { const char* e = getenv("MT_OLD_DEC"); return e && e[0] == '1'; }();
  if (swapped && !old_dec) return linear_dec(p, ep, s);
  MT_CHECK(p.dtype_in != MT_DTYPE_F16, "linear: the legacy kernel (MT_OLD_BIG / MT_OLD_DEC) has no fp16 flavour");
  int BN, rowsA, rowsB;
  CUtensorMap tmA, tmB;
  if (p.w_tiled) MT_CHECK(p.K % k_elems == 0, "linear: tile-major weights need K %% %d == 0 (K=%d)", k_elems, p.K);
  if (swapped) {
    BN = pick_bn_swapped(p.M);
    rowsA = p.N; rowsB = p.M;
    if (p.w_tmap) tmA = *p.w_tmap;
    else if (p.w_tiled) MT_TRY(make_tmap_tiled(&tmA, p.w, p.dtype_in, p.N, p.K));
    else MT_TRY(make_tmap_2d(&tmA, p.w, p.dtype_in, p.N, p.K, p.ldw, kTileA));
    MT_TRY(make_tmap_2d(&tmB, p.x, p.dtype_in, p.M, p.K, p.ldx, BN));
  } else {
    BN = 128;
    rowsA = p.M; rowsB = p.N;
    MT_TRY(make_tmap_2d(&tmA, p.x, p.dtype_in, p.M, p.K, p.ldx, kTileA));
    if (p.w_tmap) tmB = *p.w_tmap;
    else if (p.w_tiled) MT_TRY(make_tmap_tiled(&tmB, p.w, p.dtype_in, p.N, p.K));
    else MT_TRY(make_tmap_2d(&tmB, p.w, p.dtype_in, p.N, p.K, p.ldw, BN));
  }
  const int tiled = p.w_tiled ? (swapped ? 1 : 2) : 0;
  // split-K: fill ~2 CTAs per SM when the tile grid alone cannot (decode-sized problems)
  const int tiles = cdiv(rowsA, kTileA) * cdiv(rowsB, BN);
  int splits = 1;
  const long long per_split = static_cast<long long>(p.M) * p.N * 4;
  if (p.workspace && tiles < 4 * kNumSMs) {
    // pick the split count whose CTA total fills whole waves of 2 CTAs/SM best; ties -> fewer splits
    const int slots = (g_deep && BN == 64 ? 1 : 2) * kNumSMs;
    double best_eff = 0.0;
    for (int sp = 1; sp <= 16; ++sp) {
      if (sp > 1 && k_blocks / sp < 4) break;                       // keep >= 4 k-blocks per CTA
      if (per_split * sp > p.workspace_bytes) break;
      const int ctas = tiles * sp;
      const int waves = (ctas + slots - 1) / slots;
      const double eff = static_cast<double>(ctas) / (static_cast<double>(waves) * slots);
      if (eff > best_eff + 0.02) { best_eff = eff; splits = sp; }
    }
  }
  if (p.out_splits) {
    MT_CHECK(p.workspace && per_split <= p.workspace_bytes, "linear: workspace too small for partial output");
    *p.out_splits = splits;
    ep.partial = p.workspace;
  } else if (splits > 1) {
    ep.partial = p.workspace;
  }
  // weights are streamed once in decode (evict-first); activations are re-read by every tile (evict-last)
  const uint64_t hint_w = swapped ? ptx::kEvictFirst : ptx::kEvictNormal;
  const uint64_t hint_x = ptx::kEvictLast;
  const uint64_t hA = swapped ? hint_w : hint_x, hB = swapped ? hint_x : hint_w;
  // constant weights: stream the first tiles before the previous kernel has finished (PDL).  The L2
  // prefetch only pays in the decode orientation, where the kernel is HBM-bound.
  const int early = p.w_static ? (swapped ? 1 : 2) : 0;
  const long long pf = swapped ? g_pf_bytes : 0;
  if (p.dtype_in == MT_DTYPE_BF16)
    MT_TRY(dispatch_tc<1>(BN, tmA, tmB, ep, swapped, rowsA, rowsB, k_blocks, splits, hA, hB, early, pf, tiled, s));
  else
    MT_TRY(dispatch_tc<2>(BN, tmA, tmB, ep, swapped, rowsA, rowsB, k_blocks, splits, hA, hB, early, pf, tiled, s));
  if (ep.partial && !p.out_splits) MT_TRY(splitk_reduce(ep, splits, s));
  return MT_OK;
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
