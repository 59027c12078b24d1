// Persistent tcgen05/TMA GEMM for token-major problems (M > 256): LM prefill and every codec Linear.
//   y[M,N] = epilogue(x[M,K] . w[N,K]^T),  bf16 (kind::f16) or fp32 storage with TF32 math (kind::tf32).
//
// One CTA per SM walks output tiles of 128 (tokens) x 256 (features) in an M-fastest order, so the
// CTAs running at the same time share a few weight-tile columns in L2.  Per CTA:
//   warp 0        TMA producer: 4-stage ring of {A 128x128B, B 256x128B} k-blocks (48 KB / stage)
//   warp 1        TMEM allocator + single-thread tcgen05.mma issuer (UMMA 128x256, 4 K-slices / k-block)
//   warps 2..9    epilogue: TMEM -> registers -> bias/GELU/LayerScale/residual -> global
// The accumulator is double-buffered in TMEM (2 x 256 columns = all 512), so the epilogue of tile i
// overlaps the main loop of tile i+1 -- the codec's short-K GEMMs (K = 768/1280) spend as long in the
// epilogue as in the MMAs.  Warp w may only touch TMEM lanes 32*(w%4)..+31, so the 8 epilogue warps
// split a tile as (lane quarter = w%4) x (column half = (w-2)/4).
#include "gemm.cuh"
#include "gemm_epi.cuh"
#include "ptx.cuh"

namespace mt {

namespace {

constexpr int kBM = 128, kBN = 256, kKBytes = 128, kStages = 4;
constexpr int kEpiWarps = 8;                  // warps that own a transposition tile (the EW = 8 form)
constexpr int kThreads = 64 + 32 * kEpiWarps;
constexpr int kABytes = kBM * kKBytes, kBBytes = kBN * kKBytes, kStageBytes = kABytes + kBBytes;
constexpr int kBarOffset = kStages * kStageBytes;
constexpr int kEpiOffset = kBarOffset + (2 * kStages + 4) * 8 + 32;     // 8 warps x 4 KB: epilogue transposition tiles (16-byte aligned)
constexpr int kSmemTotal = kEpiOffset + kEpiWarps * kEpiTileFloats * 4 + 1024;
static_assert(kSmemTotal <= 232448, "shared memory budget");

// EW = epilogue warps: 8 (lane quarter x column half, warp-transposed stores) or 16 (lane quarter x column quarter, per-lane
// stores): at the codec's short K the tile time is the epilogue's per-warp latency, and an ALU-heavy epilogue (GELU) halves it
// with twice the warps.
template <int KIND /*1 bf16, 2 tf32, 3 fp16*/, int EW = 8>
__global__ void __launch_bounds__(64 + 32 * EW, 1)
gemm_big_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, Epi ep, int M, int N,
                int k_blocks, int m_tiles, int n_tiles, int w_early, int w_tiled) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kBarOffset);
  uint64_t* empty = full + kStages;
  uint64_t* tmem_full = empty + kStages;     // [2]
  uint64_t* tmem_empty = tmem_full + 2;      // [2]
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  constexpr int kKElems = KIND == 2 ? 32 : 64;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // Tile walk of this CTA (iteration i = 0, 1, ...).  While whole bands of gridDim.x token tiles remain, a CTA keeps ITS
  // token tile and walks the feature tiles: the 128 x K activation tile is re-read from L2 instead of HBM (the codec's
  // M = 64,000-token GEMMs streamed the activations n_tiles times: 0.99 GB of DRAM reads for 98 MB of input), and the
  // CTAs running together still share one weight tile.  The ragged last band (and every problem with fewer token tiles
  // than CTAs, e.g. prefill) is walked in the flat token-fastest order.
  const int full_bands = m_tiles / static_cast<int>(gridDim.x);
  const int rem_m = m_tiles - full_bands * static_cast<int>(gridDim.x);
  auto tile_of = [&](int i, int& mt, int& nt) -> bool {
    const int banded = full_bands * n_tiles;
    if (i < banded) { mt = (i / n_tiles) * gridDim.x + blockIdx.x; nt = i % n_tiles; return true; }
    const int t = blockIdx.x + (i - banded) * gridDim.x;
    if (t >= rem_m * n_tiles) return false;
    mt = full_bands * gridDim.x + t % rem_m; nt = t / rem_m;
    return true;
  };

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int i = 0; i < kStages; ++i) {
      ptx::mbar_init(&full[i], 1);
      ptx::mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&tmem_full[i], 1);
      ptx::mbar_init(&tmem_empty[i], EW);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_holder, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  if (threadIdx.x == 0) pdl_trigger();

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      auto load_a = [&](int st, int mt, int kb) {
        ptx::tma_load_2d(smem + st * kStageBytes, &tmA, &full[st], kb * kKElems, mt * kBM, ptx::kEvictNormal);
      };
      auto load_b = [&](int st, int nt, int kb) {
        uint8_t* sb = smem + st * kStageBytes + kABytes;
        if (w_tiled) {
#pragma unroll
          for (int r = 0; r < kBN / 128; ++r)
            ptx::tma_load_2d(sb + r * 128 * kKBytes, &tmB, &full[st], 0, ((nt * (kBN / 128) + r) * k_blocks + kb) * 128,
                             ptx::kEvictLast);
        } else {
          ptx::tma_load_2d(sb, &tmB, &full[st], kb * kKElems, nt * kBN, ptx::kEvictLast);
        }
      };
      int it = 0;
      int pre = 0;
      int mt = 0, nt = 0;
      if (w_early && tile_of(0, mt, nt)) {
        // constant weights: first ring fill before the previous kernel is known to have finished
        const int nt0 = nt;
        pre = k_blocks < kStages ? k_blocks : kStages;
        for (int i = 0; i < pre; ++i) {
          ptx::mbar_expect_tx(&full[i], kStageBytes);
          load_b(i, nt0, i);
        }
      }
      pdl_wait();
      for (int i = 0; tile_of(i, mt, nt); ++i) {
        for (int kb = 0; kb < k_blocks; ++kb, ++it) {
          const int st = it % kStages;
          if (it < pre) { load_a(st, mt, kb); continue; }
          ptx::mbar_wait(&empty[st], ((it / kStages) & 1) ^ 1);
          ptx::mbar_expect_tx(&full[st], kStageBytes);
          load_a(st, mt, kb);
          load_b(st, nt, kb);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::make_idesc(KIND == 3 ? 0 : KIND, kBM, kBN);   // a/b format: 0 F16, 1 BF16, 2 TF32
      int it = 0, mt = 0, nt = 0;
      for (int i = 0; tile_of(i, mt, nt); ++i) {
        const int buf = i & 1;
        ptx::mbar_wait(&tmem_empty[buf], ((i >> 1) & 1) ^ 1);     // epilogue drained this accumulator
        ptx::tc_fence_after();
        const uint32_t tacc = tmem_base + buf * kBN;
        for (int kb = 0; kb < k_blocks; ++kb, ++it) {
          const int st = it % kStages;
          ptx::mbar_wait(&full[st], (it / kStages) & 1);
          ptx::tc_fence_after();
          const uint32_t sa = ptx::smem_u32(smem + st * kStageBytes);
          const uint64_t adesc = ptx::make_smem_desc_sw128(sa);
          const uint64_t bdesc = ptx::make_smem_desc_sw128(sa + kABytes);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (KIND != 2) ptx::mma_bf16(tacc, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);   // kind::f16
            else ptx::mma_tf32(tacc, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
          }
          ptx::mma_commit(&empty[st]);
        }
        ptx::mma_commit(&tmem_full[buf]);
      }
    }
  } else {
    // ===== epilogue =====
    const int q = warp & 3;                    // TMEM lane quarter this warp may access
    constexpr int kCols = kBN / (EW / 4);      // columns per warp: a half (EW = 8) or a quarter (EW = 16) of the tile
    const int half = (warp - 2) >> 2;          // which of them
    float* tile = reinterpret_cast<float*>(smem + kEpiOffset) + ((warp - 2) % kEpiWarps) * kEpiTileFloats;
    const bool coalesced = EW == 8 && epi_coalesced_ok(ep);
    pdl_wait();                                // residual reads / output writes follow the previous kernel
    int mt = 0, nt = 0;
    for (int i = 0; tile_of(i, mt, nt); ++i) {
      const int buf = i & 1;
      const int m0 = mt * kBM + q * 32;
      const int n_base = nt * kBN + half * kCols;
      EpiRes res;
      if (coalesced && n_base + 32 <= N && m0 < M) epi_res_load(ep, m0, n_base, M, lane, res);   // lands behind the wait
      ptx::mbar_wait(&tmem_full[buf], (i >> 1) & 1);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * kBN + half * kCols;
      epi_tile_warp<kCols>(ep, coalesced, m0, n_base, M, N, taddr, tile, lane, res);
      // this warp is done reading the accumulator: hand it back to the MMA warp
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tmem_empty[buf]);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, 512);
}

}  // namespace

int linear_big(const LinearPlan& p, const Epi& ep, cudaStream_t s) {
  const int esz = dtype_bytes(p.dtype_in);
  const int k_elems = 128 / esz;
  const int k_blocks = cdiv(p.K, k_elems);
  CUtensorMap tmA, tmB;
  MT_TRY(make_tmap_2d(&tmA, p.x, p.dtype_in, p.M, p.K, p.ldx, kBM));
  if (p.w_tiled) {
    if (p.w_tmap) tmB = *p.w_tmap;
    else MT_TRY(make_tmap_tiled(&tmB, p.w, p.dtype_in, p.N, p.K));
  } else {
    // the cached map (if any) has a 128-row box; this kernel wants 256 rows per load
    MT_TRY(make_tmap_2d(&tmB, p.w, p.dtype_in, p.N, p.K, p.ldw, kBN));
  }
  const int m_tiles = cdiv(p.M, kBM), n_tiles = cdiv(p.N, kBN);
  const int total = m_tiles * n_tiles;
  const int grid = total < kNumSMs ? total : kNumSMs;
  const int kind = mma_kind(p.dtype_in);
  const int w_early = p.w_static ? 1 : 0, w_tiled = p.w_tiled ? 1 : 0;
  // 16 epilogue warps for the GELU epilogue (fp16 operands: the codec's linear1); mt_gemm_debug_set_pair(+32) keeps 8
  const bool ew16 = kind == 3 && ep.act == 1 && !(g_gemm_pair_mode & 32);
  static bool attr_set[4] = {false, false, false, false};
  auto go = [&](auto kern, int threads) -> int {
    bool& done = attr_set[ew16 ? 0 : kind];
    if (!done) {
      MT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
      done = true;
    }
    MT_CUDA(launch(kern, dim3(grid), dim3(threads), kSmemTotal, s, tmA, tmB, ep, p.M, p.N, k_blocks, m_tiles, n_tiles, w_early,
                   w_tiled));
    return MT_OK;
  };
  if (ew16) MT_TRY(go(gemm_big_kernel<3, 16>, 64 + 32 * 16));
  else if (kind == 1) MT_TRY(go(gemm_big_kernel<1, 8>, kThreads));
  else if (kind == 2) MT_TRY(go(gemm_big_kernel<2, 8>, kThreads));
  else MT_TRY(go(gemm_big_kernel<3, 8>, kThreads));
  count_launch();
  return MT_OK;
}

}  // namespace mt
