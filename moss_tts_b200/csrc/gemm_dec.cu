// Persistent tcgen05/TMA GEMM for the decode orientation (M = batch <= 256 tokens):
//   y[M,N] = epilogue(x[M,K] . w[N,K]^T)   with TMEM lanes = 128 weight rows, TMEM columns = tokens.
// The kernel is HBM-bound (every weight byte is read once per step), so the design goal is a weight stream
// that never pauses:
//   * one CTA per SM owns a deep TMA ring (all of shared memory) and walks a static list of work units
//     (weight tile, K split); the producer warp runs ahead into the next unit while the MMA warp and the
//     epilogue are still busy with the current one, so there is no per-unit pipeline fill/drain;
//   * the accumulator is double-buffered in TMEM, the epilogue (fp32 split-K partials or the finished
//     output) overlaps the next unit's main loop;
//   * constant weights are requested before griddepcontrol.wait (PDL): the ring is full of the first
//     weight tiles by the time the previous kernel has produced x.
// The non-persistent one-unit-per-CTA kernel (gemm_tc.cu) gave every CTA a ~5 us latency chain
// (prologue, first TMA round trip, epilogue, exit) for as little as 112 KB of weights.
#include "gemm.cuh"
#include "ptx.cuh"

#include <stdlib.h>

namespace mt {

namespace {

constexpr int kTileA = 128, kKBytes = 128;
constexpr int kEpiWarps = 8;
constexpr int kThreads = 64 + 32 * kEpiWarps;

template <int BN>
struct DecCfg {
  static constexpr int kABytes = kTileA * kKBytes;
  static constexpr int kBBytes = BN * kKBytes;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (200 * 1024) / kStageBytes;       // 16:11  32:10  64:8  128:6  256:4
  static constexpr int kBarOffset = kStages * kStageBytes;
  static constexpr int kTotal = kBarOffset + (2 * kStages + 4) * 8 + 16 + 1024;
  static constexpr uint32_t kTmemCols = 2 * BN < 32 ? 32 : 2 * BN;   // two accumulators, power of two >= 32
};

template <int BN, int KIND /*1 bf16, 2 tf32, 3 fp16*/>
__global__ void __launch_bounds__(kThreads, 1)
gemm_dec_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, Epi ep, int rowsA,
                int rowsB, int k_blocks, int tiles, int splits, int w_early, int w_tiled,
                const int* __restrict__ full_flag, int keep_lo, int keep_hi,
                unsigned long long* __restrict__ trace) {
  using C = DecCfg<BN>;
  constexpr int S = C::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + C::kBarOffset);
  uint64_t* empty = full + S;
  uint64_t* tmem_full = empty + S;           // [2]
  uint64_t* tmem_empty = tmem_full + 2;      // [2]
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  constexpr int kKElems = KIND == 2 ? 32 : 64;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int units = tiles * splits;
  // unit u = split * tiles + tile; k-range of a split as in gemm_tc_kernel
  auto kb_begin = [&](int split) { return static_cast<int>((static_cast<long long>(k_blocks) * split) / splits); };
  // Optional row-tile filter (the text head while every utterance is inside an audio span: only the
  // gen/delay-slot logits are read by the sampler, modeling_moss_tts.py:459-460): when *full_flag == 0 only
  // the weight tiles keep_lo..keep_hi are computed.  The flag is read after pdl_wait by every role.
  auto skip_tile = [&](bool restricted, int tile) { return restricted && (tile < keep_lo || tile > keep_hi); };
  // timeline instrumentation (mt_debug_set_trace; null in every product run): 8 global-timer stamps per CTA
  auto stamp = [&](int e) {
    if (trace) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      trace[blockIdx.x * 8 + e] = t;
    }
  };
  if (threadIdx.x == 0) stamp(0);

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int i = 0; i < S; ++i) {
      ptx::mbar_init(&full[i], 1);
      ptx::mbar_init(&empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&tmem_full[i], 1);
      ptx::mbar_init(&tmem_empty[i], kEpiWarps);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_holder, C::kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  if (threadIdx.x == 0) { pdl_trigger(); stamp(1); }

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      auto load_w = [&](int st, int tile, int kb) {
        uint8_t* sa = smem + st * C::kStageBytes;
        if (w_tiled) ptx::tma_load_2d(sa, &tmA, &full[st], 0, (tile * k_blocks + kb) * kTileA, ptx::kEvictFirst);
        else ptx::tma_load_2d(sa, &tmA, &full[st], kb * kKElems, tile * kTileA, ptx::kEvictFirst);
      };
      auto load_x = [&](int st, int kb) {
        ptx::tma_load_2d(smem + st * C::kStageBytes + C::kABytes, &tmB, &full[st], kb * kKElems, 0, ptx::kEvictLast);
      };
      // pass 1 (only with constant weights): fill the ring with weight tiles before the dependency wait
      int pre = 0;
      if (w_early) {
        const int depth = w_early < S ? w_early : S;
        for (int u = blockIdx.x; u < units && pre < depth; u += gridDim.x) {
          const int tile = u % tiles, split = u / tiles;
          for (int kb = kb_begin(split); kb < kb_begin(split + 1) && pre < depth; ++kb, ++pre) {
            ptx::mbar_expect_tx(&full[pre], C::kStageBytes);
            load_w(pre, tile, kb);
          }
        }
      }
      pdl_wait();
      stamp(3);
      const bool restricted = full_flag && *full_flag == 0;
      int it = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int tile = u % tiles, split = u / tiles;
        if (skip_tile(restricted, tile)) continue;
        for (int kb = kb_begin(split); kb < kb_begin(split + 1); ++kb, ++it) {
          const int st = it % S;
          if (it < pre) { load_x(st, kb); continue; }
          ptx::mbar_wait(&empty[st], ((it / S) & 1) ^ 1);
          ptx::mbar_expect_tx(&full[st], C::kStageBytes);
          load_w(st, tile, kb);
          load_x(st, kb);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::make_idesc(KIND == 3 ? 0 : KIND, kTileA, BN);
      bool restricted = false;
      if (full_flag) { pdl_wait(); restricted = *full_flag == 0; }
      int it = 0, i = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int split = u / tiles;
        if (skip_tile(restricted, u % tiles)) continue;
        const int buf = i & 1;
        ptx::mbar_wait(&tmem_empty[buf], ((i >> 1) & 1) ^ 1);
        ptx::tc_fence_after();
        const uint32_t tacc = tmem_base + buf * BN;
        const int n = kb_begin(split + 1) - kb_begin(split);
        for (int j = 0; j < n; ++j, ++it) {
          const int st = it % S;
          ptx::mbar_wait(&full[st], (it / S) & 1);
          if (it == 0) stamp(4);
          ptx::tc_fence_after();
          const uint32_t sa = ptx::smem_u32(smem + st * C::kStageBytes);
          const uint64_t adesc = ptx::make_smem_desc_sw128(sa);
          const uint64_t bdesc = ptx::make_smem_desc_sw128(sa + C::kABytes);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (KIND != 2) ptx::mma_bf16(tacc, adesc + 2 * k, bdesc + 2 * k, idesc, (j | k) != 0);
            else ptx::mma_tf32(tacc, adesc + 2 * k, bdesc + 2 * k, idesc, (j | k) != 0);
          }
          ptx::mma_commit(&empty[st]);
        }
        ptx::mma_commit(&tmem_full[buf]);
        ++i;
      }
      stamp(5);
    }
  } else {
    // ===== epilogue: warp w reads TMEM lanes 32*(w%4).., columns [(w-2)/4 * BN/2, +BN/2) =====
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;
    constexpr int CW = BN / 2;                     // columns per warp
    constexpr int CH = CW >= 32 ? 32 : CW;         // columns per tcgen05.ld (BN=16 -> 8 is not a legal shape)
    pdl_wait();
    const bool restricted = full_flag && *full_flag == 0;
    int i = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
      const int tile = u % tiles, split = u / tiles;
      if (skip_tile(restricted, tile)) continue;
      const int buf = i & 1;
      ptx::mbar_wait(&tmem_full[buf], (i >> 1) & 1);
      if (warp == 2 && lane == 0) stamp(6);
      ptx::tc_fence_after();
      const int n = tile * kTileA + q * 32 + lane;       // output feature of this lane
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * BN + half * CW;
#pragma unroll 1
      for (int c0 = 0; c0 < CW; c0 += CH) {
        uint32_t r[CH];
        if constexpr (CH == 32) ptx::tmem_ld32(taddr + c0, r);
        else ptx::tmem_ld16(taddr + c0, r);
        ptx::tmem_ld_wait();
        if (n < rowsA) {
          const int m0 = half * CW + c0;
          if (ep.partial) {
            // hot path of every decode layer: one base pointer, predicated stores (the generic epilogue below
            // unrolls into ~9k instructions and ran out of the instruction cache: 3 us per unit)
            // (32-bit element offsets: linear_dec checks that the partial buffer holds < 2^31 floats)
            float* dst = ep.partial + (static_cast<uint32_t>(split * ep.M + m0) * static_cast<uint32_t>(ep.N) + static_cast<uint32_t>(n));
            const int ldn = ep.N;
            const int valid = rowsB - m0;
            if (valid >= CH) {
#pragma unroll
              for (int j = 0; j < CH; ++j) { *dst = __uint_as_float(r[j]); dst += ldn; }
            } else {
#pragma unroll
              for (int j = 0; j < CH; ++j) { if (j < valid) *dst = __uint_as_float(r[j]); dst += ldn; }
            }
          } else {
#pragma unroll
            for (int j = 0; j < CH; ++j) {
              const int m = m0 + j;
              if (m < rowsB) epi_store(ep, m, n, epi_apply(ep, m, n, __uint_as_float(r[j])));
            }
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tmem_empty[buf]);
      ++i;
    }
    if (warp == 2 && lane == 0) stamp(7);
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, C::kTmemCols);
  if (threadIdx.x == 0) stamp(2);
}

template <int BN, int KIND>
int launch_dec(const CUtensorMap& tmA, const CUtensorMap& tmB, const Epi& ep, int rowsA, int rowsB, int k_blocks,
               int tiles, int splits, int w_early, int w_tiled, const int* full_flag, int keep_lo, int keep_hi,
               cudaStream_t s) {
  using C = DecCfg<BN>;
  auto kern = gemm_dec_kernel<BN, KIND>;
  static bool attr_set = false;
  if (!attr_set) {
    MT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kTotal));
    attr_set = true;
  }
  const int units = tiles * splits;
  const int grid = units < kNumSMs ? units : kNumSMs;
  MT_CUDA(launch(kern, dim3(grid), dim3(kThreads), C::kTotal, s, tmA, tmB, ep, rowsA, rowsB, k_blocks, tiles, splits,
                 w_early, w_tiled, full_flag, keep_lo, keep_hi, trace_slot(kNumSMs * 8)));
  count_launch();
  return MT_OK;
}

}  // namespace

// Split count for the persistent kernel: minimise the longest CTA's work, counted in k-blocks, with every
// unit charged ~2 k-blocks for its epilogue (a [128 x BN] fp32 partial tile costs about as much traffic).
int dec_pick_splits(int tiles, int k_blocks, long long per_split_bytes, long long workspace_bytes, bool own_reduce) {
  int best = 1;
  long long best_cost = -1;
  for (int sp = 1; sp <= 32; ++sp) {
    if (sp > 1 && (k_blocks / sp < 4 || per_split_bytes * sp > workspace_bytes)) break;
    const long long rounds = (static_cast<long long>(tiles) * sp + kNumSMs - 1) / kNumSMs;
    // + sp: every split is another [M, N] fp32 slab the consumer kernel has to pull through its few SMs' L2 ports
    // (add_rmsnorm: one block per row), about one k-block of time each
    long long cost = rounds * ((k_blocks + sp - 1) / sp + 2) + sp;
    // a GEMM that has to hand back a finished tensor pays a separate reduce kernel for any split (launch + a pass
    // over all slabs, ~17 us for the audio heads): only worth it when the balance gain is large
    if (own_reduce && sp > 1) cost += 24 + 2 * sp;
    if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = sp; }
  }
  return best;
}

// CTAs the decode-orientation kernel launches for a bf16 [M,K] x [N,K]^T Linear handed to a fused consumer (split-K partials):
// every CTA owns ~200 KB of shared memory, so two of these kernels can never share an SM (lm_attn.cu's early-read rule)
int dec_grid_ctas(int M, int N, int K, long long workspace_bytes) {
  const int k_blocks = cdiv(K, 64), tiles = cdiv(N, kTileA);
  const int splits = dec_pick_splits(tiles, k_blocks, static_cast<long long>(M) * N * 4, workspace_bytes, false);
  const int units = tiles * splits;
  return units < kNumSMs ? units : kNumSMs;
}

int linear_dec(const LinearPlan& p, Epi ep, cudaStream_t s) {
  const int esz = dtype_bytes(p.dtype_in);
  const int k_elems = 128 / esz;
  const int k_blocks = cdiv(p.K, k_elems);
  int BN = 16;
  while (BN < p.M) BN *= 2;
  MT_CHECK(BN <= 256, "linear_dec: M=%d exceeds 256", p.M);
  if (BN == 16) BN = 32;   // the two-warp column split needs >= 16 columns per warp
  CUtensorMap tmA, tmB;
  if (p.w_tmap) tmA = *p.w_tmap;
  else if (p.w_tiled) MT_TRY(make_tmap_tiled(&tmA, p.w, p.dtype_in, p.N, p.K));
  else MT_TRY(make_tmap_2d(&tmA, p.w, p.dtype_in, p.N, p.K, p.ldw, kTileA));
  MT_TRY(make_tmap_2d(&tmB, p.x, p.dtype_in, p.M, p.K, p.ldx, BN));
  const int tiles = cdiv(p.N, kTileA);
  const long long per_split = static_cast<long long>(p.M) * p.N * 4;
  int splits = 1;
  if (p.workspace) splits = dec_pick_splits(tiles, k_blocks, per_split, p.workspace_bytes, p.out_splits == nullptr);
  MT_CHECK(per_split / 4 * splits < (1ll << 31), "linear_dec: split-K partials exceed 2^31 elements");
  if (p.out_splits) {
    MT_CHECK(p.workspace && per_split <= p.workspace_bytes, "linear: workspace too small for partial output");
    *p.out_splits = splits;
    ep.partial = p.workspace;
  } else if (splits > 1) {
    ep.partial = p.workspace;
  }
  const int w_tiled = p.w_tiled ? 1 : 0;
  // a filtered launch decides its unit list after the dependency wait: no early weight fetch
  // w_early = number of ring stages requested before the dependency wait (capped at the ring depth).
  // Measured: any early depth other than the full ring is within +-1 %.
  const int w_early = p.w_static && !p.tile_filter_flag ? 64 : 0;
  const int keep_lo = p.tile_lo / kTileA, keep_hi = p.tile_hi / kTileA;
  if (p.tile_filter_flag) MT_CHECK(splits == 1 && !p.out_splits, "linear_dec: the row-tile filter needs an unsplit GEMM");
  const int kind = mma_kind(p.dtype_in);
#define MT_DEC(BNV)                                                                                               \
  case BNV:                                                                                                       \
    if (kind == 1) MT_TRY((launch_dec<BNV, 1>(tmA, tmB, ep, p.N, p.M, k_blocks, tiles, splits, w_early, w_tiled, p.tile_filter_flag, keep_lo, keep_hi, s)));    \
    else if (kind == 2) MT_TRY((launch_dec<BNV, 2>(tmA, tmB, ep, p.N, p.M, k_blocks, tiles, splits, w_early, w_tiled, p.tile_filter_flag, keep_lo, keep_hi, s)));       \
    else MT_TRY((launch_dec<BNV, 3>(tmA, tmB, ep, p.N, p.M, k_blocks, tiles, splits, w_early, w_tiled, p.tile_filter_flag, keep_lo, keep_hi, s)));       \
    break;
  switch (BN) {
    MT_DEC(32) MT_DEC(64) MT_DEC(128) MT_DEC(256)
    default: return fail(MT_ERR_ARG, "unsupported BN %d", BN);
  }
#undef MT_DEC
  if (ep.partial && !p.out_splits) MT_TRY(splitk_reduce(ep, splits, s));
  return MT_OK;
}

}  // namespace mt
