// Two-SM form of the token-major GEMM (gemm_big.cu): a CTA pair (cluster of 2, the two SMs of a TPC) owns a
// 256 (tokens) x 256 (features) output tile and multiplies it with tcgen05.mma.cta_group::2 (UMMA 256 x 256 x 16).
//   y[M,N] = epilogue(x[M,K] . w[N,K]^T),  bf16 / fp16 (kind::f16) or fp32 storage with TF32 math (kind::tf32).
//
// Why: the one-SM kernel pulls (128 + 256) rows x 128 B of operands into shared memory per 128 x 256 x 64 k-block; at the
// codec's short K that operand stream (26 TB/s at full tensor rate over 148 SMs), not the tensor pipe, is the limit
// (ncu: tensor pipe 48-58 %).  In a pair every CTA loads ITS 128 token rows and ITS half (128 rows) of the weight tile and
// the MMA reads the other half from the peer's shared memory: (128 + 128) rows per CTA and k-block, a third less traffic
// from L2, and six 32 KB stages instead of four 48 KB ones.
//
// Per CTA (both ranks): warp 0 TMA producer, warp 1 TMEM allocator (+ on rank 0 the single MMA-issuing thread of the
// pair), warps 2..9 epilogue of the CTA's own 128 accumulator lanes.  Barriers:
//   full[s]        rank 0 only: 1 arrival (rank 0's expect_tx of both CTAs' bytes) + the bytes of both CTAs' TMA loads
//                  (the peer's loads name rank 0's barrier: cp.async.bulk.tensor ... cta_group::2)
//   empty[s]       per CTA: tcgen05.commit.cta_group::2 multicast to both CTAs once the MMAs reading stage s are done
//   tmem_full[b]   per CTA: commit multicast after the last k-block of a tile
//   tmem_empty[b]  rank 0 only: 2 x 8 epilogue warps (the peer's arrive through the cluster address of rank 0's barrier)
#include "gemm.cuh"
#include "gemm_epi.cuh"
#include "ptx.cuh"

namespace mt {

int g_gemm_pair_mode = 0;        // mt_gemm_debug_set_pair: 0 = by shape (product), 1 = pairs wherever M allows, 4 = never; + 2 = per-lane row stores

namespace {

constexpr int kPM = 128, kPN = 128;                  // rows each CTA loads per k-block: its tokens, its half of the weight tile
constexpr int kTileM = 256, kTileN = 256;            // the pair's output tile
constexpr int kKBytes = 128, kStages = 6;
constexpr int kEpiWarps = 8;
constexpr int kThreads = 64 + 32 * kEpiWarps;
constexpr int kABytes = kPM * kKBytes, kBBytes = kPN * kKBytes, kStageBytes = kABytes + kBBytes;
constexpr int kBarOffset = kStages * kStageBytes;
constexpr int kEpiOffset = kBarOffset + (2 * kStages + 4) * 8 + 32;     // 8 warps x 4 KB: epilogue transposition tiles (16-byte aligned)
constexpr int kSmemTotal = kEpiOffset + kEpiWarps * kEpiTileFloats * 4 + 1024;
static_assert(kSmemTotal <= 232448, "shared memory budget");

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `p` (an object of this CTA's shared memory) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_rank(const void* p, uint32_t rank) {
  uint32_t a;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(a) : "r"(ptx::smem_u32(p)), "r"(rank));
  return a;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load of one CTA of a pair: data into this CTA's shared memory, bytes signalled on the barrier at `bar_cluster_addr`
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint32_t bar_cluster_addr, int c0, int c1,
                                                 uint64_t hint) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(ptx::smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "l"(hint)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t* smem_holder, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(ptx::smem_u32(smem_holder)), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish_pair() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
template <int KIND>
__device__ __forceinline__ void mma_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  if constexpr (KIND == 2) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// arrive on the barrier at this shared-memory offset in BOTH CTAs once the pair's earlier MMAs have completed
__device__ __forceinline__ void mma_commit_pair(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(ptx::smem_u32(bar)), "h"(static_cast<uint16_t>(3))
               : "memory");
}

template <int KIND /*1 bf16, 2 tf32, 3 fp16*/>
__global__ void __launch_bounds__(kThreads, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, Epi ep, int M, int N,
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
  const uint32_t rank = cluster_rank();
  const int pair = blockIdx.x >> 1, pairs = gridDim.x >> 1;
  // the pair's tile walk: gemm_big.cu's banded order with pairs in place of CTAs (a pair keeps ITS 256 tokens while whole
  // bands remain and walks the feature tiles; the ragged rest token-fastest)
  const int full_bands = m_tiles / pairs;
  const int rem_m = m_tiles - full_bands * pairs;
  auto tile_of = [&](int i, int& mt, int& nt) -> bool {
    const int banded = full_bands * n_tiles;
    if (i < banded) { mt = (i / n_tiles) * pairs + pair; nt = i % n_tiles; return true; }
    const int t = pair + (i - banded) * pairs;
    if (t >= rem_m * n_tiles) return false;
    mt = full_bands * pairs + t % rem_m; nt = t / rem_m;
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
      ptx::mbar_init(&tmem_empty[i], 2 * kEpiWarps);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc_pair(tmem_holder, 512);
    tmem_relinquish_pair();
  }
  ptx::tc_fence_before();
  cluster_sync_all();                          // both CTAs' barriers exist before either one's TMA or commit names them
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  if (threadIdx.x == 0) pdl_trigger();

  if (warp == 0) {
    // ===== TMA producer (both CTAs) =====
    if (lane == 0) {
      uint32_t full0[kStages];                 // rank 0's full barriers
#pragma unroll
      for (int i = 0; i < kStages; ++i) full0[i] = map_to_rank(&full[i], 0);
      auto load_a = [&](int st, int mt, int kb) {
        tma_load_2d_pair(smem + st * kStageBytes, &tmA, full0[st], kb * kKElems, mt * kTileM + static_cast<int>(rank) * kPM,
                         ptx::kEvictNormal);
      };
      auto load_b = [&](int st, int nt, int kb) {
        uint8_t* sb = smem + st * kStageBytes + kABytes;
        const int half_tile = nt * (kTileN / kPN) + static_cast<int>(rank);     // this CTA's 128 weight rows
        if (w_tiled) tma_load_2d_pair(sb, &tmB, full0[st], 0, (half_tile * k_blocks + kb) * 128, ptx::kEvictLast);
        else tma_load_2d_pair(sb, &tmB, full0[st], kb * kKElems, half_tile * kPN, ptx::kEvictLast);
      };
      int it = 0, pre = 0, mt = 0, nt = 0;
      if (w_early && tile_of(0, mt, nt)) {
        // constant weights: first ring fill before the previous kernel is known to have finished
        const int nt0 = nt;
        pre = k_blocks < kStages ? k_blocks : kStages;
        for (int i = 0; i < pre; ++i) {
          if (rank == 0) ptx::mbar_expect_tx(&full[i], 2 * kStageBytes);
          load_b(i, nt0, i);
        }
      }
      pdl_wait();
      for (int i = 0; tile_of(i, mt, nt); ++i) {
        for (int kb = 0; kb < k_blocks; ++kb, ++it) {
          const int st = it % kStages;
          if (it < pre) { load_a(st, mt, kb); continue; }
          ptx::mbar_wait(&empty[st], ((it / kStages) & 1) ^ 1);
          if (rank == 0) ptx::mbar_expect_tx(&full[st], 2 * kStageBytes);
          load_a(st, mt, kb);
          load_b(st, nt, kb);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one thread of rank 0 for the pair =====
    if (lane == 0 && rank == 0) {
      constexpr uint32_t idesc = ptx::make_idesc(KIND == 3 ? 0 : KIND, kTileM, kTileN);
      int it = 0, mt = 0, nt = 0;
      for (int i = 0; tile_of(i, mt, nt); ++i) {
        const int buf = i & 1;
        ptx::mbar_wait(&tmem_empty[buf], ((i >> 1) & 1) ^ 1);     // both CTAs' epilogues drained this accumulator
        ptx::tc_fence_after();
        const uint32_t tacc = tmem_base + buf * kTileN;
        for (int kb = 0; kb < k_blocks; ++kb, ++it) {
          const int st = it % kStages;
          ptx::mbar_wait(&full[st], (it / kStages) & 1);
          ptx::tc_fence_after();
          const uint32_t sa = ptx::smem_u32(smem + st * kStageBytes);
          const uint64_t adesc = ptx::make_smem_desc_sw128(sa);
          const uint64_t bdesc = ptx::make_smem_desc_sw128(sa + kABytes);
#pragma unroll
          for (int k = 0; k < 4; ++k) mma_pair<KIND>(tacc, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
          mma_commit_pair(&empty[st]);
        }
        mma_commit_pair(&tmem_full[buf]);
      }
    }
  } else {
    // ===== epilogue (both CTAs): this CTA's 128 tokens x 256 features =====
    const int q = warp & 3;                    // TMEM lane quarter this warp may access
    const int half = (warp - 2) >> 2;          // column half
    uint32_t tmem_empty0[2];
    tmem_empty0[0] = map_to_rank(&tmem_empty[0], 0);
    tmem_empty0[1] = map_to_rank(&tmem_empty[1], 0);
    float* tile = reinterpret_cast<float*>(smem + kEpiOffset) + (warp - 2) * kEpiTileFloats;
    const bool coalesced = epi_coalesced_ok(ep);
    pdl_wait();                                // residual reads / output writes follow the previous kernel
    int mt = 0, nt = 0;
    for (int i = 0; tile_of(i, mt, nt); ++i) {
      const int buf = i & 1;
      const int m0 = mt * kTileM + static_cast<int>(rank) * kPM + q * 32;
      const int n_base = nt * kTileN + half * (kTileN / 2);
      EpiRes res;
      if (coalesced && n_base + 32 <= N && m0 < M) epi_res_load(ep, m0, n_base, M, lane, res);   // lands behind the wait
      ptx::mbar_wait(&tmem_full[buf], (i >> 1) & 1);
      ptx::tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + buf * kTileN + half * (kTileN / 2);
      epi_tile_warp<kTileN / 2>(ep, coalesced, m0, n_base, M, N, taddr, tile, lane, res);
      // this warp is done reading the accumulator: hand it back to the pair's MMA thread
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(tmem_empty0[buf]);
    }
  }

  // the peer's shared memory and barriers are in use until the leader's last MMA / commit has landed
  ptx::tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_pair(tmem_base, 512);
}

template <int KIND>
int launch_pair(const CUtensorMap& tmA, const CUtensorMap& tmB, const Epi& ep, int M, int N, int k_blocks, int m_tiles,
                int n_tiles, int w_early, int w_tiled, cudaStream_t s) {
  auto kern = gemm_pair_kernel<KIND>;
  static bool attr_set = false;
  if (!attr_set) {
    MT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
    attr_set = true;
  }
  const int total = m_tiles * n_tiles;
  const int pairs = total < kNumSMs / 2 ? total : kNumSMs / 2;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * pairs); cfg.blockDim = dim3(kThreads); cfg.dynamicSmemBytes = kSmemTotal; cfg.stream = s;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  MT_CUDA(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, ep, M, N, k_blocks, m_tiles, n_tiles, w_early, w_tiled));
  count_launch();
  return MT_OK;
}

}  // namespace

// Which token-major Linears run on CTA pairs.  A pair needs at least one full 256-token tile (below that half of every MMA
// multiplies zero padding).  Measured on the codec's and the prefill's shapes (profiles/r02_gemm_pair_ab.jsonl): with long K
// (the main loop dominates: prefill, the codec's linear2) pairs are 1-3 % faster, the third less operand traffic showing; with
// the codec's short K = 768 / 1280 the tile time is the epilogue's and coupling two CTAs' epilogues per tile costs 5-10 %.
bool linear_pair_applies(const LinearPlan& p) {
  if (p.M < kTileM || (g_gemm_pair_mode & 4)) return false;
  if (g_gemm_pair_mode & 1) return true;
  return p.K >= 2048 && p.M >= 2 * kTileM;
}

int linear_pair(const LinearPlan& p, const Epi& ep, cudaStream_t s) {
  const int esz = dtype_bytes(p.dtype_in);
  const int k_elems = 128 / esz;
  const int k_blocks = cdiv(p.K, k_elems);
  CUtensorMap tmA, tmB;
  MT_TRY(make_tmap_2d(&tmA, p.x, p.dtype_in, p.M, p.K, p.ldx, kPM));
  if (p.w_tmap) tmB = *p.w_tmap;                                      // cached maps have the 128-row box this kernel loads
  else if (p.w_tiled) MT_TRY(make_tmap_tiled(&tmB, p.w, p.dtype_in, p.N, p.K));
  else MT_TRY(make_tmap_2d(&tmB, p.w, p.dtype_in, p.N, p.K, p.ldw, kPN));
  const int m_tiles = cdiv(p.M, kTileM), n_tiles = cdiv(p.N, kTileN);
  const int w_early = p.w_static ? 1 : 0, w_tiled = p.w_tiled ? 1 : 0;
  const int kind = mma_kind(p.dtype_in);
  if (kind == 1) return launch_pair<1>(tmA, tmB, ep, p.M, p.N, k_blocks, m_tiles, n_tiles, w_early, w_tiled, s);
  if (kind == 2) return launch_pair<2>(tmA, tmB, ep, p.M, p.N, k_blocks, m_tiles, n_tiles, w_early, w_tiled, s);
  return launch_pair<3>(tmA, tmB, ep, p.M, p.N, k_blocks, m_tiles, n_tiles, w_early, w_tiled, s);
}

}  // namespace mt
