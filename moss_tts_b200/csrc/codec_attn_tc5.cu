// Codec attention on the 5th-generation tensor cores (reference MossAudioTokenizerMultiheadAttention.forward,
// moss_audio_tokenizer/modeling_moss_audio_tokenizer.py:627-666 with the interleaved-pair RoPE of :255-300): causal
// sliding-window attention, head_dim 64, IEEE-half operands, fp32 accumulation and softmax.
//
// FlashAttention forward restated for Blackwell: S = Q.K^T and O_j = P_j.V_j are tcgen05.mma instructions issued by one
// thread, their accumulators live in tensor memory, Q / K / V tiles arrive by TMA straight out of the in_proj GEMM's
// [tokens, 3d] output (one tensor map, three column offsets; 128-byte swizzle), and the probabilities go back to the tensor
// core THROUGH tensor memory (A operand of P.V read from TMEM), never through shared memory or registers of another warp.
//
// CTA = 128 consecutive queries of one (sample, head) walking 64-key tiles; FOUR CTAs per SM (128 TMEM columns, 48 KB of shared
// memory, < 86 registers each), so four independent S -> softmax -> P.V chains hide each other's hand-off latencies:
//   warps 0-3  softmax: thread t owns TMEM lane t = query row t.  Per tile: S (fp32, 64 columns) is read twice (row max, then
//              p = exp2(s * scale*log2e - m)) and P is written as packed halves OVER the first 32 columns of S (a column is
//              overwritten only after this thread has read it).  O stays in tensor memory: P_j.V_j ACCUMULATES into it, and
//              the running maximum m is only raised -- with O and the row sum rescaled by 2^(m_old - m_new) through a TMEM
//              load / store -- when a tile exceeds it by more than 2^8 (p <= 256 is harmless in half precision); the final
//              o = O / l is exact for any such stale maximum
//   warp 4     TMA producer: Q once, then K_j / V_j into a 2-stage ring (2 x 8 KB per stage)
//   warp 5     TMEM allocator + MMA issuer: P_j.V_j, then S_{j+1} = Q.K_{j+1}^T (tcgen05.mma instructions of one thread execute
//              in issue order, so S_{j+1} overwrites the S / P columns only after P_j.V_j has consumed them)
// V is consumed as the MN-major B operand (a [64 keys x 64 dims] tile is 64 keys of one 128-byte row each: exactly the
// canonical MN-major SWIZZLE_128B layout with an 8-key group stride of 1024 B), so no transpose of V is ever made.
// q and k must already carry their rotary embedding (codec_rope_qk_f16 rotates them in place after the in_proj GEMM).
//
// visible(q, k) = lo(q) <= k <= q with lo(q) = max(q - context + 1, chunk bound, 0): the same closed form as
// codec_attention_f16_kernel (codec_attn.cu), which stays the kernel of the incremental (ring-cache) streaming path.
#include "codec_ops.cuh"
#include "ptx.cuh"

namespace mt {

namespace {

constexpr int kBQ = 128, kBK = 64, kD = 64, kStages = 2;
constexpr int kQBytes = kBQ * kD * 2;                          // 16 KB: 128 rows of 128 bytes
constexpr int kTileBytes = kBK * kD * 2;                       // 8 KB: 64 rows of 128 bytes
constexpr int kThreads = 192;
constexpr int kOffK = kQBytes, kOffV = kOffK + kStages * kTileBytes, kOffBar = kOffV + kStages * kTileBytes;
constexpr int kNumBars = 1 + 3 * kStages + 3;
constexpr int kSmemBytes = kOffBar + kNumBars * 8 + 16 + 1024;      // + alignment slack: ~50 KB, four CTAs per SM
constexpr uint32_t kTmemCols = 128;
constexpr uint32_t kColS = 0, kColP = 0, kColO = 64;            // TMEM columns: S x 64 (P over its first 32) | O x 64
constexpr float kRescaleThreshold = 8.f;                        // log2 units: the maximum is raised only past 2^8

__device__ __forceinline__ int vis_lo5(int t, int T, int context, int chunk) {
  int lo = t - context + 1;
  if (chunk > 0 && T > chunk) {
    const int start = (t / chunk) * chunk;
    const int len = min(chunk, T - start);
    lo = max(lo, start + len - context);
  }
  return max(lo, 0);
}

__device__ __forceinline__ void mma_f16_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
// A operand (P) read from tensor memory
__device__ __forceinline__ void mma_f16_ts(uint32_t d, uint32_t a_tmem, uint64_t b, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
               ::"r"(d), "r"(a_tmem), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
               "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
               ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
                 "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
               : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__global__ void __launch_bounds__(kThreads, 4)
codec_attention_tc5_kernel(const __grid_constant__ CUtensorMap tmq, const __grid_constant__ CUtensorMap tmkv,
                           __half* __restrict__ out, int T, int heads, int context, int chunk, float scale_log2e) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* q_full = bars;
  uint64_t* k_full = q_full + 1;
  uint64_t* v_full = k_full + kStages;
  uint64_t* kv_empty = v_full + kStages;
  uint64_t* s_full = kv_empty + kStages;     // MMA -> softmax
  uint64_t* p_full = s_full + 1;             // softmax -> MMA (S read, P written, O rescaled if needed)
  uint64_t* o_full = p_full + 1;             // MMA -> softmax: P_j.V_j has been added to O
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(o_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // heavy tiles (late queries see the most keys) first
  const int q0 = (static_cast<int>(gridDim.x) - 1 - static_cast<int>(blockIdx.x)) * kBQ;
  const int h = blockIdx.y, b = blockIdx.z;
  const int d = heads * kD;
  const int tok0 = b * T;
  const int last_q = min(q0 + kBQ, T) - 1;
  const int k_begin = (vis_lo5(q0, T, context, chunk) / kBK) * kBK;
  const int n_tiles = (last_q - k_begin) / kBK + 1;

  if (warp == 4 && lane == 0) {
    ptx::prefetch_tmap(&tmq);
    ptx::prefetch_tmap(&tmkv);
    ptx::mbar_init(q_full, 1);
    for (int i = 0; i < kStages; ++i) { ptx::mbar_init(&k_full[i], 1); ptx::mbar_init(&v_full[i], 1); ptx::mbar_init(&kv_empty[i], 1); }
    ptx::mbar_init(s_full, 1);
    ptx::mbar_init(p_full, 4);
    ptx::mbar_init(o_full, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 5) {
    ptx::tmem_alloc(tmem_holder, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  if (threadIdx.x == 0) pdl_trigger();

  if (warp == 4) {
    // ===== TMA producer =====
    if (lane == 0) {
      pdl_wait();
      ptx::mbar_expect_tx(q_full, kQBytes);
      ptx::tma_load_2d(smem, &tmq, q_full, h * kD, tok0 + q0, ptx::kEvictNormal);
      for (int j = 0; j < n_tiles; ++j) {
        const int st = j % kStages;
        if (j >= kStages) ptx::mbar_wait_suspend(&kv_empty[st], ((j / kStages) & 1) ^ 1);
        const int row = tok0 + k_begin + j * kBK;
        ptx::mbar_expect_tx(&k_full[st], kTileBytes);
        ptx::tma_load_2d(smem + kOffK + st * kTileBytes, &tmkv, &k_full[st], d + h * kD, row, ptx::kEvictLast);
        ptx::mbar_expect_tx(&v_full[st], kTileBytes);
        ptx::tma_load_2d(smem + kOffV + st * kTileBytes, &tmkv, &v_full[st], 2 * d + h * kD, row, ptx::kEvictLast);
      }
    }
  } else if (warp == 5) {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint32_t idesc_s = ptx::make_idesc(0, kBQ, kBK);                 // F16 x F16 -> F32, both K-major
      constexpr uint32_t idesc_o = ptx::make_idesc(0, kBQ, kD) | (1u << 16);     // B (= V) MN-major
      const uint64_t qdesc = ptx::make_smem_desc_sw128(ptx::smem_u32(smem));
      auto issue_s = [&](int j) {
        const int st = j % kStages;
        ptx::mbar_wait_suspend(&k_full[st], (j / kStages) & 1);
        ptx::tc_fence_after();
        const uint64_t kdesc = ptx::make_smem_desc_sw128(ptx::smem_u32(smem + kOffK + st * kTileBytes));
#pragma unroll
        for (int k = 0; k < kD / 16; ++k) mma_f16_ss(tmem_base + kColS, qdesc + 2 * k, kdesc + 2 * k, idesc_s, k != 0);
        ptx::mma_commit(s_full);
      };
      ptx::mbar_wait_suspend(q_full, 0);
      issue_s(0);
      for (int j = 0; j < n_tiles; ++j) {
        const int st = j % kStages;
        ptx::mbar_wait_suspend(&v_full[st], (j / kStages) & 1);
        ptx::mbar_wait_suspend(p_full, j & 1);
        ptx::tc_fence_after();
        const uint64_t vdesc = ptx::make_smem_desc_sw128(ptx::smem_u32(smem + kOffV + st * kTileBytes));
#pragma unroll
        for (int k = 0; k < kBK / 16; ++k)      // 16 keys per instruction: 8 packed-half columns of P, 2 KB of V rows
          mma_f16_ts(tmem_base + kColO, tmem_base + kColP + 8 * k, vdesc + 128 * k, idesc_o, (j | k) != 0);
        ptx::mma_commit(o_full);
        ptx::mma_commit(&kv_empty[st]);
        if (j + 1 < n_tiles) issue_s(j + 1);    // in issue order behind P_j.V_j: the S / P columns are free by then
      }
    }
  } else {
    // ===== softmax + output: thread = query row = TMEM lane =====
    const int row = q0 + warp * 32 + lane;
    const int lo = vis_lo5(min(row, T - 1), T, context, chunk);
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
    const uint32_t s_addr = lane_addr + kColS;
    float m = -INFINITY, l = 0.f;
    // a tile needs the element mask when it touches the diagonal or the window edge of some row of the CTA
    const int lo_max = vis_lo5(min(q0 + kBQ - 1, T - 1), T, context, chunk);

    for (int j = 0; j < n_tiles; ++j) {
      const int k0 = k_begin + j * kBK;
      const bool masked = (k0 + kBK - 1 > q0) || (k0 < lo_max);
      ptx::mbar_wait_suspend(s_full, j & 1);
      ptx::tc_fence_after();
      // pass 1: row maximum of the visible scores
      float mx = -INFINITY;
#pragma unroll
      for (int c0 = 0; c0 < kBK; c0 += 32) {
        uint32_t r[32];
        ptx::tmem_ld32(s_addr + c0, r);
        ptx::tmem_ld_wait();
        if (masked) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const int k = k0 + c0 + i;
            if (k >= lo && k <= row) mx = fmaxf(mx, __uint_as_float(r[i]));
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(r[i]));
        }
      }
      mx *= scale_log2e;
      // raise the running maximum only when this tile exceeds it by more than 2^8: O (in TMEM) and l are rescaled then
      const bool raise = mx > m + kRescaleThreshold || (m == -INFINITY && mx > -INFINITY);
      if (__any_sync(0xffffffffu, raise)) {
        const float alpha = raise ? ex2(m - mx) : 1.f;          // m = -inf: 0 (nothing accumulated yet)
        if (raise) { m = mx; l *= alpha; }
        if (j > 0) {
          ptx::mbar_wait_suspend(o_full, (j - 1) & 1);                  // P_{j-1}.V_{j-1} has landed in O
          ptx::tc_fence_after();
#pragma unroll
          for (int c0 = 0; c0 < kD; c0 += 16) {
            uint32_t r[16];
            ptx::tmem_ld16(lane_addr + kColO + c0, r);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
            tmem_st16(lane_addr + kColO + c0, r);
          }
        }
      }
      const float m_use = m == -INFINITY ? 0.f : m;             // a row that has seen nothing yet: everything stays zero
      // pass 2: p = 2^(s * scale * log2(e) - m), packed halves -> TMEM (the A operand of P.V)
      float sum = 0.f;
#pragma unroll
      for (int c0 = 0; c0 < kBK; c0 += 32) {
        uint32_t r[32];
        ptx::tmem_ld32(s_addr + c0, r);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = ex2(__uint_as_float(r[i]) * scale_log2e - m_use);
          float p1 = ex2(__uint_as_float(r[i + 1]) * scale_log2e - m_use);
          if (masked) {
            const int k = k0 + c0 + i;
            if (k < lo || k > row) p0 = 0.f;
            if (k + 1 < lo || k + 1 > row) p1 = 0.f;
          }
          sum += p0 + p1;
          r[i >> 1] = pack_f16(p0, p1);                      // packed in place: r[0..15] become the 16 P columns
        }
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) pk[i] = r[i];
        tmem_st16(lane_addr + kColP + (c0 >> 1), pk);        // over S columns this thread has already consumed
      }
      tmem_st_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(p_full);
      l += sum;
    }
    // ---- o = O / l
    ptx::mbar_wait_suspend(o_full, (n_tiles - 1) & 1);
    ptx::tc_fence_after();
    pdl_wait();                               // `out` may still be read by the kernel before the producer of qkv
    const float inv = 1.f / l;
    __half* dst = out + static_cast<long long>(tok0 + row) * d + h * kD;
#pragma unroll
    for (int c0 = 0; c0 < kD; c0 += 16) {
      uint32_t r[16];
      ptx::tmem_ld16(lane_addr + kColO + c0, r);
      ptx::tmem_ld_wait();
      if (row < T) {
        uint4 v0, v1;
        v0.x = pack_f16(__uint_as_float(r[0]) * inv, __uint_as_float(r[1]) * inv);
        v0.y = pack_f16(__uint_as_float(r[2]) * inv, __uint_as_float(r[3]) * inv);
        v0.z = pack_f16(__uint_as_float(r[4]) * inv, __uint_as_float(r[5]) * inv);
        v0.w = pack_f16(__uint_as_float(r[6]) * inv, __uint_as_float(r[7]) * inv);
        v1.x = pack_f16(__uint_as_float(r[8]) * inv, __uint_as_float(r[9]) * inv);
        v1.y = pack_f16(__uint_as_float(r[10]) * inv, __uint_as_float(r[11]) * inv);
        v1.z = pack_f16(__uint_as_float(r[12]) * inv, __uint_as_float(r[13]) * inv);
        v1.w = pack_f16(__uint_as_float(r[14]) * inv, __uint_as_float(r[15]) * inv);
        *reinterpret_cast<uint4*>(dst + c0) = v0;
        *reinterpret_cast<uint4*>(dst + c0 + 8) = v1;
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 5) ptx::tmem_dealloc(tmem_base, kTmemCols);
}

// interleaved-pair RoPE (:255-300) of q and k in place: qkv [M, 3d] half, table row = token index inside its sample
__global__ void codec_rope_qk_f16_kernel(__half* __restrict__ qkv, int T, int d, const float4* __restrict__ rope, long long total) {
  pdl_trigger();
  pdl_wait();
  const int g8 = 2 * d / 8;                       // 8-half groups of the q | k part of a row
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long tok = idx / g8;
    const int c8 = static_cast<int>(idx % g8);
    const int t = static_cast<int>(tok % T);
    const int c = c8 % (kD / 8);                  // group inside the head
    __half* p = qkv + tok * 3 * d + 8 * c8;
    const uint4 raw = *reinterpret_cast<const uint4*>(p);
    const float4 cs0 = rope[static_cast<long long>(t) * (kD / 4) + 2 * c], cs1 = rope[static_cast<long long>(t) * (kD / 4) + 2 * c + 1];
    const float2 p0 = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
    const float2 p1 = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
    const float2 p2 = __half22float2(*reinterpret_cast<const __half2*>(&raw.z));
    const float2 p3 = __half22float2(*reinterpret_cast<const __half2*>(&raw.w));
    uint4 o;
    o.x = pack_f16(p0.x * cs0.x - p0.y * cs0.y, p0.x * cs0.y + p0.y * cs0.x);
    o.y = pack_f16(p1.x * cs0.z - p1.y * cs0.w, p1.x * cs0.w + p1.y * cs0.z);
    o.z = pack_f16(p2.x * cs1.x - p2.y * cs1.y, p2.x * cs1.y + p2.y * cs1.x);
    o.w = pack_f16(p3.x * cs1.z - p3.y * cs1.w, p3.x * cs1.w + p3.y * cs1.z);
    *reinterpret_cast<uint4*>(p) = o;
  }
}

}  // namespace

bool codec_attention_tc5_supported(int hd, long long rows, int d) {
  return hd == kD && rows > 0 && (static_cast<long long>(3) * d * 2) % 16 == 0;
}

// qkv is rotated IN PLACE (q and k), then attended: out [B*T, heads*64] half
int codec_attention_tc5(void* qkv, void* out, int B, int T, int heads, int hd, int context, int chunk, cudaStream_t s,
                        const float* rope_table, bool pre_rotated) {
  MT_CHECK(hd == kD && rope_table, "codec_attention_tc5: head_dim 64 and a RoPE table are required");
  const int d = heads * kD;
  const long long M = static_cast<long long>(B) * T;
  if (!pre_rotated) {
    const long long total = M * (2 * d / 8);
    int blocks = static_cast<int>(cdivll(total, 256));
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    MT_CUDA(launch(codec_rope_qk_f16_kernel, dim3(blocks), dim3(256), 0, s, reinterpret_cast<__half*>(qkv), T, d,
                   reinterpret_cast<const float4*>(rope_table), total));
    count_launch();
  }
  // tensor maps per (buffer, shape): Q boxes of 128 rows, K / V boxes of 64 rows (the engine calls with the same qkv buffer
  // for every layer of a stage)
  struct Cached { const void* base; long long M; int d; CUtensorMap tmq, tmkv; };
  static thread_local Cached cache[8];
  static thread_local int next = 0;
  const Cached* tm = nullptr;
  for (auto& c : cache)
    if (c.base == qkv && c.M == M && c.d == d) { tm = &c; break; }
  if (!tm) {
    Cached& c = cache[next++ % 8];
    c.base = nullptr;
    MT_TRY(make_tmap_2d(&c.tmq, qkv, MT_DTYPE_F16, M, 3ll * d, 3ll * d, kBQ));
    MT_TRY(make_tmap_2d(&c.tmkv, qkv, MT_DTYPE_F16, M, 3ll * d, 3ll * d, kBK));
    c.base = qkv; c.M = M; c.d = d;
    tm = &c;
  }
  static bool attr_set = false;
  if (!attr_set) {
    MT_CUDA(cudaFuncSetAttribute(codec_attention_tc5_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_set = true;
  }
  const float scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(hd));
  MT_CUDA(launch(codec_attention_tc5_kernel, dim3(cdiv(T, kBQ), heads, B), dim3(kThreads), kSmemBytes, s, tm->tmq, tm->tmkv,
                 reinterpret_cast<__half*>(out), T, heads, context, chunk, scale_log2e));
  count_launch();
  return MT_OK;
}

}  // namespace mt
