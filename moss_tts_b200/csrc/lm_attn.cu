// Single-token (decode) GQA attention fused with the q/k RMSNorm + RoPE + KV-append step.
// One 4-warp block per (batch row, kv head) -- all B*n_kv blocks are resident at once -- keys split over the warps.
//
//   prep   q (G heads), k, v rows of the new token, one warp per head: split-K partial sum -> bf16,
//          q/k RMSNorm + RoPE (same arithmetic as qkv_post_kernel, lm_ops.cu), k/v appended to the cache
//          (reference: Qwen3Attention.forward modeling_qwen3.py:252-291, cache_utils.py:102-121)
//   pass 1 S = Q.K^T on the tensor cores: mma.sync m16n8k16 bf16, A = the G query heads (rows >= G are
//          zero), B fragments read straight from the K cache (one fully used 32-byte sector per key row
//          and k-step); s = bf16(bf16(acc) * scale), masked keys -> -inf             (:196-217)
//   soft   fp32 max / exp / sum over the visible keys (one warp per head), p = bf16(e / sum)
//   pass 2 O = P.V, again mma.sync; V is read as 32-bit words (two adjacent dims of one key) and the two
//          halves of four words are permuted into the B fragments of TWO n-tiles holding the even and the
//          odd dims, so every V sector is used exactly once and no transpose through shared memory is
//          needed.  Per-warp partial outputs are summed in warp order; out = bf16(O).
// Rounding points equal the reference's eager path, like attention_kernel (lm_ops.cu), which stays the
// kernel for prefill (T > 1) and for contexts whose scores do not fit in shared memory.
#include "lm_ops.cuh"

#include <cooperative_groups.h>
#include <stdlib.h>

namespace mt {

namespace {

__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

constexpr int kMaxG = 8, kWarps = 4;

// q | per-warp partial outputs | (cluster variant only) per-CTA max, sum and partial output read by the peer CTAs
template <int D>
__host__ __device__ constexpr size_t attn_fixed_bytes(bool cl = false) {
  return static_cast<size_t>(kMaxG) * D * 2 + static_cast<size_t>(kWarps) * kMaxG * D * 4 +
         (cl ? static_cast<size_t>(2 * kMaxG + kMaxG * D) * 4 : 0);
}

// CL = true: a thread-block cluster of `cs` CTAs shares one (batch row, kv head); CTA r owns the keys
// [r * per, (r + 1) * per) (per = a multiple of 16), the softmax max / sum and the partial outputs are exchanged
// through distributed shared memory.  For small batches with long contexts (B * n_kv blocks would leave most SMs idle
// while every block walks thousands of keys alone).
template <int D, bool CL>
__global__ void __launch_bounds__(kWarps * 32, 4)
attn_decode_kernel(const bf16* __restrict__ qkv, const float* __restrict__ partials, int nsplit, long long split_stride,
                   const bf16* __restrict__ q_gain, const bf16* __restrict__ k_gain, const bf16* __restrict__ rope_cos,
                   const bf16* __restrict__ rope_sin, bf16* __restrict__ k_cache, bf16* __restrict__ v_cache,
                   const uint8_t* __restrict__ mask, long long mask_ld, bf16* __restrict__ out, int n_heads, int n_kv,
                   const int* __restrict__ pos_dev, int pos_host, int spad, int max_seq, float eps, float scale,
                   unsigned long long* __restrict__ trace, int exp_flags, int cs, int early_pos) {
  extern __shared__ __align__(16) uint8_t sm_raw[];
  constexpr int E = D / 32;
  const int G = n_heads / n_kv;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  // q [kMaxG][D] bf16 | partial outputs [kWarps][kMaxG][D] f32 | scores [G][spad] f32 | probabilities [G][spad] bf16
  bf16* sq = reinterpret_cast<bf16*>(sm_raw);
  float* red = reinterpret_cast<float*>(sm_raw + kMaxG * D * 2);
  float* cl_max = red + kWarps * kMaxG * D;            // [kMaxG]       (cluster variant)
  float* cl_sum = cl_max + kMaxG;                      // [kMaxG]
  float* cl_o = cl_sum + kMaxG;                        // [kMaxG][D]
  float* sc = reinterpret_cast<float*>(sm_raw + attn_fixed_bytes<D>(CL));
  bf16* sp = reinterpret_cast<bf16*>(sm_raw + attn_fixed_bytes<D>(CL) + static_cast<size_t>(G) * spad * 4);
  const int item = CL ? blockIdx.x / cs : blockIdx.x;
  const int rank = CL ? blockIdx.x % cs : 0;           // == cluster.block_rank() for a 1-D grid of (cs, 1, 1) clusters

  // timeline instrumentation (mt_debug_set_trace; null in every product run)
  auto stamp = [&](int e) {
    if (trace && threadIdx.x == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      trace[blockIdx.x * 8 + e] = t;
    }
  };
  stamp(0);
  pdl_trigger();
  const int b = item / n_kv, kvh = item % n_kv;
  const int heads_total = n_heads + 2 * n_kv;
  const long long row_w = static_cast<long long>(heads_total) * D;
  constexpr int kSlots = (kMaxG + 2 + kWarps - 1) / kWarps;
  long long idx0[kSlots];
  bool live[kSlots];
#pragma unroll
  for (int sl = 0; sl < kSlots; ++sl) {
    const int i = warp + sl * kWarps;
    live[sl] = i < G + 2;
    const int hh = i < G ? kvh * G + i : (i == G ? n_heads + kvh : n_heads + n_kv + kvh);
    idx0[sl] = b * row_w + static_cast<long long>(hh) * D + lane;
  }
  float gq[E], gk[E];
#pragma unroll
  for (int e = 0; e < E; ++e) {
    gq[e] = bf2f(q_gain[lane + 32 * e]);
    gk[e] = bf2f(k_gain[lane + 32 * e]);
  }
  bf16* kbase = k_cache + (static_cast<long long>(b) * n_kv + kvh) * max_seq * D;
  bf16* vbase = v_cache + (static_cast<long long>(b) * n_kv + kvh) * max_seq * D;

  // ---- everything that depends only on the POSITION (not on this step's q/k/v): the position itself, the rotary
  // row, the K/V rows of the earlier positions.  With `early_pos` set all of it is requested BEFORE the dependency wait,
  // so these round trips -- position -> rotary row / first K tile from HBM -- overlap the producer's tail instead of
  // sitting behind it.  The caller sets it only when the writers of that data are KNOWN to have finished although this
  // kernel has not waited yet (see attn_decode in lm_ops.cuh): with programmatic dependent launch a chain of small
  // kernels can be resident, all waiting, while a kernel many launches back is still running.
  int pos = 0, nkeys = 0, kbeg = 0, kend = 0, kfirst = 0;
  bool owner = false, early_k = false;
  float rc[E], rs[E];
  // The dot product over the head dim does not care in which MMA k-slot a dim sits as long as Q uses the same
  // order: lane (g, t4) reads 16 contiguous bytes (dims 32c + 8 t4 .. +7) of its key row per chunk c, and k-step
  // s = 2c + h takes dims +4h, +4h+1 as k-slots (2 t4, 2 t4 + 1) and dims +4h+2, +4h+3 as slots (2 t4 + 8, + 9).
  // 8 LDG.128 per 16-key tile instead of 32 LDG.32 that each touched 8 rows (the LSU was the bottleneck).
  uint4 kq[2][D / 32];
  auto load_k = [&](int k0, int limit) {
#pragma unroll
    for (int n = 0; n < 2; ++n) {
      const int key = min(k0 + 8 * n + g, limit);
      const uint4* kr = reinterpret_cast<const uint4*>(kbase + static_cast<long long>(key) * D) + t4;
#pragma unroll
      for (int c = 0; c < D / 32; ++c) kq[n][c] = kr[4 * c];
    }
  };
  auto position_work = [&]() {
    pos = pos_dev ? *pos_dev : pos_host;
    nkeys = pos + 1;
    kbeg = 0; kend = nkeys;                             // this CTA's keys
    if constexpr (CL) {
      const int per = (((nkeys + cs - 1) / cs) + 15) & ~15;
      kbeg = min(rank * per, nkeys);
      kend = min(kbeg + per, nkeys);
    }
    owner = kbeg <= pos && pos < kend;                  // the CTA that appends (and reads) row `pos`
    // the K/V rows of this (row, kv head) come from HBM (the cache is far larger than L2 and 15 GB of weights
    // pass through between two visits): optionally ask L2 for all of them now
    // (one bulk prefetch per tensor: the rows [0, pos) of a (row, kv head) are contiguous)
    if (!CL && threadIdx.x < 2 && pos > 0 && !(exp_flags & 1)) {
      const bf16* p = threadIdx.x ? vbase : kbase;
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(pos * D * 2) : "memory");
    }
#pragma unroll
    for (int e = 0; e < E; ++e) {
      if (rope_cos) {
        const int jh = (lane + 32 * e) % (D / 2);
        rc[e] = bf2f(rope_cos[static_cast<long long>(pos) * (D / 2) + jh]);
        rs[e] = bf2f(rope_sin[static_cast<long long>(pos) * (D / 2) + jh]);
      } else {
        rc[e] = 1.f; rs[e] = 0.f;
      }
    }
    // first score tile of this warp: its K rows are requested now and land while the prep below runs.
    // Row `pos` (the token being appended) does not exist yet: clamp to pos-1 and patch after the barrier.
    kfirst = kbeg + 16 * warp;
    early_k = kfirst < kend && !(exp_flags & 2);
    if (early_k) load_k(kfirst, max(pos - 1, 0));
  };
  const bool early = early_pos && !(exp_flags & 4);
  if (early) position_work();
  pdl_wait();
  stamp(1);

  // ---- the q/k/v inputs of this block: a warp owns heads warp, warp + kWarps, ... (G query heads, then the k and the v
  // head); lane owns elements lane + 32e.  Every load of the first four splits is in flight before anything else is asked
  // of the memory system.
  float t0[kSlots][4][E];
#pragma unroll
  for (int sl = 0; sl < kSlots; ++sl)
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int e = 0; e < E; ++e) {
        if (partials) t0[sl][u][e] = (live[sl] && u < nsplit) ? __ldcg(partials + u * split_stride + idx0[sl] + 32 * e) : 0.f;
        else t0[sl][u][e] = (live[sl] && u == 0) ? bf2f(qkv[idx0[sl] + 32 * e]) : 0.f;
      }
  if (!early) position_work();

  // ---- prep: sum the split-K partials (split order 0, 1, 2, ... kept), RMSNorm, RoPE, append k/v
  float xs[kSlots][E];
  if (partials) {
    float a[kSlots][E];
#pragma unroll
    for (int sl = 0; sl < kSlots; ++sl)
#pragma unroll
      for (int e = 0; e < E; ++e) a[sl][e] = 0.f;
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (u < nsplit) {
#pragma unroll
        for (int sl = 0; sl < kSlots; ++sl)
#pragma unroll
          for (int e = 0; e < E; ++e) a[sl][e] += t0[sl][u][e];
      }
    for (int s0 = 4; s0 < nsplit; s0 += 4) {
      float t[kSlots][4][E];
#pragma unroll
      for (int sl = 0; sl < kSlots; ++sl)
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
          for (int e = 0; e < E; ++e)
            t[sl][u][e] = (live[sl] && s0 + u < nsplit) ? __ldcg(partials + (s0 + u) * split_stride + idx0[sl] + 32 * e) : 0.f;
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (s0 + u < nsplit) {
#pragma unroll
          for (int sl = 0; sl < kSlots; ++sl)
#pragma unroll
            for (int e = 0; e < E; ++e) a[sl][e] += t[sl][u][e];
        }
    }
#pragma unroll
    for (int sl = 0; sl < kSlots; ++sl)
#pragma unroll
      for (int e = 0; e < E; ++e) xs[sl][e] = rbf(a[sl][e]);
    if (xs[0][0] != 12345.678f) stamp(7);
  } else {
#pragma unroll
    for (int sl = 0; sl < kSlots; ++sl)
#pragma unroll
      for (int e = 0; e < E; ++e) xs[sl][e] = t0[sl][0][e];
  }
#pragma unroll
  for (int sl = 0; sl < kSlots; ++sl) {
    const int i = warp + sl * kWarps;
    if (i >= G + 2) break;
    float x[E];
#pragma unroll
    for (int e = 0; e < E; ++e) x[e] = xs[sl][e];
    if (i <= G) {
      float ss = 0.f;
#pragma unroll
      for (int e = 0; e < E; ++e) ss += x[e] * x[e];
      ss = warp_sum(ss);
      const float inv = rsqrtf(ss / static_cast<float>(D) + eps);
#pragma unroll
      for (int e = 0; e < E; ++e) x[e] = rbf((i < G ? gq[e] : gk[e]) * rbf(x[e] * inv));
      if (rope_cos) {                // MossTTSLocal's depth transformer has no rotary embedding
        float y[E];
#pragma unroll
        for (int e = 0; e < E; ++e) {
          const int j = lane + 32 * e;
          const float partner = x[(e + E / 2) % E];
          const float rot = j < D / 2 ? -partner : partner;
          y[e] = rbf(rbf(x[e] * rc[e]) + rbf(rot * rs[e]));
        }
#pragma unroll
        for (int e = 0; e < E; ++e) x[e] = y[e];
      }
    }
    bf16* dst = i < G ? sq + i * D : (i == G ? kbase : vbase) + static_cast<long long>(pos) * D;
    if (i < G || owner) {
#pragma unroll
      for (int e = 0; e < E; ++e) dst[lane + 32 * e] = f2bf(x[e]);
    }
  }
  for (int i = G * D + threadIdx.x; i < kMaxG * D; i += blockDim.x) sq[i] = f2bf(0.f);   // unused query rows
  __syncthreads();       // block-scope barrier also orders the appended k/v row before the reads below
  stamp(2);

  // ---- A fragments of Q: rows 0..7 = query heads (rows 8..15 are zero)
  uint32_t qa[D / 16][2];
#pragma unroll
  for (int s = 0; s < D / 16; ++s) {
    const int d0 = 32 * (s >> 1) + 8 * t4 + 4 * (s & 1);
    qa[s][0] = *reinterpret_cast<const uint32_t*>(sq + g * D + d0);
    qa[s][1] = *reinterpret_cast<const uint32_t*>(sq + g * D + d0 + 2);
  }
  const uint8_t* mrow = mask + b * mask_ld;

  // ---- pass 1: scores, 16 keys per warp and iteration
  for (int k0 = kfirst; k0 < kend; k0 += 16 * kWarps) {
    if (k0 != kfirst || !early_k) {
      load_k(k0, kend - 1);
    } else {
#pragma unroll
      for (int n = 0; n < 2; ++n) {
        if (k0 + 8 * n + g >= pos) {          // this lane's row is the one the prep just wrote
          const uint4* kr = reinterpret_cast<const uint4*>(kbase + static_cast<long long>(pos) * D) + t4;
#pragma unroll
          for (int c = 0; c < D / 32; ++c) kq[n][c] = kr[4 * c];
        }
      }
    }
    bool vis[2][2];
#pragma unroll
    for (int n = 0; n < 2; ++n) {
      const int ka = k0 + 8 * n + 2 * t4;
      vis[n][0] = ka < kend && mrow[min(ka, kend - 1)];
      vis[n][1] = ka + 1 < kend && mrow[min(ka + 1, kend - 1)];
    }
    float c[2][4];
#pragma unroll
    for (int n = 0; n < 2; ++n) c[n][0] = c[n][1] = c[n][2] = c[n][3] = 0.f;
#pragma unroll
    for (int s = 0; s < D / 16; ++s) {
#pragma unroll
      for (int n = 0; n < 2; ++n) {
        const uint4 w = kq[n][s >> 1];
        mma_bf16_16816(c[n], qa[s][0], 0u, qa[s][1], 0u, (s & 1) ? w.z : w.x, (s & 1) ? w.w : w.y);
      }
    }
    if (g < G) {
#pragma unroll
      for (int n = 0; n < 2; ++n) {
        const int ka = k0 + 8 * n + 2 * t4;
        const float s0 = vis[n][0] ? rbf(rbf(c[n][0]) * scale) : -INFINITY;
        const float s1 = vis[n][1] ? rbf(rbf(c[n][1]) * scale) : -INFINITY;
        *reinterpret_cast<float2*>(sc + g * spad + (ka - kbeg)) = make_float2(s0, s1);
      }
    }
  }
  __syncthreads();
  stamp(3);

  // ---- softmax over the visible keys (fp32), one warp per head; p = bf16(e / sum), keys past the end get 0.
  // Cluster variant: the row max and the row sum are combined over the CTAs (in rank order) before they are used.
  const int nloc = kend - kbeg;
  const int kpad = (nloc + 15) & ~15;
  constexpr int kHS = (kMaxG + kWarps - 1) / kWarps;
  float mxs[kHS], sums[kHS];
#pragma unroll
  for (int j = 0; j < kHS; ++j) {
    const int h = warp + j * kWarps;
    float mx = -INFINITY;
    if (h < G)
      for (int k = lane; k < nloc; k += 32) mx = fmaxf(mx, sc[h * spad + k]);
    mxs[j] = warp_max(mx);
  }
  if constexpr (CL) {
    cooperative_groups::cluster_group cluster = cooperative_groups::this_cluster();
#pragma unroll
    for (int j = 0; j < kHS; ++j)
      if (warp + j * kWarps < G && lane == 0) cl_max[warp + j * kWarps] = mxs[j];
    cluster.sync();
#pragma unroll
    for (int j = 0; j < kHS; ++j) {
      const int h = warp + j * kWarps;
      if (h < G) {
        float m = -INFINITY;
        for (int r = 0; r < cs; ++r) m = fmaxf(m, cluster.map_shared_rank(cl_max, r)[h]);
        mxs[j] = m;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < kHS; ++j) {
    const int h = warp + j * kWarps;
    float sum = 0.f;
    if (h < G) {
      float* row = sc + h * spad;
      for (int k = lane; k < nloc; k += 32) {
        const float sv = row[k];
        const float e = (sv == -INFINITY) ? 0.f : __expf(sv - mxs[j]);
        row[k] = e;
        sum += e;
      }
    }
    sums[j] = warp_sum(sum);
  }
  if constexpr (CL) {
    cooperative_groups::cluster_group cluster = cooperative_groups::this_cluster();
#pragma unroll
    for (int j = 0; j < kHS; ++j)
      if (warp + j * kWarps < G && lane == 0) cl_sum[warp + j * kWarps] = sums[j];
    cluster.sync();
#pragma unroll
    for (int j = 0; j < kHS; ++j) {
      const int h = warp + j * kWarps;
      if (h < G) {
        float t = 0.f;
        for (int r = 0; r < cs; ++r) t += cluster.map_shared_rank(cl_sum, r)[h];
        sums[j] = t;
      }
    }
  }
#pragma unroll
  for (int j = 0; j < kHS; ++j) {
    const int h = warp + j * kWarps;
    if (h < G) {
      const float* row = sc + h * spad;
      bf16* prow = sp + h * spad;
      const float sum = sums[j];
      for (int k = lane; k < kpad; k += 32) prow[k] = f2bf((k < nloc && sum > 0.f) ? row[k] / sum : 0.f);
    }
  }
  __syncthreads();
  stamp(4);

  // ---- pass 2: O = P.V, 16 keys per warp and iteration.  Lane (g, t4) reads 16 contiguous bytes (dims
  // 64 pp + 8 g .. +7) of its four key rows; word w of that chunk feeds dim-pair tile p = 4 pp + w, whose "even"
  // n-tile holds dim 64 pp + 8 n + 2 w in fragment column n and whose "odd" n-tile holds that dim + 1.
  float oe[D / 16][4], oo[D / 16][4];
#pragma unroll
  for (int p = 0; p < D / 16; ++p) {
    oe[p][0] = oe[p][1] = oe[p][2] = oe[p][3] = 0.f;
    oo[p][0] = oo[p][1] = oo[p][2] = oo[p][3] = 0.f;
  }
  for (int k0 = 16 * warp; k0 < kpad; k0 += 16 * kWarps) {          // k0 is relative to kbeg
    const uint32_t pa0 = g < G ? *reinterpret_cast<const uint32_t*>(sp + g * spad + k0 + 2 * t4) : 0u;
    const uint32_t pa2 = g < G ? *reinterpret_cast<const uint32_t*>(sp + g * spad + k0 + 2 * t4 + 8) : 0u;
    const int kk = kbeg + k0 + 2 * t4;
    const int r0 = min(kk, kend - 1), r1 = min(kk + 1, kend - 1);
    const int r2 = min(kk + 8, kend - 1), r3 = min(kk + 9, kend - 1);
    const uint4* v0 = reinterpret_cast<const uint4*>(vbase + static_cast<long long>(r0) * D) + g;
    const uint4* v1 = reinterpret_cast<const uint4*>(vbase + static_cast<long long>(r1) * D) + g;
    const uint4* v2 = reinterpret_cast<const uint4*>(vbase + static_cast<long long>(r2) * D) + g;
    const uint4* v3 = reinterpret_cast<const uint4*>(vbase + static_cast<long long>(r3) * D) + g;
    uint32_t w[D / 16][4];
#pragma unroll
    for (int pp = 0; pp < D / 64; ++pp) {       // 16 contiguous bytes (dims 64 pp + 8 g .. +7) of each of the four keys
      const uint4 a0 = v0[8 * pp], a1 = v1[8 * pp], a2 = v2[8 * pp], a3 = v3[8 * pp];
      w[4 * pp + 0][0] = a0.x; w[4 * pp + 0][1] = a1.x; w[4 * pp + 0][2] = a2.x; w[4 * pp + 0][3] = a3.x;
      w[4 * pp + 1][0] = a0.y; w[4 * pp + 1][1] = a1.y; w[4 * pp + 1][2] = a2.y; w[4 * pp + 1][3] = a3.y;
      w[4 * pp + 2][0] = a0.z; w[4 * pp + 2][1] = a1.z; w[4 * pp + 2][2] = a2.z; w[4 * pp + 2][3] = a3.z;
      w[4 * pp + 3][0] = a0.w; w[4 * pp + 3][1] = a1.w; w[4 * pp + 3][2] = a2.w; w[4 * pp + 3][3] = a3.w;
    }
#pragma unroll
    for (int p = 0; p < D / 16; ++p) {
      mma_bf16_16816(oe[p], pa0, 0u, pa2, 0u, __byte_perm(w[p][0], w[p][1], 0x5410), __byte_perm(w[p][2], w[p][3], 0x5410));
      mma_bf16_16816(oo[p], pa0, 0u, pa2, 0u, __byte_perm(w[p][0], w[p][1], 0x7632), __byte_perm(w[p][2], w[p][3], 0x7632));
    }
  }
  // tile p = 4 pp + w, fragment column n (= the g of the lane that loaded it) holds dims 64 pp + 8 n + 2 w (even
  // tile) and + 1 (odd tile); this lane's accumulators are row g, columns 2 t4 and 2 t4 + 1
  if (g < G) {
#pragma unroll
    for (int p = 0; p < D / 16; ++p) {
      const int d0 = 64 * (p >> 2) + 16 * t4 + 2 * (p & 3);
      *reinterpret_cast<float2*>(red + (warp * kMaxG + g) * D + d0) = make_float2(oe[p][0], oo[p][0]);
      *reinterpret_cast<float2*>(red + (warp * kMaxG + g) * D + d0 + 8) = make_float2(oe[p][1], oo[p][1]);
    }
  }
  __syncthreads();
  stamp(5);
  if constexpr (CL) {
    // CTA partial = warp partials in warp order; output = CTA partials in rank order (rank 0 writes)
    cooperative_groups::cluster_group cluster = cooperative_groups::this_cluster();
    for (int i = threadIdx.x; i < G * D; i += blockDim.x) {
      const int h = i / D, d = i % D;
      float v = 0.f;
#pragma unroll
      for (int w = 0; w < kWarps; ++w) v += red[(w * kMaxG + h) * D + d];
      cl_o[i] = v;
    }
    cluster.sync();
    if (rank == 0) {
      for (int i = threadIdx.x; i < G * D; i += blockDim.x) {
        float v = 0.f;
        for (int r = 0; r < cs; ++r) v += cluster.map_shared_rank(cl_o, r)[i];
        out[(static_cast<long long>(b) * n_heads + kvh * G) * D + i] = f2bf(v);
      }
    }
    cluster.sync();            // peers keep their shared memory alive until rank 0 has read it
  } else {
    for (int i = threadIdx.x; i < G * D; i += blockDim.x) {
      const int h = i / D, d = i % D;
      float v = 0.f;
#pragma unroll
      for (int w = 0; w < kWarps; ++w) v += red[(w * kMaxG + h) * D + d];
      out[(static_cast<long long>(b) * n_heads + kvh * G + h) * D + d] = f2bf(v);
    }
  }
  stamp(6);
}

}  // namespace

// CTAs per (batch row, kv head): 1 while the grid fills the GPU by itself, else a power of two <= 8 (the portable
// cluster size) that leaves every CTA at least 128 keys of the longest context.  MT_ATTN_CLUSTER or
// mt_debug_set_attn_cluster override.
static int g_attn_cluster = -1;       // -1: read MT_ATTN_CLUSTER on first use; 0: automatic; 1/2/4/8: forced
void set_attn_cluster(int cs) { g_attn_cluster = cs; }

int attn_decode_cluster(int B, int n_kv, int max_keys) {
  if (g_attn_cluster < 0) g_attn_cluster = exp_env("MT_ATTN_CLUSTER", 0);
  const int forced = g_attn_cluster;
  if (forced == 1 || forced == 2 || forced == 4 || forced == 8) return forced;
  int cs = 1;
  while (cs < 8 && B * n_kv * cs < 148 && max_keys / (2 * cs) >= 128) cs *= 2;
  return cs;
}

static size_t attn_smem_bytes(int G, int D, int max_keys, int cs, int* spad_out) {
  const int per = cs > 1 ? ((((max_keys + cs - 1) / cs) + 15) & ~15) : max_keys;
  const int spad = (per + 31) & ~31;
  if (spad_out) *spad_out = spad;
  const size_t fixed = D == 128 ? attn_fixed_bytes<128>(cs > 1) : attn_fixed_bytes<64>(cs > 1);
  return fixed + static_cast<size_t>(G) * spad * 6;
}

bool attn_early_reads_safe(int ctas_qkv, int ctas_between) { return ctas_qkv + ctas_between > kNumSMs + 8; }

bool attn_decode_fits(int B, int n_heads, int n_kv, int D, int max_keys) {
  if ((D != 64 && D != 128) || n_heads % n_kv || n_heads / n_kv > kMaxG) return false;
  return attn_smem_bytes(n_heads / n_kv, D, max_keys, attn_decode_cluster(B, n_kv, max_keys), nullptr) <= 200 * 1024;
}

int attn_decode(const bf16* qkv, const float* partials, int nsplit, long long split_stride, const bf16* q_gain,
                const bf16* k_gain, const bf16* rope_cos, const bf16* rope_sin, bf16* k_cache, bf16* v_cache,
                const uint8_t* mask, long long mask_ld, bf16* out, int B, int n_heads, int n_kv, int D,
                const int* pos_dev, int pos_host, int max_keys, int max_seq, float eps, cudaStream_t s, int early_pos) {
  const int G = n_heads / n_kv;
  const int cs = attn_decode_cluster(B, n_kv, max_keys);
  int spad = 0;
  const size_t smem = attn_smem_bytes(G, D, max_keys, cs, &spad);
  MT_CHECK(smem <= 200 * 1024, "attn_decode: %d keys do not fit in shared memory", max_keys);
  auto kern = cs > 1 ? (D == 128 ? attn_decode_kernel<128, true> : attn_decode_kernel<64, true>)
                     : (D == 128 ? attn_decode_kernel<128, false> : attn_decode_kernel<64, false>);
  static const int exp_flags = exp_env("MT_ATTN_EXP", 1);   // bit 0: no bulk L2 prefetch of the K/V rows (measured 1.3 % faster over 150 steps, also when issued before the dependency wait), bit 1: no early K tile, bit 2: no reads before the dependency wait
  static size_t smem_set[4] = {0, 0, 0, 0};
  size_t& cur = smem_set[(D == 128) + 2 * (cs > 1)];
  if (smem > 48 * 1024 && smem > cur) {
    MT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    cur = 200 * 1024;
  }
  const float scale = 1.0f / sqrtf(static_cast<float>(D));
  unsigned long long* trace = trace_slot(B * n_kv * cs * 8);
  if (cs > 1) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(B * n_kv * cs); cfg.blockDim = dim3(kWarps * 32); cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    MT_CUDA(cudaLaunchKernelEx(&cfg, kern, qkv, partials, nsplit, split_stride, q_gain, k_gain, rope_cos, rope_sin, k_cache,
                               v_cache, mask, mask_ld, out, n_heads, n_kv, pos_dev, pos_host, spad, max_seq, eps, scale,
                               trace, exp_flags, cs, early_pos));
  } else {
    MT_CUDA(launch(kern, dim3(B * n_kv), dim3(kWarps * 32), smem, s, qkv, partials, nsplit, split_stride, q_gain, k_gain,
                   rope_cos, rope_sin, k_cache, v_cache, mask, mask_ld, out, n_heads, n_kv, pos_dev, pos_host, spad,
                   max_seq, eps, scale, trace, exp_flags, 1, early_pos));
  }
  count_launch();
  return MT_OK;
}

}  // namespace mt
