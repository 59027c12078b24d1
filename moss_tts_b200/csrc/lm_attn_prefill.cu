// Prefill attention of the Qwen3 backbone on tensor cores (mma.sync.m16n8k16 bf16), any number of keys.
// Reference semantics (transformers models/qwen3/modeling_qwen3.py eager_attention_forward :214-238 as called from
// moss_tts_delay/modeling_moss_tts.py:264-276): scores = bf16(bf16(q.k) * scale) + mask, softmax in fp32,
// probabilities rounded to bf16, out = bf16(P.V).  The rounding of P needs the finished row sum, so the kernel
// makes two passes over the keys: pass 1 = row max and sum (online), pass 2 recomputes the scores, forms
// p = bf16(exp(s - max) / sum) in registers and feeds it straight into the P.V MMAs (the C fragment layout of
// two adjacent 8-key tiles is the A fragment layout of one 16-key step).  Nothing of size T x S is ever stored.
//
// One block per (16-query tile, kv head, batch row); warp g of the block owns query head kvh * G + g, so the
// G warps share every staged K/V chunk (64 keys, rows padded to D + 8 elements: conflict-free LDS / ldmatrix).
#include "lm_ops.cuh"

#include <stdlib.h>

namespace mt {

namespace {

constexpr int kChunk = 64;

__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], const void* smem_ptr) {
  const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(smem_ptr));
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(a));
}

template <int D, int G>
__global__ void __launch_bounds__(G * 32)
attn_prefill_kernel(const bf16* __restrict__ q, const bf16* __restrict__ k_cache, const bf16* __restrict__ v_cache,
                    const uint8_t* __restrict__ mask, long long mask_ld, bf16* __restrict__ out, int T, int n_heads,
                    int n_kv, int past_len, const int* __restrict__ past_len_dev, int max_seq, float scale) {
  constexpr int LD = D + 8;                     // padded row (elements)
  __shared__ __align__(16) bf16 sK[kChunk * LD];
  __shared__ __align__(16) bf16 sV[kChunk * LD];
  __shared__ uint8_t sM[kChunk];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gr = lane >> 2, t4 = lane & 3;
  const int t0 = blockIdx.x * 16, kvh = blockIdx.y, b = blockIdx.z;
  pdl_trigger();
  pdl_wait();
  if (past_len_dev) past_len = *past_len_dev;
  const int head = kvh * G + warp;
  const int last_t = min(t0 + 15, T - 1);
  const int nkeys = past_len + last_t + 1;      // keys any query of this tile may see
  const bf16* kbase = k_cache + (static_cast<long long>(b) * n_kv + kvh) * max_seq * D;
  const bf16* vbase = v_cache + (static_cast<long long>(b) * n_kv + kvh) * max_seq * D;
  const uint8_t* mrow = mask + b * mask_ld;

  // ---- Q fragments of this warp's head: rows gr and gr + 8 of the tile (queries t0 + row)
  const int tq[2] = {t0 + gr, t0 + gr + 8};
  const int posq[2] = {past_len + tq[0], past_len + tq[1]};
  uint32_t qa[D / 16][4];
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int t = min(tq[r], T - 1);
    const uint32_t* qp = reinterpret_cast<const uint32_t*>(q + ((static_cast<long long>(b) * T + t) * n_heads + head) * D) + t4;
#pragma unroll
    for (int s = 0; s < D / 16; ++s) {
      qa[s][r] = qp[8 * s];           // k = 16 s + 2 t4, +1
      qa[s][r + 2] = qp[8 * s + 4];   // k = 16 s + 8 + 2 t4, +1
    }
  }

  auto stage = [&](int c0, bool with_v) {
    // 64 rows x D elements of K (and V) -> padded shared rows; row indices clamped to written positions
    constexpr int kVecPerRow = D / 8;
    for (int i = threadIdx.x; i < kChunk * kVecPerRow; i += G * 32) {
      const int r = i / kVecPerRow, v = i % kVecPerRow;
      const int key = min(c0 + r, nkeys - 1);
      *reinterpret_cast<uint4*>(sK + r * LD + 8 * v) = *reinterpret_cast<const uint4*>(kbase + static_cast<long long>(key) * D + 8 * v);
      if (with_v)
        *reinterpret_cast<uint4*>(sV + r * LD + 8 * v) = *reinterpret_cast<const uint4*>(vbase + static_cast<long long>(key) * D + 8 * v);
    }
    for (int i = threadIdx.x; i < kChunk; i += G * 32) sM[i] = (c0 + i < nkeys) ? mrow[c0 + i] : 0;
  };

  // scores of one chunk: sc[j][0..3] = rows (gr, gr, gr+8, gr+8) x keys (8j + 2 t4, +1, same, +1); -inf where masked
  auto scores = [&](int c0, float (&sc)[kChunk / 8][4]) {
#pragma unroll
    for (int j = 0; j < kChunk / 8; ++j) {
      float c[4] = {0.f, 0.f, 0.f, 0.f};
      const uint32_t* kp = reinterpret_cast<const uint32_t*>(sK + (8 * j + gr) * LD) + t4;
#pragma unroll
      for (int s = 0; s < D / 16; ++s) mma16816(c, qa[s][0], qa[s][1], qa[s][2], qa[s][3], kp[8 * s], kp[8 * s + 4]);
      const int key = c0 + 8 * j + 2 * t4;
      const bool m0 = sM[8 * j + 2 * t4] != 0, m1 = sM[8 * j + 2 * t4 + 1] != 0;
      sc[j][0] = (m0 && key <= posq[0]) ? rbf(rbf(c[0]) * scale) : -INFINITY;
      sc[j][1] = (m1 && key + 1 <= posq[0]) ? rbf(rbf(c[1]) * scale) : -INFINITY;
      sc[j][2] = (m0 && key <= posq[1]) ? rbf(rbf(c[2]) * scale) : -INFINITY;
      sc[j][3] = (m1 && key + 1 <= posq[1]) ? rbf(rbf(c[3]) * scale) : -INFINITY;
    }
  };

  // ---- pass 1: row max and sum of exp (online; the four lanes of a quad hold disjoint key columns of a row)
  float mrun[2] = {-INFINITY, -INFINITY}, lrun[2] = {0.f, 0.f};
  for (int c0 = 0; c0 < nkeys; c0 += kChunk) {
    __syncthreads();
    stage(c0, false);
    __syncthreads();
    float sc[kChunk / 8][4];
    scores(c0, sc);
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      float cm = -INFINITY;
#pragma unroll
      for (int j = 0; j < kChunk / 8; ++j) cm = fmaxf(cm, fmaxf(sc[j][2 * r], sc[j][2 * r + 1]));
      cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, 1));
      cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, 2));
      const float mn = fmaxf(mrun[r], cm);
      if (mn == -INFINITY) continue;
      float add = 0.f;
#pragma unroll
      for (int j = 0; j < kChunk / 8; ++j) {
        add += (sc[j][2 * r] == -INFINITY) ? 0.f : __expf(sc[j][2 * r] - mn);
        add += (sc[j][2 * r + 1] == -INFINITY) ? 0.f : __expf(sc[j][2 * r + 1] - mn);
      }
      lrun[r] = lrun[r] * ((mrun[r] == -INFINITY) ? 0.f : __expf(mrun[r] - mn)) + add;
      mrun[r] = mn;
    }
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    lrun[r] += __shfl_xor_sync(0xffffffffu, lrun[r], 1);
    lrun[r] += __shfl_xor_sync(0xffffffffu, lrun[r], 2);
  }

  // ---- pass 2: p = bf16(exp(s - max) / sum), O += P.V
  float o[D / 8][4];
#pragma unroll
  for (int jd = 0; jd < D / 8; ++jd) o[jd][0] = o[jd][1] = o[jd][2] = o[jd][3] = 0.f;
  for (int c0 = 0; c0 < nkeys; c0 += kChunk) {
    __syncthreads();
    stage(c0, true);
    __syncthreads();
    float sc[kChunk / 8][4];
    scores(c0, sc);
    uint32_t pa[kChunk / 8][2];          // [j][0] = row gr, [j][1] = row gr + 8: keys (8j + 2 t4, +1) as packed bf16
#pragma unroll
    for (int j = 0; j < kChunk / 8; ++j) {
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const float s0 = sc[j][2 * r], s1 = sc[j][2 * r + 1];
        const float p0 = (s0 == -INFINITY || !(lrun[r] > 0.f)) ? 0.f : __expf(s0 - mrun[r]) / lrun[r];
        const float p1 = (s1 == -INFINITY || !(lrun[r] > 0.f)) ? 0.f : __expf(s1 - mrun[r]) / lrun[r];
        pa[j][r] = pack_bf16(p0, p1);
      }
    }
#pragma unroll
    for (int kk = 0; kk < kChunk / 16; ++kk) {
      // A fragment of P for keys [16 kk, 16 kk + 16): tiles 2kk (low 8 keys) and 2kk + 1 (high 8 keys)
      const uint32_t a0 = pa[2 * kk][0], a1 = pa[2 * kk][1], a2 = pa[2 * kk + 1][0], a3 = pa[2 * kk + 1][1];
#pragma unroll
      for (int jd = 0; jd < D / 8; jd += 2) {
        // matrices: (keys lo, dims jd), (keys hi, dims jd), (keys lo, dims jd + 1), (keys hi, dims jd + 1)
        const int mi = lane >> 3, mr = lane & 7;
        const bf16* vp = sV + (16 * kk + (mi & 1) * 8 + mr) * LD + 8 * (jd + (mi >> 1));
        uint32_t vb[4];
        ldmatrix_x4_trans(vb, vp);
        mma16816(o[jd], a0, a1, a2, a3, vb[0], vb[1]);
        mma16816(o[jd + 1], a0, a1, a2, a3, vb[2], vb[3]);
      }
    }
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    if (tq[r] >= T) continue;
    uint32_t* op = reinterpret_cast<uint32_t*>(out + ((static_cast<long long>(b) * T + tq[r]) * n_heads + head) * D) + t4;
#pragma unroll
    for (int jd = 0; jd < D / 8; ++jd) op[4 * jd] = pack_bf16(o[jd][2 * r], o[jd][2 * r + 1]);
  }
}

template <int D, int G>
int launch_prefill(const bf16* q, const bf16* k_cache, const bf16* v_cache, const uint8_t* mask, long long mask_ld, bf16* out,
                   int B, int T, int n_heads, int n_kv, int past_len, const int* past_len_dev, int max_seq, cudaStream_t s) {
  dim3 grid((T + 15) / 16, n_kv, B);
  MT_CUDA(launch(attn_prefill_kernel<D, G>, grid, dim3(G * 32), 0, s, q, k_cache, v_cache, mask, mask_ld, out, T, n_heads,
                 n_kv, past_len, past_len_dev, max_seq, 1.0f / sqrtf(static_cast<float>(D))));
  count_launch();
  return MT_OK;
}

}  // namespace

bool attention_tc_supported(int n_heads, int n_kv, int D) {
  const int G = n_kv > 0 ? n_heads / n_kv : 0;
  static const bool off = exp_env("MT_OLD_PREFILL_ATTN", 0) != 0;   // A/B: the SIMT two-pass kernel
  return !off && (D == 64 || D == 128) && (G == 1 || G == 2 || G == 4 || G == 8);
}

int attention_tc(const bf16* q, const bf16* k_cache, const bf16* v_cache, const uint8_t* mask, long long mask_ld, bf16* out,
                 int B, int T, int n_heads, int n_kv, int D, int past_len, const int* past_len_dev, int max_seq,
                 cudaStream_t s) {
  const int G = n_heads / n_kv;
#define MT_PF(DD, GG)                                                                                              \
  if (D == DD && G == GG)                                                                                           \
    return launch_prefill<DD, GG>(q, k_cache, v_cache, mask, mask_ld, out, B, T, n_heads, n_kv, past_len, past_len_dev, \
                                  max_seq, s);
  MT_PF(128, 1) MT_PF(128, 2) MT_PF(128, 4) MT_PF(128, 8) MT_PF(64, 1) MT_PF(64, 2) MT_PF(64, 4) MT_PF(64, 8)
#undef MT_PF
  return fail(MT_ERR_UNSUPPORTED, "attention_tc: head_dim %d with group size %d not supported", D, G);
}

}  // namespace mt
