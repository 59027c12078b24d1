// Codec decoder attention on the tensor cores (reference MossAudioTokenizerMultiheadAttention.forward,
// moss_audio_tokenizer/modeling_moss_audio_tokenizer.py:627-666): causal sliding-window attention,
// head_dim 64, fp32 storage, TF32 mma (m16n8k8) with fp32 accumulate and an online softmax.
//
// CTA = 4 warps = 64 consecutive queries of one (batch, head); each warp owns 16 query rows.  Key/value
// tiles of 64 keys are staged in shared memory (row pitch 68 floats: the fragment reads of both Q.K^T and
// P.V are bank-conflict free) already rounded to TF32.  The probabilities never leave registers: the
// m16n8k8 accumulator layout of S (thread holds keys 2t, 2t+1 of rows g, g+8) is reused directly as the A
// fragment of P.V by enumerating the keys of a k-step in the order {2t | 2t+1}, and V's B fragment is
// read in the same order.
//
// visible(q, k) = lo(q) <= k <= q,  lo(q) = max(q-context+1, chunk bound, 0); the chunk bound reproduces
// the ring KV cache of decode(chunk_duration) (:413-480, :1728-1750): start(q)+len(q)-context.
#include "codec_ops.cuh"

namespace mt {

namespace {

constexpr int kQ = 64, kK = 64, kD = 64, kPitch = 68, kWarps = 4;

__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void mma_tf32_16x8x8(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ int vis_lo(int t, int T, int context, int chunk) {
  int lo = t - context + 1;
  if (chunk > 0 && T > chunk) {
    const int start = (t / chunk) * chunk;
    const int len = min(chunk, T - start);
    lo = max(lo, start + len - context);
  }
  return max(lo, 0);
}

__global__ void __launch_bounds__(kWarps * 32)
codec_attention_tc_kernel(const float* __restrict__ qkv, float* __restrict__ out, int T, int heads, int context,
                          int chunk, float scale, const float4* __restrict__ rope, int out_f16) {
  __shared__ __align__(16) uint32_t sk[kK * kPitch];
  __shared__ __align__(16) uint32_t sv[kK * kPitch];
  pdl_trigger();
  pdl_wait();
  const int q0 = blockIdx.x * kQ, h = blockIdx.y, b = blockIdx.z;
  const int d = heads * kD;
  const long long tok0 = static_cast<long long>(b) * T;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;

  // ---- Q tile -> smem (coalesced) -> A fragments (scaled, TF32)
  for (int i = threadIdx.x; i < kQ * kD / 4; i += kWarps * 32) {
    const int r = i / (kD / 4), c = i % (kD / 4);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (q0 + r < T) {
      v = reinterpret_cast<const float4*>(qkv + (tok0 + q0 + r) * 3 * d + h * kD)[c];
      if (rope) {       // interleaved-pair RoPE (:255-300) applied on the way in: (cos, sin) of pairs 2c and 2c + 1
        const float4 cs = rope[static_cast<long long>(q0 + r) * (kD / 4) + c];
        v = make_float4(v.x * cs.x - v.y * cs.y, v.x * cs.y + v.y * cs.x, v.z * cs.z - v.w * cs.w, v.z * cs.w + v.w * cs.z);
      }
    }
    uint32_t* dst = sk + r * kPitch + 4 * c;
    dst[0] = to_tf32(v.x * scale); dst[1] = to_tf32(v.y * scale); dst[2] = to_tf32(v.z * scale); dst[3] = to_tf32(v.w * scale);
  }
  __syncthreads();
  uint32_t qa[kD / 8][4];
  {
    const uint32_t* base = sk + (warp * 16) * kPitch;
#pragma unroll
    for (int s = 0; s < kD / 8; ++s) {
      qa[s][0] = base[g * kPitch + 8 * s + t4];
      qa[s][1] = base[(g + 8) * kPitch + 8 * s + t4];
      qa[s][2] = base[g * kPitch + 8 * s + t4 + 4];
      qa[s][3] = base[(g + 8) * kPitch + 8 * s + t4 + 4];
    }
  }
  const int row0 = q0 + warp * 16 + g, row1 = row0 + 8;         // the two query rows this thread holds
  const int lo0 = vis_lo(min(row0, T - 1), T, context, chunk), lo1 = vis_lo(min(row1, T - 1), T, context, chunk);
  float o[kD / 8][4];
#pragma unroll
  for (int n = 0; n < kD / 8; ++n) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f; }
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

  const int last_q = min(q0 + kQ, T) - 1;
  const int k_begin = (vis_lo(q0, T, context, chunk) / kK) * kK;
  const int lo_last = vis_lo(last_q, T, context, chunk);
  for (int k0 = k_begin; k0 <= last_q; k0 += kK) {
    __syncthreads();                                            // previous tile (or the Q staging) is consumed
    for (int i = threadIdx.x; i < kK * kD / 4; i += kWarps * 32) {
      const int r = i / (kD / 4), c = i % (kD / 4);
      float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
      if (k0 + r < T) {
        const float* base = qkv + (tok0 + k0 + r) * 3 * d + h * kD;
        kv = reinterpret_cast<const float4*>(base + d)[c];
        vv = reinterpret_cast<const float4*>(base + 2 * d)[c];
        if (rope) {
          const float4 cs = rope[static_cast<long long>(k0 + r) * (kD / 4) + c];
          kv = make_float4(kv.x * cs.x - kv.y * cs.y, kv.x * cs.y + kv.y * cs.x, kv.z * cs.z - kv.w * cs.w, kv.z * cs.w + kv.w * cs.z);
        }
      }
      *reinterpret_cast<uint4*>(sk + r * kPitch + 4 * c) = make_uint4(to_tf32(kv.x), to_tf32(kv.y), to_tf32(kv.z), to_tf32(kv.w));
      *reinterpret_cast<uint4*>(sv + r * kPitch + 4 * c) = make_uint4(to_tf32(vv.x), to_tf32(vv.y), to_tf32(vv.z), to_tf32(vv.w));
    }
    __syncthreads();

    // ---- S = Q.K^T for 16 rows x 64 keys
    float sc[kK / 8][4];
#pragma unroll
    for (int n = 0; n < kK / 8; ++n) {
      sc[n][0] = sc[n][1] = sc[n][2] = sc[n][3] = 0.f;
      const uint32_t* kr = sk + (8 * n + g) * kPitch + t4;
#pragma unroll
      for (int s = 0; s < kD / 8; ++s) mma_tf32_16x8x8(sc[n], qa[s], kr[8 * s], kr[8 * s + 4]);
    }
    // ---- mask (edge tiles only), row max
    const bool interior = k0 + kK - 1 <= q0 && k0 >= lo_last;
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int n = 0; n < kK / 8; ++n) {
      if (!interior) {
        const int ka = k0 + 8 * n + 2 * t4, kb = ka + 1;
        if (ka < lo0 || ka > row0) sc[n][0] = -INFINITY;
        if (kb < lo0 || kb > row0) sc[n][1] = -INFINITY;
        if (ka < lo1 || ka > row1) sc[n][2] = -INFINITY;
        if (kb < lo1 || kb > row1) sc[n][3] = -INFINITY;
      }
      mx0 = fmaxf(mx0, fmaxf(sc[n][0], sc[n][1]));
      mx1 = fmaxf(mx1, fmaxf(sc[n][2], sc[n][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float nm0 = fmaxf(m0, mx0), nm1 = fmaxf(m1, mx1);
    // a row whose keys so far are all masked keeps m = -inf; exp(-inf - -inf) must not poison it
    const float c0 = nm0 == -INFINITY ? 1.f : __expf(m0 - nm0), c1 = nm1 == -INFINITY ? 1.f : __expf(m1 - nm1);
    const float b0 = nm0 == -INFINITY ? 0.f : nm0, b1 = nm1 == -INFINITY ? 0.f : nm1;
    m0 = nm0; m1 = nm1;
    l0 *= c0; l1 *= c1;
#pragma unroll
    for (int n = 0; n < kD / 8; ++n) { o[n][0] *= c0; o[n][1] *= c0; o[n][2] *= c1; o[n][3] *= c1; }
    // ---- P = exp(S - m), row sums; P (TF32) is the A operand of P.V with keys taken in the order {2t | 2t+1}
    float s0 = 0.f, s1 = 0.f;
    uint32_t pa[kK / 8][4];
#pragma unroll
    for (int n = 0; n < kK / 8; ++n) {
      const float p0 = __expf(sc[n][0] - b0), p1 = __expf(sc[n][1] - b0);
      const float p2 = __expf(sc[n][2] - b1), p3 = __expf(sc[n][3] - b1);
      s0 += p0 + p1; s1 += p2 + p3;
      pa[n][0] = to_tf32(p0); pa[n][1] = to_tf32(p2); pa[n][2] = to_tf32(p1); pa[n][3] = to_tf32(p3);
    }
    l0 += s0; l1 += s1;
    // ---- O += P.V
#pragma unroll
    for (int n = 0; n < kD / 8; ++n) {
      const uint32_t* vr = sv + (2 * t4) * kPitch + 8 * n + g;
#pragma unroll
      for (int s = 0; s < kK / 8; ++s) mma_tf32_16x8x8(o[n], pa[s], vr[(8 * s) * kPitch], vr[(8 * s + 1) * kPitch]);
    }
  }
  // ---- normalise and store
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = l0 > 0.f ? 1.f / l0 : 0.f, i1 = l1 > 0.f ? 1.f / l1 : 0.f;
#pragma unroll
  for (int n = 0; n < kD / 8; ++n) {
    const int col = h * kD + 8 * n + 2 * t4;
    if (out_f16) {
      __half* oh = reinterpret_cast<__half*>(out);
      if (row0 < T) *reinterpret_cast<uint32_t*>(oh + (tok0 + row0) * d + col) = pack_f16(o[n][0] * i0, o[n][1] * i0);
      if (row1 < T) *reinterpret_cast<uint32_t*>(oh + (tok0 + row1) * d + col) = pack_f16(o[n][2] * i1, o[n][3] * i1);
    } else {
      if (row0 < T) *reinterpret_cast<float2*>(out + (tok0 + row0) * d + col) = make_float2(o[n][0] * i0, o[n][1] * i0);
      if (row1 < T) *reinterpret_cast<float2*>(out + (tok0 + row1) * d + col) = make_float2(o[n][2] * i1, o[n][3] * i1);
    }
  }
}

}  // namespace

int codec_attention_tc(const float* qkv, float* out, int B, int T, int heads, int hd, int context, int chunk,
                       cudaStream_t s, const float* rope_table, int out_f16) {
  if (hd != kD) return fail(MT_ERR_UNSUPPORTED, "codec attention supports head_dim 64 only (got %d)", hd);
  dim3 grid(cdiv(T, kQ), heads, B);
  MT_CUDA(launch(codec_attention_tc_kernel, grid, dim3(kWarps * 32), 0, s, qkv, out, T, heads, context, chunk,
                 1.0f / sqrtf(static_cast<float>(hd)), reinterpret_cast<const float4*>(rope_table), out_f16));
  count_launch();
  return MT_OK;
}

}  // namespace mt

// =============================================================================================
// fp16 flavour (layer_gemm_f16 codecs): q/k/v arrive as IEEE half from the in_proj GEMM's epilogue, the output
// leaves as half for the out_proj GEMM.  FlashAttention-2 structure on mma.sync.m16n8k16 (fp32 accumulate):
// CTA = 8 warps = 128 consecutive queries of one (batch, head), 16 query rows per warp; 64-key K / V tiles are
// staged once per CTA in shared memory (row pitch 72 halves = 144 B: every ldmatrix phase touches 8 rows that
// fall into 8 different 16-byte bank groups), K with the interleaved-pair RoPE applied in fp32 on the way in.
// B fragments come from ldmatrix.x4 (K: two k-steps per instruction; V: .trans, two 8-dim column groups per
// instruction) -- 32 ldmatrix per warp and tile instead of the 256 LDS.32 of the TF32 kernel -- and half as
// many MMAs, each doing twice the work.  Same mantissa as the TF32 path (10 explicit bits), same online softmax.
// =============================================================================================
namespace mt {
namespace {

constexpr int kHQ = 128, kHWarps = 8, kHPitch = 72;

__device__ __forceinline__ void mma_f16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], const __half* p) {
  const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(p));
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], const __half* p) {
  const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(p));
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(a));
}
// 8 halves (dims 8c .. 8c+7 of one head row) rotated by the interleaved-pair RoPE of position `pos`, times `mul`
__device__ __forceinline__ uint4 rope8(uint4 raw, const float4* __restrict__ rope, long long pos, int c, float mul) {
  const float4 cs0 = rope[pos * (kD / 4) + 2 * c], cs1 = rope[pos * (kD / 4) + 2 * c + 1];
  const float2 p0 = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
  const float2 p1 = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
  const float2 p2 = __half22float2(*reinterpret_cast<const __half2*>(&raw.z));
  const float2 p3 = __half22float2(*reinterpret_cast<const __half2*>(&raw.w));
  uint4 o;
  o.x = pack_f16((p0.x * cs0.x - p0.y * cs0.y) * mul, (p0.x * cs0.y + p0.y * cs0.x) * mul);
  o.y = pack_f16((p1.x * cs0.z - p1.y * cs0.w) * mul, (p1.x * cs0.w + p1.y * cs0.z) * mul);
  o.z = pack_f16((p2.x * cs1.x - p2.y * cs1.y) * mul, (p2.x * cs1.y + p2.y * cs1.x) * mul);
  o.w = pack_f16((p3.x * cs1.z - p3.y * cs1.w) * mul, (p3.x * cs1.w + p3.y * cs1.z) * mul);
  return o;
}

// RING = true: streaming push.  Queries are the T new tokens (absolute positions pos0 + row); keys live in the layer's ring
// cache (already rotated) at slot = position % cap; `context` carries cap and the visible range of every row starts at
// max(pos0 + T - cap, 0).  Key positions, not rows, drive the tile loop.
template <bool RING>
__global__ void __launch_bounds__(kHWarps * 32, 2)
codec_attention_f16_kernel(const __half* __restrict__ qkv, __half* __restrict__ out, int T, int heads, int context,
                           int chunk, float scale, const float4* __restrict__ rope, const __half* __restrict__ ring,
                           int pos0) {
  __shared__ __align__(16) __half sm[2 * kK * kHPitch];        // K tile | V tile (both: the 128-row Q tile first)
  __half* sk = sm;
  __half* sv = sm + kK * kHPitch;
  pdl_trigger();
  pdl_wait();
  const int q0 = blockIdx.x * kHQ, h = blockIdx.y, b = blockIdx.z;
  const int d = heads * kD;
  const long long tok0 = static_cast<long long>(b) * T;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;

  // ---- Q tile (RoPE, 1/sqrt(d) folded in) -> smem -> A fragments
  for (int i = threadIdx.x; i < kHQ * (kD / 8); i += kHWarps * 32) {
    const int r = i / (kD / 8), c = i % (kD / 8);
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (q0 + r < T) {
      v = *reinterpret_cast<const uint4*>(qkv + (tok0 + q0 + r) * 3 * d + h * kD + 8 * c);
      v = rope8(v, rope, q0 + r, c, scale);
    }
    *reinterpret_cast<uint4*>(sm + r * kHPitch + 8 * c) = v;
  }
  __syncthreads();
  uint32_t qa[kD / 16][4];
  {
    // A fragment of k-step s: a0 (row g, cols 16s+2t4), a1 (row g+8), a2 (row g, cols +8), a3 (row g+8, cols +8)
    const __half* base = sm + (warp * 16 + (lane & 15)) * kHPitch + 8 * (lane >> 4);
#pragma unroll
    for (int s = 0; s < kD / 16; ++s) ldsm_x4(qa[s], base + 16 * s);
  }
  const int row0 = q0 + warp * 16 + g, row1 = row0 + 8;
  const int ring_lo = max(pos0 + T - context, 0);
  // from here on rows and keys are compared as POSITIONS (pos0 = 0 outside streaming)
  const int p0 = pos0 + row0, p1 = pos0 + row1;
  const int lo0 = RING ? ring_lo : vis_lo(min(row0, T - 1), T, context, chunk);
  const int lo1 = RING ? ring_lo : vis_lo(min(row1, T - 1), T, context, chunk);
  const int wfirst = q0 + warp * 16, wlast = pos0 + min(wfirst + 15, T - 1);   // this warp's query rows / last position
  const int wlo = RING ? ring_lo : vis_lo(min(wfirst, T - 1), T, context, chunk);   // lowest key any of them sees
  float o[kD / 8][4];
#pragma unroll
  for (int n = 0; n < kD / 8; ++n) { o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f; }
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

  const int last_q = pos0 + min(q0 + kHQ, T) - 1;                 // position of the CTA's last query
  const int k_begin = ((RING ? ring_lo : vis_lo(q0, T, context, chunk)) / kK) * kK;
  const int lo_last = RING ? ring_lo : vis_lo(last_q, T, context, chunk);
  for (int k0 = k_begin; k0 <= last_q; k0 += kK) {
    __syncthreads();                                            // previous tile (or the Q staging) is consumed
    for (int i = threadIdx.x; i < kK * (kD / 8); i += kHWarps * 32) {
      const int r = i / (kD / 8), c = i % (kD / 8);
      uint4 kv = make_uint4(0u, 0u, 0u, 0u), vv = kv;
      if constexpr (RING) {
        const int kp = k0 + r;
        if (kp >= ring_lo && kp < pos0 + T) {
          const __half* base = ring + ((static_cast<long long>(b) * context + kp % context) * 2) * d + h * kD + 8 * c;
          kv = *reinterpret_cast<const uint4*>(base);
          vv = *reinterpret_cast<const uint4*>(base + d);
        }
      } else if (k0 + r < T) {
        const __half* base = qkv + (tok0 + k0 + r) * 3 * d + h * kD + 8 * c;
        kv = rope8(*reinterpret_cast<const uint4*>(base + d), rope, k0 + r, c, 1.f);
        vv = *reinterpret_cast<const uint4*>(base + 2 * d);
      }
      *reinterpret_cast<uint4*>(sk + r * kHPitch + 8 * c) = kv;
      *reinterpret_cast<uint4*>(sv + r * kHPitch + 8 * c) = vv;
    }
    __syncthreads();
    if (wfirst >= T || k0 > wlast || k0 + kK - 1 < wlo) continue;      // nothing visible to this warp in this tile

    // ---- S = Q.K^T for 16 rows x 64 keys
    float sc[kK / 8][4];
#pragma unroll
    for (int n = 0; n < kK / 8; ++n) {
      sc[n][0] = sc[n][1] = sc[n][2] = sc[n][3] = 0.f;
      // matrices: keys 8n..8n+7 x dims {0-7, 8-15, 16-23, 24-31} (+32 for the second instruction)
      const __half* kr = sk + (8 * n + (lane & 7)) * kHPitch + 8 * (lane >> 3);
      uint32_t kb[4];
      ldsm_x4(kb, kr);
      mma_f16_16816(sc[n], qa[0], kb[0], kb[1]);
      mma_f16_16816(sc[n], qa[1], kb[2], kb[3]);
      ldsm_x4(kb, kr + 32);
      mma_f16_16816(sc[n], qa[2], kb[0], kb[1]);
      mma_f16_16816(sc[n], qa[3], kb[2], kb[3]);
    }
    // ---- mask (edge tiles only), row max
    const bool interior = k0 + kK - 1 <= pos0 + q0 && k0 >= lo_last;
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int n = 0; n < kK / 8; ++n) {
      if (!interior) {
        const int ka = k0 + 8 * n + 2 * t4, kb2 = ka + 1;
        if (ka < lo0 || ka > p0) sc[n][0] = -INFINITY;
        if (kb2 < lo0 || kb2 > p0) sc[n][1] = -INFINITY;
        if (ka < lo1 || ka > p1) sc[n][2] = -INFINITY;
        if (kb2 < lo1 || kb2 > p1) sc[n][3] = -INFINITY;
      }
      mx0 = fmaxf(mx0, fmaxf(sc[n][0], sc[n][1]));
      mx1 = fmaxf(mx1, fmaxf(sc[n][2], sc[n][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    const float nm0 = fmaxf(m0, mx0), nm1 = fmaxf(m1, mx1);
    const float c0 = nm0 == -INFINITY ? 1.f : __expf(m0 - nm0), c1 = nm1 == -INFINITY ? 1.f : __expf(m1 - nm1);
    const float b0 = nm0 == -INFINITY ? 0.f : nm0, b1 = nm1 == -INFINITY ? 0.f : nm1;
    m0 = nm0; m1 = nm1;
    l0 *= c0; l1 *= c1;
#pragma unroll
    for (int n = 0; n < kD / 8; ++n) { o[n][0] *= c0; o[n][1] *= c0; o[n][2] *= c1; o[n][3] *= c1; }
    // ---- P = exp(S - m) as half A fragments: key groups 2j, 2j+1 of S are k-step j of P.V
    float s0 = 0.f, s1 = 0.f;
    uint32_t pa[kK / 16][4];
#pragma unroll
    for (int n = 0; n < kK / 8; ++n) {
      const float p0 = __expf(sc[n][0] - b0), p1 = __expf(sc[n][1] - b0);
      const float p2 = __expf(sc[n][2] - b1), p3 = __expf(sc[n][3] - b1);
      s0 += p0 + p1; s1 += p2 + p3;
      pa[n >> 1][(n & 1) * 2 + 0] = pack_f16(p0, p1);
      pa[n >> 1][(n & 1) * 2 + 1] = pack_f16(p2, p3);
    }
    l0 += s0; l1 += s1;
    // ---- O += P.V: matrices (keys 16j..+7 | 16j+8..+15) x (dims 8n..+7 | 8n+8..+15), transposed on load
#pragma unroll
    for (int j = 0; j < kK / 16; ++j) {
      const __half* vr = sv + (16 * j + (lane & 15)) * kHPitch + 8 * (lane >> 4);
#pragma unroll
      for (int n = 0; n < kD / 8; n += 2) {
        uint32_t vb[4];
        ldsm_x4_trans(vb, vr + 8 * n);
        mma_f16_16816(o[n], pa[j], vb[0], vb[1]);
        mma_f16_16816(o[n + 1], pa[j], vb[2], vb[3]);
      }
    }
  }
  // ---- normalise and store
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = l0 > 0.f ? 1.f / l0 : 0.f, i1 = l1 > 0.f ? 1.f / l1 : 0.f;
#pragma unroll
  for (int n = 0; n < kD / 8; ++n) {
    const int col = h * kD + 8 * n + 2 * t4;
    if (row0 < T) *reinterpret_cast<uint32_t*>(out + (tok0 + row0) * d + col) = pack_f16(o[n][0] * i0, o[n][1] * i0);
    if (row1 < T) *reinterpret_cast<uint32_t*>(out + (tok0 + row1) * d + col) = pack_f16(o[n][2] * i1, o[n][3] * i1);
  }
}

}  // namespace

int codec_attention_f16(const void* qkv, void* out, int B, int T, int heads, int hd, int context, int chunk,
                        cudaStream_t s, const float* rope_table) {
  if (hd != kD) return fail(MT_ERR_UNSUPPORTED, "codec attention supports head_dim 64 only (got %d)", hd);
  MT_CHECK(rope_table, "codec_attention_f16: the RoPE table is required");
  dim3 grid(cdiv(T, kHQ), heads, B);
  MT_CUDA(launch(codec_attention_f16_kernel<false>, grid, dim3(kHWarps * 32), 0, s, reinterpret_cast<const __half*>(qkv),
                 reinterpret_cast<__half*>(out), T, heads, context, chunk, 1.0f / sqrtf(static_cast<float>(hd)),
                 reinterpret_cast<const float4*>(rope_table), static_cast<const __half*>(nullptr), 0));
  count_launch();
  return MT_OK;
}

namespace {
// new tokens' k (rotated: table row t = position pos0 + t) and v -> ring slots (pos0 + t) % cap; 8 halves per thread
__global__ void codec_kv_append_kernel(const __half* __restrict__ qkv, __half* __restrict__ ring, int T, int d, int cap,
                                       int pos0, const float4* __restrict__ rope, long long total) {
  pdl_trigger();
  pdl_wait();
  const int d8 = d / 8;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long tok = idx / d8;                 // b * T + t
    const int c8 = static_cast<int>(idx % d8);      // 8-half column group inside the d-wide k (or v) vector
    const int b = static_cast<int>(tok / T), t = static_cast<int>(tok % T);
    const __half* src = qkv + tok * 3 * d + 8 * c8;
    const uint4 k = rope8(*reinterpret_cast<const uint4*>(src + d), rope, t, c8 % (kD / 8), 1.f);
    const uint4 v = *reinterpret_cast<const uint4*>(src + 2 * d);
    __half* dst = ring + ((static_cast<long long>(b) * cap + (pos0 + t) % cap) * 2) * d + 8 * c8;
    *reinterpret_cast<uint4*>(dst) = k;
    *reinterpret_cast<uint4*>(dst + d) = v;
  }
}
}  // namespace

int codec_kv_append_f16(const void* qkv, void* ring, int B, int T, int heads, int cap, long long pos0, const float* rope_table,
                        cudaStream_t s) {
  MT_CHECK(rope_table && T <= cap, "codec_kv_append_f16: a push of %d tokens exceeds the ring capacity %d", T, cap);
  MT_CHECK(pos0 + T < (1ll << 31), "codec stream position overflow");
  const int d = heads * kD;
  const long long total = static_cast<long long>(B) * T * (d / 8);
  int blocks = static_cast<int>(cdivll(total, 256));
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  MT_CUDA(launch(codec_kv_append_kernel, dim3(blocks), dim3(256), 0, s, reinterpret_cast<const __half*>(qkv),
                 reinterpret_cast<__half*>(ring), T, d, cap, static_cast<int>(pos0), reinterpret_cast<const float4*>(rope_table), total));
  count_launch();
  return MT_OK;
}

int codec_attention_ring_f16(const void* qkv, const void* ring, void* out, int B, int T, int heads, int cap, long long pos0,
                             cudaStream_t s, const float* rope_table) {
  MT_CHECK(rope_table && ring, "codec_attention_ring_f16: null argument");
  dim3 grid(cdiv(T, kHQ), heads, B);
  MT_CUDA(launch(codec_attention_f16_kernel<true>, grid, dim3(kHWarps * 32), 0, s, reinterpret_cast<const __half*>(qkv),
                 reinterpret_cast<__half*>(out), T, heads, cap, 0, 1.0f / sqrtf(static_cast<float>(kD)),
                 reinterpret_cast<const float4*>(rope_table), reinterpret_cast<const __half*>(ring), static_cast<int>(pos0)));
  count_launch();
  return MT_OK;
}

}  // namespace mt
