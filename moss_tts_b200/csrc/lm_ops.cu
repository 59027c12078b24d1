// Non-GEMM kernels of the MossTTSDelay decode/prefill step.  Every kernel rounds to bf16 exactly
// where the reference's PyTorch ops return bf16 tensors (see oracle/lm.py), so greedy ids can match.
//   a1  embed_sum            modeling_moss_tts.py:196-213
//   a2  add_rmsnorm          Qwen3RMSNorm modeling_qwen3.py:59-64 + residual adds :305-335
//       qkv_post             q/k RMSNorm + RoPE :263-268,:160-181 + KV append cache_utils.py:102-121
//       attention            eager_attention_forward :196-217 (GQA, additive mask, fp32 softmax)
//       silu_mul             Qwen3MLP :81-83
#include "lm_ops.cuh"

namespace mt {

// ---------------------------------------------------------------------------------------------
// a1: x[m,:] = E_text[id0] + sum_i emb_ext[i][id_{i+1}]   (left-to-right, bf16 after every add)
// ---------------------------------------------------------------------------------------------
__global__ void embed_sum_kernel(const int64_t* __restrict__ ids, const bf16* __restrict__ emb_text,
                                 const bf16* __restrict__ emb_ext, bf16* __restrict__ out, int H, int n_vq,
                                 int audio_rows, int vocab, int row_stride) {
  const int m = blockIdx.x;
  const int64_t* row = ids + static_cast<long long>(m) * row_stride;
  pdl_trigger();
  pdl_wait();
  for (int c = threadIdx.x * 8; c < H; c += blockDim.x * 8) {
    long long t0 = row[0];
    t0 = t0 < 0 ? 0 : (t0 >= vocab ? vocab - 1 : t0);
    uint4 a = *reinterpret_cast<const uint4*>(emb_text + t0 * H + c);
    float acc[8] = {bits_lo(a.x), bits_hi(a.x), bits_lo(a.y), bits_hi(a.y),
                    bits_lo(a.z), bits_hi(a.z), bits_lo(a.w), bits_hi(a.w)};
    // eight table rows are requested before the first add (the add order stays left to right)
    for (int i0 = 0; i0 < n_vq; i0 += 8) {
      uint4 bb[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (i0 + u < n_vq) {
          long long t = row[1 + i0 + u];
          t = t < 0 ? 0 : (t >= audio_rows ? audio_rows - 1 : t);
          bb[u] = *reinterpret_cast<const uint4*>(emb_ext + (static_cast<long long>(i0 + u) * audio_rows + t) * H + c);
        }
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (i0 + u < n_vq) {
          const uint4 b = bb[u];
          const float e[8] = {bits_lo(b.x), bits_hi(b.x), bits_lo(b.y), bits_hi(b.y),
                              bits_lo(b.z), bits_hi(b.z), bits_lo(b.w), bits_hi(b.w)};
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] = rbf(acc[j] + e[j]);
        }
      }
    }
    uint4 o;
    o.x = pack_bf16(acc[0], acc[1]); o.y = pack_bf16(acc[2], acc[3]);
    o.z = pack_bf16(acc[4], acc[5]); o.w = pack_bf16(acc[6], acc[7]);
    *reinterpret_cast<uint4*>(out + static_cast<long long>(m) * H + c) = o;
  }
}

int embed_sum(const int64_t* ids, const bf16* emb_text, const bf16* emb_ext, bf16* out, int M, int H, int n_vq,
              int audio_rows, int vocab, cudaStream_t s, int row_stride) {
  if (row_stride <= 0) row_stride = 1 + n_vq;
  MT_CHECK(H % 8 == 0, "hidden size must be a multiple of 8");
  int threads = H / 8 < 256 ? ((H / 8 + 31) / 32) * 32 : 256;
  MT_CUDA(launch(embed_sum_kernel, dim3(M), dim3(threads), 0, s, ids, emb_text, emb_ext, out, H, n_vq, audio_rows, vocab,
                 row_stride));
  count_launch();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// x_new = residual + bf16(delta)      (delta = bf16 tensor, or the sum of fp32 split-K partials)
// normed = gain * bf16(x_new * rsqrt(mean(x_new^2) + eps))
// One block per token row.
// ---------------------------------------------------------------------------------------------
// v[0..3] += sum over splits (in split order) of the float4 at p + s * stride.  Eight splits' loads are issued before
// the first add: a rolled `for s` loop costs one L2 round trip per split.
__device__ __forceinline__ void sum_splits4(const float* __restrict__ p, long long stride, int nsplit, float (&v)[4]) {
  for (int s0 = 0; s0 < nsplit; s0 += 8) {
    float4 t[8];
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (s0 + u < nsplit) t[u] = __ldcg(reinterpret_cast<const float4*>(p + (s0 + u) * stride));
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (s0 + u < nsplit) { v[0] += t[u].x; v[1] += t[u].y; v[2] += t[u].z; v[3] += t[u].w; }
  }
}

// 512 threads at most: a 1024-thread block at 64 registers takes the whole register file of an SM, so it could
// neither start next to a resident GEMM CTA nor let the next GEMM's CTA start its early weight fetch beside it.
__global__ void __launch_bounds__(512)
add_rmsnorm_generic_kernel(const bf16* __restrict__ residual, const bf16* __restrict__ delta,
                   const float* __restrict__ partials, int nsplit, long long split_stride,
                   bf16* __restrict__ x_out, const bf16* __restrict__ gain, bf16* __restrict__ normed, int H,
                   float eps, unsigned long long* __restrict__ trace) {
  extern __shared__ float srow[];            // H floats + 33 scratch
  float* scratch = srow + H;
  const long long m = blockIdx.x;
  float ss = 0.f;
  auto stamp = [&](int e) {                  // timeline instrumentation (mt_debug_set_trace; null in product runs)
    if (trace && threadIdx.x == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      trace[blockIdx.x * 8 + e] = t;
    }
  };
  stamp(0);
  pdl_trigger();
  pdl_wait();
  stamp(1);
  // 4 elements per thread and iteration: every split's float4 is an independent 16-byte load
#pragma unroll 2
  for (int c = threadIdx.x * 4; c < H; c += blockDim.x * 4) {
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (partials) {
      sum_splits4(partials + m * H + c, split_stride, nsplit, v);
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = rbf(v[j]);
    } else if (delta) {
      const uint2 d = *reinterpret_cast<const uint2*>(delta + m * H + c);
      v[0] = bits_lo(d.x); v[1] = bits_hi(d.x); v[2] = bits_lo(d.y); v[3] = bits_hi(d.y);
    }
    if (residual) {
      const uint2 r = *reinterpret_cast<const uint2*>(residual + m * H + c);
      v[0] = rbf(bits_lo(r.x) + v[0]); v[1] = rbf(bits_hi(r.x) + v[1]);
      v[2] = rbf(bits_lo(r.y) + v[2]); v[3] = rbf(bits_hi(r.y) + v[3]);
    }
    if (x_out) {
      uint2 o; o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]);
      *reinterpret_cast<uint2*>(x_out + m * H + c) = o;
    }
    *reinterpret_cast<float4*>(srow + c) = make_float4(v[0], v[1], v[2], v[3]);
    ss += v[0] * v[0] + v[1] * v[1] + v[2] * v[2] + v[3] * v[3];
  }
  stamp(2);
  ss = block_sum(ss, scratch);
  stamp(3);
  const float inv = rsqrtf(ss / static_cast<float>(H) + eps);
  if (normed) {
    for (int c = threadIdx.x * 4; c < H; c += blockDim.x * 4) {
      const float4 x = *reinterpret_cast<const float4*>(srow + c);
      const uint2 g = *reinterpret_cast<const uint2*>(gain + c);
      uint2 o;
      o.x = pack_bf16(bits_lo(g.x) * rbf(x.x * inv), bits_hi(g.x) * rbf(x.y * inv));
      o.y = pack_bf16(bits_lo(g.y) * rbf(x.z * inv), bits_hi(g.y) * rbf(x.w * inv));
      *reinterpret_cast<uint2*>(normed + m * H + c) = o;
    }
  }
  stamp(4);
}

// Same operation for H <= 2 * 4 * blockDim.x (every LM shape here): each thread owns at most two float4 columns and
// issues EVERY global load it will need (split-K partials of both columns, residual, gain) before the first use.
// The generic kernel above pays four dependent L2 round trips (partials, residual, partials, residual) plus one for
// the gain; this one pays one.
__global__ void __launch_bounds__(512)
add_rmsnorm_kernel(const bf16* __restrict__ residual, const bf16* __restrict__ delta,
                   const float* __restrict__ partials, int nsplit, long long split_stride,
                   bf16* __restrict__ x_out, const bf16* __restrict__ gain, bf16* __restrict__ normed, int H,
                   float eps, unsigned long long* __restrict__ trace) {
  __shared__ float scratch[40];
  const long long m = blockIdx.x;
  auto stamp = [&](int e) {                  // timeline instrumentation (mt_debug_set_trace; null in product runs)
    if (trace && threadIdx.x == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
      trace[blockIdx.x * 8 + e] = t;
    }
  };
  stamp(0);
  pdl_trigger();
  pdl_wait();
  stamp(1);
  constexpr int kIt = 2, kBatch = 8;
  float4 t[kIt][kBatch];
  uint2 rr[kIt], dd[kIt], gg[kIt];
  bool ok[kIt];
#pragma unroll
  for (int it = 0; it < kIt; ++it) {
    const int c = (threadIdx.x + it * blockDim.x) * 4;
    ok[it] = c < H;
    rr[it] = dd[it] = gg[it] = make_uint2(0u, 0u);
    if (ok[it]) {
      if (partials) {
#pragma unroll
        for (int u = 0; u < kBatch; ++u)
          if (u < nsplit) t[it][u] = __ldcg(reinterpret_cast<const float4*>(partials + u * split_stride + m * H + c));
      } else if (delta) {
        dd[it] = *reinterpret_cast<const uint2*>(delta + m * H + c);
      }
      if (residual) rr[it] = *reinterpret_cast<const uint2*>(residual + m * H + c);
      if (normed) gg[it] = *reinterpret_cast<const uint2*>(gain + c);
    }
  }
  float v[kIt][4];
  float ss = 0.f;
#pragma unroll
  for (int it = 0; it < kIt; ++it) {
    v[it][0] = v[it][1] = v[it][2] = v[it][3] = 0.f;
    if (!ok[it]) continue;
    const int c = (threadIdx.x + it * blockDim.x) * 4;
    if (partials) {
#pragma unroll
      for (int u = 0; u < kBatch; ++u)
        if (u < nsplit) { v[it][0] += t[it][u].x; v[it][1] += t[it][u].y; v[it][2] += t[it][u].z; v[it][3] += t[it][u].w; }
      for (int s = kBatch; s < nsplit; ++s) {          // more than eight splits: rare, plain loop
        const float4 x = __ldcg(reinterpret_cast<const float4*>(partials + s * split_stride + m * H + c));
        v[it][0] += x.x; v[it][1] += x.y; v[it][2] += x.z; v[it][3] += x.w;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) v[it][j] = rbf(v[it][j]);
    } else if (delta) {
      v[it][0] = bits_lo(dd[it].x); v[it][1] = bits_hi(dd[it].x); v[it][2] = bits_lo(dd[it].y); v[it][3] = bits_hi(dd[it].y);
    }
    if (residual) {
      v[it][0] = rbf(bits_lo(rr[it].x) + v[it][0]); v[it][1] = rbf(bits_hi(rr[it].x) + v[it][1]);
      v[it][2] = rbf(bits_lo(rr[it].y) + v[it][2]); v[it][3] = rbf(bits_hi(rr[it].y) + v[it][3]);
    }
    if (x_out) {
      uint2 o; o.x = pack_bf16(v[it][0], v[it][1]); o.y = pack_bf16(v[it][2], v[it][3]);
      *reinterpret_cast<uint2*>(x_out + m * H + c) = o;
    }
    ss += v[it][0] * v[it][0] + v[it][1] * v[it][1] + v[it][2] * v[it][2] + v[it][3] * v[it][3];
  }
  stamp(2);
  ss = block_sum(ss, scratch);
  stamp(3);
  const float inv = rsqrtf(ss / static_cast<float>(H) + eps);
  if (normed) {
#pragma unroll
    for (int it = 0; it < kIt; ++it) {
      if (!ok[it]) continue;
      const int c = (threadIdx.x + it * blockDim.x) * 4;
      uint2 o;
      o.x = pack_bf16(bits_lo(gg[it].x) * rbf(v[it][0] * inv), bits_hi(gg[it].x) * rbf(v[it][1] * inv));
      o.y = pack_bf16(bits_lo(gg[it].y) * rbf(v[it][2] * inv), bits_hi(gg[it].y) * rbf(v[it][3] * inv));
      *reinterpret_cast<uint2*>(normed + m * H + c) = o;
    }
  }
  stamp(4);
}

// Prefill form (finished bf16 delta, thousands of rows): one 256-thread block per row, 16 elements per thread
// (two 16-byte loads of delta and of the residual, all issued before the first use), results kept in registers
// between the two passes.  Needs H <= 4096; larger rows take the generic kernel.
__global__ void __launch_bounds__(256)
add_rmsnorm_rows_kernel(const bf16* __restrict__ residual, const bf16* __restrict__ delta, bf16* __restrict__ x_out,
                        const bf16* __restrict__ gain, bf16* __restrict__ normed, int H, float eps) {
  __shared__ float scratch[40];
  const long long m = blockIdx.x;
  pdl_trigger();
  pdl_wait();
  uint4 dd[2], rr[2], gg[2];
  bool ok[2];
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const int c = (threadIdx.x + it * 256) * 8;
    ok[it] = c < H;
    dd[it] = rr[it] = gg[it] = make_uint4(0u, 0u, 0u, 0u);
    if (ok[it]) {
      dd[it] = *reinterpret_cast<const uint4*>(delta + m * H + c);
      if (residual) rr[it] = *reinterpret_cast<const uint4*>(residual + m * H + c);
      if (normed) gg[it] = *reinterpret_cast<const uint4*>(gain + c);
    }
  }
  float v[2][8];
  float ss = 0.f;
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const uint32_t dw[4] = {dd[it].x, dd[it].y, dd[it].z, dd[it].w}, rw[4] = {rr[it].x, rr[it].y, rr[it].z, rr[it].w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float a = bits_lo(dw[j]), b = bits_hi(dw[j]);
      if (residual) { a = rbf(bits_lo(rw[j]) + a); b = rbf(bits_hi(rw[j]) + b); }
      v[it][2 * j] = a; v[it][2 * j + 1] = b;
      ss += a * a + b * b;
    }
    if (ok[it] && x_out) {
      const int c = (threadIdx.x + it * 256) * 8;
      *reinterpret_cast<uint4*>(x_out + m * H + c) = make_uint4(pack_bf16(v[it][0], v[it][1]), pack_bf16(v[it][2], v[it][3]),
                                                               pack_bf16(v[it][4], v[it][5]), pack_bf16(v[it][6], v[it][7]));
    }
  }
  ss = block_sum(ss, scratch);
  const float inv = rsqrtf(ss / static_cast<float>(H) + eps);
  if (normed) {
#pragma unroll
    for (int it = 0; it < 2; ++it) {
      if (!ok[it]) continue;
      const int c = (threadIdx.x + it * 256) * 8;
      const uint32_t gw[4] = {gg[it].x, gg[it].y, gg[it].z, gg[it].w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        o[j] = pack_bf16(bits_lo(gw[j]) * rbf(v[it][2 * j] * inv), bits_hi(gw[j]) * rbf(v[it][2 * j + 1] * inv));
      *reinterpret_cast<uint4*>(normed + m * H + c) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

int add_rmsnorm(const bf16* residual, const bf16* delta, const float* partials, int nsplit, long long split_stride,
                bf16* x_out, const bf16* gain, bf16* normed, int M, int H, float eps, cudaStream_t s) {
  MT_CHECK(H % 4 == 0, "hidden size must be a multiple of 4");
  const size_t smem = (H + 40) * sizeof(float);
  int threads = ((H / 4 + 31) / 32) * 32;
  if (threads > 512) threads = 512;
  if (!partials && delta && M > 256 && H % 8 == 0 && H <= 4096) {
    MT_CUDA(launch(add_rmsnorm_rows_kernel, dim3(M), dim3(256), 0, s, residual, delta, x_out, gain, normed, H, eps));
  } else if (H <= 2 * 4 * threads) {
    MT_CUDA(launch(add_rmsnorm_kernel, dim3(M), dim3(threads), 0, s, residual, delta, partials, nsplit, split_stride,
                   x_out, gain, normed, H, eps, M <= 1024 ? trace_slot(M * 8) : nullptr));
  } else {
    MT_CUDA(launch(add_rmsnorm_generic_kernel, dim3(M), dim3(threads), smem, s, residual, delta, partials, nsplit,
                   split_stride, x_out, gain, normed, H, eps, static_cast<unsigned long long*>(nullptr)));
  }
  count_launch();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// qkv_post: one warp per (token, head).  Input row layout [q heads | k heads | v heads] x D.
//   q,k: RMSNorm over D with gain, then RoPE:  out = bf16(bf16(x*cos) + bf16(rot_half(x)*sin))
//   q -> q_out[m, head, :];  k,v -> cache[(b, kvh, pos, :)]
// Lane l owns elements {l, l+32, ...}; the RoPE partner of element j is j +- D/2, which is the
// same lane for D = 64 and D = 128.
// ---------------------------------------------------------------------------------------------
template <int D>
__global__ void qkv_post_kernel(const bf16* __restrict__ qkv, const float* __restrict__ partials, int nsplit,
                                long long split_stride, const bf16* __restrict__ q_gain,
                                const bf16* __restrict__ k_gain, const bf16* __restrict__ rope_cos,
                                const bf16* __restrict__ rope_sin, bf16* __restrict__ q_out,
                                bf16* __restrict__ k_cache, bf16* __restrict__ v_cache, int M, int T, int n_heads,
                                int n_kv, int past_len, const int* __restrict__ past_len_dev, int max_seq,
                                float eps) {
  constexpr int E = D / 32;
  const long long warp = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int heads_total = n_heads + 2 * n_kv;
  const long long m = warp / heads_total;
  const int hh = static_cast<int>(warp % heads_total);
  pdl_trigger();
  pdl_wait();
  if (m >= M) return;
  const int b = static_cast<int>(m / T), t = static_cast<int>(m % T);
  if (past_len_dev) past_len = *past_len_dev;
  const int pos = past_len + t;
  const long long row_w = static_cast<long long>(heads_total) * D;
  float x[E];
#pragma unroll
  for (int e = 0; e < E; ++e) {
    const long long idx = m * row_w + hh * D + lane + 32 * e;
    if (partials) {
      float a = 0.f;
      for (int s = 0; s < nsplit; ++s) a += partials[s * split_stride + idx];
      x[e] = rbf(a);
    } else {
      x[e] = bf2f(qkv[idx]);
    }
  }
  if (hh < n_heads + n_kv) {
    const bf16* gain = hh < n_heads ? q_gain : k_gain;
    float ss = 0.f;
#pragma unroll
    for (int e = 0; e < E; ++e) ss += x[e] * x[e];
    ss = warp_sum(ss);
    const float inv = rsqrtf(ss / static_cast<float>(D) + eps);
#pragma unroll
    for (int e = 0; e < E; ++e) x[e] = rbf(bf2f(gain[lane + 32 * e]) * rbf(x[e] * inv));
    float y[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const int j = lane + 32 * e;
      const int jh = j % (D / 2);
      const float c = bf2f(rope_cos[static_cast<long long>(pos) * (D / 2) + jh]);
      const float sn = bf2f(rope_sin[static_cast<long long>(pos) * (D / 2) + jh]);
      const float partner = x[(e + E / 2) % E];
      const float rot = j < D / 2 ? -partner : partner;
      y[e] = rbf(rbf(x[e] * c) + rbf(rot * sn));
    }
#pragma unroll
    for (int e = 0; e < E; ++e) x[e] = y[e];
  }
  if (hh < n_heads) {
#pragma unroll
    for (int e = 0; e < E; ++e) q_out[(m * n_heads + hh) * D + lane + 32 * e] = f2bf(x[e]);
  } else {
    const bool is_k = hh < n_heads + n_kv;
    const int kvh = is_k ? hh - n_heads : hh - n_heads - n_kv;
    bf16* dst = (is_k ? k_cache : v_cache) + ((static_cast<long long>(b) * n_kv + kvh) * max_seq + pos) * D;
#pragma unroll
    for (int e = 0; e < E; ++e) dst[lane + 32 * e] = f2bf(x[e]);
  }
}

int qkv_post(const bf16* qkv, const float* partials, int nsplit, long long split_stride, const bf16* q_gain,
             const bf16* k_gain, const bf16* rope_cos, const bf16* rope_sin, bf16* q_out, bf16* k_cache,
             bf16* v_cache, int B, int T, int n_heads, int n_kv, int D, int past_len, const int* past_len_dev,
             int max_seq, float eps, cudaStream_t s) {
  const int heads_total = n_heads + 2 * n_kv;
  const int M = B * T;
  const long long warps = static_cast<long long>(M) * heads_total;
  const int threads = 128;
  const unsigned grid = static_cast<unsigned>(cdivll(warps * 32, threads));
  if (D == 128)
    MT_CUDA(launch(qkv_post_kernel<128>, dim3(grid), dim3(threads), 0, s, qkv, partials, nsplit, split_stride, q_gain,
                   k_gain, rope_cos, rope_sin, q_out, k_cache, v_cache, M, T, n_heads, n_kv, past_len, past_len_dev,
                   max_seq, eps));
  else if (D == 64)
    MT_CUDA(launch(qkv_post_kernel<64>, dim3(grid), dim3(threads), 0, s, qkv, partials, nsplit, split_stride, q_gain,
                   k_gain, rope_cos, rope_sin, q_out, k_cache, v_cache, M, T, n_heads, n_kv, past_len, past_len_dev,
                   max_seq, eps));
  else
    return fail(MT_ERR_UNSUPPORTED, "head_dim %d not supported (64 or 128)", D);
  count_launch();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// GQA attention with the reference's eager rounding points.  One block per (query token, kv head,
// batch row); the G = n_heads/n_kv query heads of the group share every K/V row read.
//   s   = bf16(bf16(q.k) * scale)                     (matmul -> bf16, then * scaling -> bf16)
//   p   = bf16(softmax_fp32(s over visible keys))     (masked keys get exactly 0)
//   out = bf16(sum_k p_k * v_k)                        (fp32 accumulate)
// visible(k) = k <= pos(query) and attn_mask[b,k].  A query with no visible key (left padding)
// writes zeros.  Phase 1: one thread per key; phase 2: one thread per output dim.
// ---------------------------------------------------------------------------------------------
constexpr int kAttThreads = 256;

template <int D, int G>
__global__ void __launch_bounds__(kAttThreads)
attention_kernel(const bf16* __restrict__ q, const bf16* __restrict__ k_cache, const bf16* __restrict__ v_cache,
                 const uint8_t* __restrict__ mask, long long mask_ld, bf16* __restrict__ out, int T, int n_heads,
                 int n_kv, int past_len, const int* __restrict__ past_len_dev, int max_seq, float scale) {
  extern __shared__ float sm[];
  constexpr int NW = kAttThreads / 32;
  constexpr int E = D / 32;               // output dims per lane in phase 2
  const int t = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  pdl_trigger();
  pdl_wait();
  if (past_len_dev) past_len = *past_len_dev;
  const int nkeys = past_len + t + 1;
  const int npad = (nkeys + 3) & ~3;
  float* sq = sm;                         // [G][D]
  float* red = sm + G * D;                // [NW][G][D] partial outputs
  float* sc = red + NW * G * D;           // [G][npad] scores / probabilities
  float* scratch = sc + G * npad;         // 33
  const long long m = static_cast<long long>(b) * T + t;
  for (int i = threadIdx.x; i < G * D; i += blockDim.x) {
    const int g = i / D, d = i % D;
    sq[i] = bf2f(q[(m * n_heads + kvh * G + g) * D + d]);
  }
  __syncthreads();
  const bf16* kbase = k_cache + (static_cast<long long>(b) * n_kv + kvh) * max_seq * D;
  const bf16* vbase = v_cache + (static_cast<long long>(b) * n_kv + kvh) * max_seq * D;
  const uint8_t* mrow = mask + b * mask_ld;
  float lmax[G];
#pragma unroll
  for (int g = 0; g < G; ++g) lmax[g] = -INFINITY;
  // phase 1: one thread per key, 16-byte loads along the K row
  for (int k = threadIdx.x; k < nkeys; k += blockDim.x) {
    if (mrow[k]) {
      float acc[G];
#pragma unroll
      for (int g = 0; g < G; ++g) acc[g] = 0.f;
      const uint4* kr = reinterpret_cast<const uint4*>(kbase + static_cast<long long>(k) * D);
#pragma unroll 4
      for (int d8 = 0; d8 < D / 8; ++d8) {
        const uint4 u = kr[d8];
        const float kv[8] = {bits_lo(u.x), bits_hi(u.x), bits_lo(u.y), bits_hi(u.y),
                             bits_lo(u.z), bits_hi(u.z), bits_lo(u.w), bits_hi(u.w)};
#pragma unroll
        for (int g = 0; g < G; ++g) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[g] = fmaf(sq[g * D + d8 * 8 + j], kv[j], acc[g]);
        }
      }
#pragma unroll
      for (int g = 0; g < G; ++g) {
        const float sv = rbf(rbf(acc[g]) * scale);
        sc[g * npad + k] = sv;
        lmax[g] = fmaxf(lmax[g], sv);
      }
    } else {
#pragma unroll
      for (int g = 0; g < G; ++g) sc[g * npad + k] = -INFINITY;
    }
  }
  float gmax[G], gsum[G];
#pragma unroll
  for (int g = 0; g < G; ++g) gmax[g] = block_max(lmax[g], scratch);
  float lsum[G];
#pragma unroll
  for (int g = 0; g < G; ++g) lsum[g] = 0.f;
  for (int k = threadIdx.x; k < nkeys; k += blockDim.x) {
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const float sv = sc[g * npad + k];
      const float e = (sv == -INFINITY) ? 0.f : __expf(sv - gmax[g]);
      sc[g * npad + k] = e;
      lsum[g] += e;
    }
  }
#pragma unroll
  for (int g = 0; g < G; ++g) gsum[g] = block_sum(lsum[g], scratch);
  for (int k = threadIdx.x; k < nkeys; k += blockDim.x) {
#pragma unroll
    for (int g = 0; g < G; ++g) sc[g * npad + k] = gsum[g] > 0.f ? rbf(sc[g * npad + k] / gsum[g]) : 0.f;
  }
  __syncthreads();
  // phase 2: warp w takes keys w, w+NW, ...; lane owns E consecutive dims (one 4/8-byte load per V row)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float acc[G][E];
#pragma unroll
  for (int g = 0; g < G; ++g)
#pragma unroll
    for (int e = 0; e < E; ++e) acc[g][e] = 0.f;
#pragma unroll 4
  for (int k = warp; k < nkeys; k += NW) {
    float vv[E];
    if constexpr (E == 4) {
      const uint2 u = *reinterpret_cast<const uint2*>(vbase + static_cast<long long>(k) * D + lane * 4);
      vv[0] = bits_lo(u.x); vv[1] = bits_hi(u.x); vv[2] = bits_lo(u.y); vv[3] = bits_hi(u.y);
    } else {
      const uint32_t u = *reinterpret_cast<const uint32_t*>(vbase + static_cast<long long>(k) * D + lane * 2);
      vv[0] = bits_lo(u); vv[1] = bits_hi(u);
    }
#pragma unroll
    for (int g = 0; g < G; ++g) {
      const float p = sc[g * npad + k];
#pragma unroll
      for (int e = 0; e < E; ++e) acc[g][e] = fmaf(p, vv[e], acc[g][e]);
    }
  }
#pragma unroll
  for (int g = 0; g < G; ++g)
#pragma unroll
    for (int e = 0; e < E; ++e) red[(warp * G + g) * D + lane * E + e] = acc[g][e];
  __syncthreads();
  for (int i = threadIdx.x; i < G * D; i += blockDim.x) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) v += red[w * G * D + i];
    const int g = i / D, d = i % D;
    out[(m * n_heads + kvh * G + g) * D + d] = f2bf(v);
  }
}

template <int D, int G>
static int launch_attention(const bf16* q, const bf16* k_cache, const bf16* v_cache, const uint8_t* mask,
                            long long mask_ld, bf16* out, int B, int T, int n_heads, int n_kv, int past_len,
                            const int* past_len_dev, int max_keys, int max_seq, cudaStream_t s) {
  const int npad = (max_keys + 3) & ~3;
  const size_t smem = (static_cast<size_t>(G) * D * (1 + kAttThreads / 32) + static_cast<size_t>(G) * npad + 40) * sizeof(float);
  MT_CHECK(smem <= 200 * 1024, "attention: %d keys exceed the single-pass shared-memory budget", max_keys);
  auto kern = attention_kernel<D, G>;
  static size_t smem_set = 0;
  if (smem > 48 * 1024 && smem > smem_set) {
    MT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    smem_set = 200 * 1024;
  }
  dim3 grid(T, n_kv, B);
  MT_CUDA(launch(kern, grid, dim3(kAttThreads), smem, s, q, k_cache, v_cache, mask, mask_ld, out, T, n_heads, n_kv,
                 past_len, past_len_dev, max_seq, 1.0f / sqrtf(static_cast<float>(D))));
  count_launch();
  return MT_OK;
}

int attention(const bf16* q, const bf16* k_cache, const bf16* v_cache, const uint8_t* mask, long long mask_ld,
              bf16* out, int B, int T, int n_heads, int n_kv, int D, int past_len, const int* past_len_dev,
              int max_keys, int max_seq, cudaStream_t s) {
  if (attention_tc_supported(n_heads, n_kv, D))
    return attention_tc(q, k_cache, v_cache, mask, mask_ld, out, B, T, n_heads, n_kv, D, past_len, past_len_dev, max_seq, s);
  const int G = n_heads / n_kv;
#define MT_ATT(DD, GG)                                                                                   \
  if (D == DD && G == GG)                                                                                \
    return launch_attention<DD, GG>(q, k_cache, v_cache, mask, mask_ld, out, B, T, n_heads, n_kv, past_len, \
                                    past_len_dev, max_keys, max_seq, s);
  MT_ATT(128, 1) MT_ATT(128, 2) MT_ATT(128, 4) MT_ATT(128, 8)
  MT_ATT(64, 1) MT_ATT(64, 2) MT_ATT(64, 4) MT_ATT(64, 8)
#undef MT_ATT
  return fail(MT_ERR_UNSUPPORTED, "attention: head_dim %d with group size %d not supported", D, G);
}

// ---------------------------------------------------------------------------------------------
// act[m,i] = bf16(bf16(silu(gate)) * up),  gate/up = columns [0,I) / [I,2I) of the fused GEMM output
// ---------------------------------------------------------------------------------------------
__global__ void silu_mul_kernel(const bf16* __restrict__ gu, const float* __restrict__ partials, int nsplit,
                                long long split_stride, bf16* __restrict__ act, long long M, int I) {
  const int I4 = I / 4;
  const long long total = M * I4;
  pdl_trigger();
  pdl_wait();
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long m = idx / I4;
    const int i = static_cast<int>(idx % I4) * 4;
    float a[4] = {0.f, 0.f, 0.f, 0.f}, c[4] = {0.f, 0.f, 0.f, 0.f};
    const float* p = partials + m * 2 * I + i;
    for (int s0 = 0; s0 < nsplit; s0 += 4) {        // gate and up of four splits in flight together
      float4 x[4], y[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (s0 + u < nsplit) {
          x[u] = __ldcg(reinterpret_cast<const float4*>(p + (s0 + u) * split_stride));
          y[u] = __ldcg(reinterpret_cast<const float4*>(p + (s0 + u) * split_stride + I));
        }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (s0 + u < nsplit) {
          a[0] += x[u].x; a[1] += x[u].y; a[2] += x[u].z; a[3] += x[u].w;
          c[0] += y[u].x; c[1] += y[u].y; c[2] += y[u].z; c[3] += y[u].w;
        }
    }
    float r[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float g = rbf(a[j]), u = rbf(c[j]);
      r[j] = rbf(g / (1.f + expf(-g))) * u;
    }
    uint2 o; o.x = pack_bf16(r[0], r[1]); o.y = pack_bf16(r[2], r[3]);
    *reinterpret_cast<uint2*>(act + m * I + i) = o;
  }
}

// Prefill form (finished bf16 gate|up): a pure streaming op, 8 outputs per thread and iteration, two iterations'
// 16-byte loads in flight (the one-uint2-per-iteration loop ran at a quarter of the HBM rate).
__global__ void silu_mul_bf16_kernel(const bf16* __restrict__ gu, bf16* __restrict__ act, long long M, int I) {
  const int I8 = I / 8;
  const long long total = M * I8;
  pdl_trigger();
  pdl_wait();
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long idx0 = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx0 < total; idx0 += 2 * stride) {
    uint4 x[2], y[2];
    long long m[2]; int i[2]; bool ok[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const long long idx = idx0 + u * stride;
      ok[u] = idx < total;
      m[u] = ok[u] ? idx / I8 : 0;
      i[u] = ok[u] ? static_cast<int>(idx % I8) * 8 : 0;
      if (ok[u]) {
        x[u] = *reinterpret_cast<const uint4*>(gu + m[u] * 2 * I + i[u]);
        y[u] = *reinterpret_cast<const uint4*>(gu + m[u] * 2 * I + I + i[u]);
      }
    }
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (!ok[u]) continue;
      const uint32_t gx[4] = {x[u].x, x[u].y, x[u].z, x[u].w}, ux[4] = {y[u].x, y[u].y, y[u].z, y[u].w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float g0 = bits_lo(gx[j]), g1 = bits_hi(gx[j]);
        o[j] = pack_bf16(rbf(g0 / (1.f + expf(-g0))) * bits_lo(ux[j]), rbf(g1 / (1.f + expf(-g1))) * bits_hi(ux[j]));
      }
      *reinterpret_cast<uint4*>(act + m[u] * I + i[u]) = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

int silu_mul(const bf16* gu, const float* partials, int nsplit, long long split_stride, bf16* act, int M, int I,
             cudaStream_t s) {
  MT_CHECK(I % 8 == 0, "intermediate size must be a multiple of 8");
  if (partials) {
    const long long total = static_cast<long long>(M) * (I / 4);
    int blocks = static_cast<int>(cdivll(total, 256));
    if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
    MT_CUDA(launch(silu_mul_kernel, dim3(blocks), dim3(256), 0, s, gu, partials, nsplit, split_stride, act,
                   static_cast<long long>(M), I));
  } else {
    const long long total = static_cast<long long>(M) * (I / 8);
    int blocks = static_cast<int>(cdivll(total, 512));
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    MT_CUDA(launch(silu_mul_bf16_kernel, dim3(blocks), dim3(256), 0, s, gu, act, static_cast<long long>(M), I));
  }
  count_launch();
  return MT_OK;
}

// audio_logits[..., last] = -inf  (modeling_moss_tts.py:298-299)
__global__ void fill_audio_pad_kernel(bf16* logits, long long rows, int head_size) {
  const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  pdl_trigger();
  pdl_wait();
  if (i < rows) logits[i * head_size + head_size - 1] = f2bf(-INFINITY);
}
int fill_audio_pad(bf16* logits, long long rows, int head_size, cudaStream_t s) {
  MT_CUDA(launch(fill_audio_pad_kernel, dim3(static_cast<unsigned>(cdivll(rows, 256))), dim3(256), 0, s, logits, rows,
                 head_size));
  count_launch();
  return MT_OK;
}

// gather the last token of every batch row: dst[b,:] = src[b*T + T-1, :]
__global__ void gather_last_kernel(const bf16* __restrict__ src, bf16* __restrict__ dst, int T, int H) {
  const int b = blockIdx.x;
  pdl_trigger();
  pdl_wait();
  for (int c = threadIdx.x; c < H; c += blockDim.x)
    dst[static_cast<long long>(b) * H + c] = src[(static_cast<long long>(b) * T + T - 1) * H + c];
}
int gather_last(const bf16* src, bf16* dst, int B, int T, int H, cudaStream_t s) {
  MT_CUDA(launch(gather_last_kernel, dim3(B), dim3(256), 0, s, src, dst, T, H));
  count_launch();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// MossTTSLocal helpers
//   moss_rmsnorm: MossTTSRMSNorm (reference moss_tts_local/modeling_moss_tts.py:34-44) - unlike Qwen3RMSNorm every
//     operation runs in the working dtype: p = bf16(x*x); m = bf16(sum p / H)  (fp32 accumulate); r = bf16(rsqrt(bf16(m + eps)));
//     y = bf16(bf16(x*r) * gain).  x comes as bf16 or as split-K partials of the producing Linear.
//   embed_rows: out[b,:] = table[ids[b*stride + col], :]   (embedding_list[c] lookup of the token just sampled, :417-420)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024)
moss_rmsnorm_kernel(const bf16* __restrict__ x, const float* __restrict__ partials, int nsplit, long long split_stride,
                    const bf16* __restrict__ gain, bf16* __restrict__ out, int H, float eps) {
  extern __shared__ float srow[];
  float* scratch = srow + H;
  const long long m = blockIdx.x;
  pdl_trigger();
  pdl_wait();
  float ss = 0.f;
  for (int c = threadIdx.x; c < H; c += blockDim.x) {
    float v;
    if (partials) {
      float a = 0.f;
      for (int s = 0; s < nsplit; ++s) a += partials[s * split_stride + m * H + c];
      v = rbf(a);
    } else {
      v = bf2f(x[m * H + c]);
    }
    srow[c] = v;
    ss += rbf(v * v);
  }
  ss = block_sum(ss, scratch);
  const float mean = rbf(ss / static_cast<float>(H));      // torch's CPU/CUDA mean: fp32 sum, fp32 divide, one rounding
  const float r = rbf(1.0f / sqrtf(rbf(mean + eps)));
  for (int c = threadIdx.x; c < H; c += blockDim.x)
    out[m * H + c] = f2bf(rbf(srow[c] * r) * bf2f(gain[c]));
}

int moss_rmsnorm(const bf16* x, const float* partials, int nsplit, long long split_stride, const bf16* gain, bf16* out,
                 int M, int H, float eps, cudaStream_t s) {
  const size_t smem = (H + 40) * sizeof(float);
  int threads = ((H + 31) / 32) * 32;
  if (threads > 1024) threads = 1024;
  MT_CUDA(launch(moss_rmsnorm_kernel, dim3(M), dim3(threads), smem, s, x, partials, nsplit, split_stride, gain, out, H, eps));
  count_launch();
  return MT_OK;
}

__global__ void embed_rows_kernel(const int64_t* __restrict__ ids, int stride, int col, const bf16* __restrict__ table,
                                  long long rows, bf16* __restrict__ out, int H) {
  const int b = blockIdx.x;
  pdl_trigger();
  pdl_wait();
  long long id = ids[static_cast<long long>(b) * stride + col];
  id = id < 0 ? 0 : (id >= rows ? rows - 1 : id);
  const uint4* src = reinterpret_cast<const uint4*>(table + id * H);
  uint4* dst = reinterpret_cast<uint4*>(out + static_cast<long long>(b) * H);
  for (int c = threadIdx.x; c < H / 8; c += blockDim.x) dst[c] = src[c];
}

int embed_rows(const int64_t* ids, int stride, int col, const bf16* table, long long rows, bf16* out, int B, int H,
               cudaStream_t s) {
  MT_CHECK(H % 8 == 0, "hidden size must be a multiple of 8");
  MT_CUDA(launch(embed_rows_kernel, dim3(B), dim3(128), 0, s, ids, stride, col, table, rows, out, H));
  count_launch();
  return MT_OK;
}

}  // namespace mt
