// MOSS-Audio-Tokenizer decode-side kernels (fp32), reference
// moss_audio_tokenizer/modeling_moss_audio_tokenizer.py:
//   a10 RVQ dequantise    ResidualLFQ.decode_codes :1276-1283, LFQ.decode_code :1065-1074, WNConv1d :967-969
//   a11 LayerNorm :237-247, interleaved RoPE apply_rope :255-300, causal sliding-window attention :627-666
//       (plus the RingKVCache visibility of chunked decode :413-480,:1728-1750)
//   a12 patch upsample    PatchedPretransform.decode :947-953
// Activations are token-major [B*T, C]; the reference's (B,C,T) transposes disappear.
#include "codec_ops.cuh"

namespace mt {

// ---------------------------------------------------------------------------------------------
// table[q][code][c] = sum_r Wo[c][r] * (sum_d Wq[q][r][d] * C[q][code][d] + bq[q][r]),
// W = g * v / ||v|| (weight_norm over the input dim).  One block per (code, quantizer); one-time.
// ---------------------------------------------------------------------------------------------
__global__ void rvq_fold_kernel(const float* __restrict__ codebook, const float* __restrict__ q_g,
                                const float* __restrict__ q_v, const float* __restrict__ q_b,
                                const float* __restrict__ o_g, const float* __restrict__ o_v, int cs, int cd, int rvq,
                                int code_dim, float* __restrict__ table) {
  extern __shared__ float emb[];   // rvq
  const int code = blockIdx.x, q = blockIdx.y;
  const float* cb = codebook + (static_cast<long long>(q) * cs + code) * cd;
  for (int r = threadIdx.x; r < rvq; r += blockDim.x) {
    const float* v = q_v + (static_cast<long long>(q) * rvq + r) * cd;
    float nn = 0.f, dot = 0.f;
    for (int d = 0; d < cd; ++d) { nn += v[d] * v[d]; dot += v[d] * cb[d]; }
    emb[r] = q_g[q * rvq + r] * dot / sqrtf(nn) + q_b[q * rvq + r];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < code_dim; c += blockDim.x) {
    const float* v = o_v + static_cast<long long>(c) * rvq;
    float nn = 0.f, dot = 0.f;
    for (int r = 0; r < rvq; ++r) { nn += v[r] * v[r]; dot += v[r] * emb[r]; }
    table[(static_cast<long long>(q) * cs + code) * code_dim + c] = o_g[c] * dot / sqrtf(nn);
  }
}

int rvq_fold_tables(const float* codebook, const float* q_g, const float* q_v, const float* q_b, const float* o_g,
                    const float* o_v, int nq, int cs, int cd, int rvq, int code_dim, float* table, cudaStream_t s) {
  dim3 grid(cs, nq);
  rvq_fold_kernel<<<grid, 256, rvq * sizeof(float), s>>>(codebook, q_g, q_v, q_b, o_g, o_v, cs, cd, rvq, code_dim, table);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// out[frame, :] = bias + sum_{q<nq} table[q][codes[q, frame]][:]     (quantizer order, fp32)
// HBM/L2-bound gather-sum: nq * code_dim * 4 B gathered + code_dim * 4 B written per frame (101,376 B at nq = 32,
// code_dim = 768).  One block = kRvqFrames consecutive frames, one float4 lane per 4 channels.  The block's
// nq x kRvqFrames codes are staged once into shared memory (each quantizer's codes for the block are one 32 B
// sector), then every thread keeps 8 independent 16 B table gathers in flight per frame (the table rows of 8
// quantizers), summed in quantizer order.  Channel-major output ([B, code_dim, T], the reference's layout for
// decode_codes) writes one float4 of 4 consecutive frames per channel when T % 4 == 0.
// ---------------------------------------------------------------------------------------------
constexpr int kRvqFrames = 4;
__global__ void __launch_bounds__(256)
rvq_dequant_kernel(const int64_t* __restrict__ codes, const float* __restrict__ table, const float* __restrict__ bias,
                   int nq, long long q_stride, long long frames, int T, int cs, int code_dim, int channel_major,
                   float* __restrict__ out) {
  extern __shared__ int s_code[];   // [nq][kRvqFrames]
  const long long f0 = blockIdx.x * static_cast<long long>(kRvqFrames);
  const int nf = frames - f0 < kRvqFrames ? static_cast<int>(frames - f0) : kRvqFrames;
  for (int i = threadIdx.x; i < nq * kRvqFrames; i += blockDim.x) {
    const int q = i / kRvqFrames, j = i % kRvqFrames;
    long long code = j < nf ? codes[q * q_stride + f0 + j] : 0;
    code = code < 0 ? 0 : (code >= cs ? cs - 1 : code);
    s_code[i] = static_cast<int>(code);
  }
  __syncthreads();
  const int c4 = threadIdx.x;
  if (c4 * 4 >= code_dim) return;
  const float4 b = *reinterpret_cast<const float4*>(bias + c4 * 4);
  const float* tcol = table + c4 * 4;
  float4 acc[kRvqFrames];
#pragma unroll
  for (int j = 0; j < kRvqFrames; ++j) {
    acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (j >= nf) continue;
    for (int q0 = 0; q0 < nq; q0 += 8) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int q = q0 + u;
        v[u] = q < nq ? __ldg(reinterpret_cast<const float4*>(
                            tcol + (static_cast<long long>(q) * cs + s_code[q * kRvqFrames + j]) * code_dim))
                      : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) { acc[j].x += v[u].x; acc[j].y += v[u].y; acc[j].z += v[u].z; acc[j].w += v[u].w; }
    }
    acc[j].x += b.x; acc[j].y += b.y; acc[j].z += b.z; acc[j].w += b.w;
  }
  if (!channel_major) {
#pragma unroll
    for (int j = 0; j < kRvqFrames; ++j)
      if (j < nf) __stcs(reinterpret_cast<float4*>(out + (f0 + j) * code_dim + c4 * 4), acc[j]);
  } else if ((T & 3) == 0 && nf == kRvqFrames) {      // 4 consecutive frames of one utterance: 16 B per channel
    const long long bidx = f0 / T, t = f0 % T;
    float* o = out + (bidx * code_dim + c4 * 4) * T + t;
    __stcs(reinterpret_cast<float4*>(o), make_float4(acc[0].x, acc[1].x, acc[2].x, acc[3].x));
    __stcs(reinterpret_cast<float4*>(o + T), make_float4(acc[0].y, acc[1].y, acc[2].y, acc[3].y));
    __stcs(reinterpret_cast<float4*>(o + 2ll * T), make_float4(acc[0].z, acc[1].z, acc[2].z, acc[3].z));
    __stcs(reinterpret_cast<float4*>(o + 3ll * T), make_float4(acc[0].w, acc[1].w, acc[2].w, acc[3].w));
  } else {
#pragma unroll
    for (int j = 0; j < kRvqFrames; ++j) {
      if (j >= nf) continue;
      const long long bidx = (f0 + j) / T, t = (f0 + j) % T;
      float* o = out + (bidx * code_dim + c4 * 4) * T + t;
      o[0] = acc[j].x; o[T] = acc[j].y; o[2ll * T] = acc[j].z; o[3ll * T] = acc[j].w;
    }
  }
}

static int rvq_launch(const int64_t* codes, const float* table, const float* bias, int nq, long long q_stride,
                      long long frames, int T, int cs, int code_dim, int channel_major, float* out, cudaStream_t s) {
  MT_CHECK(code_dim % 4 == 0 && code_dim / 4 <= 256, "rvq_dequant: code_dim must be a multiple of 4 and <= 1024");
  MT_CHECK(nq > 0 && nq <= 1024, "rvq_dequant: bad quantizer count %d", nq);
  if (frames == 0) return MT_OK;
  const int threads = ((code_dim / 4 + 31) / 32) * 32;
  const long long blocks = (frames + kRvqFrames - 1) / kRvqFrames;
  rvq_dequant_kernel<<<static_cast<unsigned>(blocks), threads, nq * kRvqFrames * sizeof(int), s>>>(
      codes, table, bias, nq, q_stride, frames, T, cs, code_dim, channel_major, out);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

int rvq_dequant(const int64_t* codes, const float* table, const float* bias, int nq, int B, int T, int cs, int code_dim,
                int channel_major, float* out, cudaStream_t s) {
  const long long frames = static_cast<long long>(B) * T;
  return rvq_launch(codes, table, bias, nq, frames, frames, T, cs, code_dim, channel_major, out, s);
}

// token-major output for `frames` consecutive frames of a [nq, q_stride] code matrix
int rvq_dequant_strided(const int64_t* codes, const float* table, const float* bias, int nq, long long q_stride,
                        long long frames, int cs, int code_dim, float* out, cudaStream_t s) {
  return rvq_launch(codes, table, bias, nq, q_stride, frames, 1, cs, code_dim, 0, out, s);
}

// ---------------------------------------------------------------------------------------------
// LayerNorm (eps 1e-5, affine), one block per token, two-pass in fp32
// ---------------------------------------------------------------------------------------------
__global__ void layernorm_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                                 float* __restrict__ out, int d, float eps, int out_f16) {
  extern __shared__ float srow[];
  float* scratch = srow + d;
  const long long m = blockIdx.x;
  float s = 0.f;
  for (int c = threadIdx.x; c < d; c += blockDim.x) { const float v = x[m * d + c]; srow[c] = v; s += v; }
  const float mean = block_sum(s, scratch) / d;
  float ss = 0.f;
  for (int c = threadIdx.x; c < d; c += blockDim.x) { const float t = srow[c] - mean; ss += t * t; }
  const float rstd = rsqrtf(block_sum(ss, scratch) / d + eps);
  for (int c = threadIdx.x; c < d; c += blockDim.x) {
    const float y = (srow[c] - mean) * rstd * w[c] + b[c];
    if (out_f16) reinterpret_cast<__half*>(out)[m * d + c] = __float2half_rn(y);
    else out[m * d + c] = y;
  }
}
// Warp-per-token form for d <= 32 * 4 * kLnVec (the codec's 768 / 1280): the row lives in registers (float4 loads, all
// issued before the first use), mean / variance by warp shuffles, no shared memory and no block barriers; 8 tokens per
// block.  Same two-pass fp32 formula as above (the summation order differs).
constexpr int kLnVec = 10;
template <int NV>
__global__ void __launch_bounds__(256)
layernorm_warp_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                      float* __restrict__ out, long long M, int d, float eps, int out_f16) {
  const int lane = threadIdx.x & 31;
  const long long m = blockIdx.x * 8ll + (threadIdx.x >> 5);
  pdl_trigger();
  pdl_wait();
  if (m >= M) return;
  const int nvec = d / 4;
  const float4* xr = reinterpret_cast<const float4*>(x + m * d);
  float4 v[NV];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + 32 * i;
    v[i] = c < nvec ? xr[c] : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  const float mean = warp_sum(s) / d;
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    if (lane + 32 * i < nvec) {
      const float a = v[i].x - mean, bb = v[i].y - mean, c2 = v[i].z - mean, e = v[i].w - mean;
      ss += (a * a + bb * bb) + (c2 * c2 + e * e);
    }
  }
  const float rstd = rsqrtf(warp_sum(ss) / d + eps);
  float4* orow = reinterpret_cast<float4*>(out + m * d);
  const float4* w4 = reinterpret_cast<const float4*>(w);
  const float4* b4 = reinterpret_cast<const float4*>(b);
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) {
      const float4 ww = w4[c], bb = b4[c];
      const float4 y = make_float4((v[i].x - mean) * rstd * ww.x + bb.x, (v[i].y - mean) * rstd * ww.y + bb.y,
                                   (v[i].z - mean) * rstd * ww.z + bb.z, (v[i].w - mean) * rstd * ww.w + bb.w);
      if (out_f16) reinterpret_cast<uint2*>(reinterpret_cast<__half*>(out) + m * d)[c] = make_uint2(pack_f16(y.x, y.y), pack_f16(y.z, y.w));
      else orow[c] = y;
    }
  }
}
int layernorm(const float* x, const float* w, const float* b, float* out, long long M, int d, cudaStream_t s, int out_f16) {
  if (d % 4 == 0 && d <= 32 * 4 * kLnVec) {
    const unsigned blocks = static_cast<unsigned>((M + 7) / 8);
    if (d <= 32 * 4 * 6) MT_CUDA(launch(layernorm_warp_kernel<6>, dim3(blocks), dim3(256), 0, s, x, w, b, out, M, d, 1e-5f, out_f16));
    else MT_CUDA(launch(layernorm_warp_kernel<kLnVec>, dim3(blocks), dim3(256), 0, s, x, w, b, out, M, d, 1e-5f, out_f16));
    count_launch();
    return MT_OK;
  }
  layernorm_kernel<<<static_cast<unsigned>(M), 256, (d + 40) * sizeof(float), s>>>(x, w, b, out, d, 1e-5f, out_f16);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// In-place interleaved-pair RoPE on the q and k thirds of qkv [B*T, 3, heads, hd]; position = t.
// freq_j = exp(-ln(max_period) * 2j / hd)
// ---------------------------------------------------------------------------------------------
__global__ void codec_rope_kernel(float* __restrict__ qkv, long long M, int T, int d, int hd, float log_period) {
  const long long total = M * d;   // pairs over q and k: 2 * (d/2) per token
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long m = idx / d;
    const int r = static_cast<int>(idx % d);
    const int which = r / (d / 2);            // 0 = q, 1 = k
    const int pr = r % (d / 2);               // pair index inside the d-wide vector
    const int j = pr % (hd / 2);
    const int t = static_cast<int>(m % T);
    const float freq = expf(static_cast<float>(j) * (-log_period * 2.f / hd));
    float sn, cs;
    sincosf(freq * static_cast<float>(t), &sn, &cs);
    float2* p = reinterpret_cast<float2*>(qkv + m * 3 * d + static_cast<long long>(which) * d + 2 * pr);
    const float2 v = *p;
    *p = make_float2(v.x * cs - v.y * sn, v.x * sn + v.y * cs);
  }
}
__global__ void codec_rope_table_kernel(float2* __restrict__ table, int T, int hd, float log_period, long long pos0) {
  const int total = T * (hd / 2);
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
    const int t = idx / (hd / 2), j = idx % (hd / 2);
    const float freq = expf(static_cast<float>(j) * (-log_period * 2.f / hd));     // same expressions as codec_rope_kernel
    float sn, cs;
    sincosf(freq * static_cast<float>(pos0 + t), &sn, &cs);     // fp32 angle of the absolute position, as the reference
    table[idx] = make_float2(cs, sn);
  }
}
int codec_rope_table(float* table, int T, int hd, float max_period, cudaStream_t s, long long pos0) {
  const int total = T * (hd / 2);
  codec_rope_table_kernel<<<(total + 255) / 256, 256, 0, s>>>(reinterpret_cast<float2*>(table), T, hd, logf(max_period), pos0);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

int codec_rope(float* qkv, long long M, int T, int d, int hd, float max_period, cudaStream_t s) {
  const long long total = M * d;
  int blocks = static_cast<int>(cdivll(total, 256));
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  codec_rope_kernel<<<blocks, 256, 0, s>>>(qkv, M, T, d, hd, logf(max_period));
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// Causal sliding-window attention, head_dim 64, fp32, one thread per query.
// Block = 128 consecutive queries of one (batch, head); K/V tiles of 32 keys are staged in shared
// memory and every thread reads the same K/V element at the same time (a broadcast, conflict-free).
// Each thread keeps q[64], o[64] and the online-softmax (m, l) in registers: no cross-thread traffic.
// visible(q, k) = lo(q) <= k <= q, lo(q) = max(q-context+1, chunk bound, 0); the chunk bound
// reproduces the ring cache of decode(chunk_duration): start(q)+len(q)-context.
// ---------------------------------------------------------------------------------------------
constexpr int kAttQ = 128, kAttK = 32, kHD = 64;

__device__ __forceinline__ int vis_lo(int t, int T, int context, int chunk) {
  int lo = t - context + 1;
  if (chunk > 0 && T > chunk) {
    const int start = (t / chunk) * chunk;
    const int len = min(chunk, T - start);
    lo = max(lo, start + len - context);
  }
  return max(lo, 0);
}

__global__ void __launch_bounds__(kAttQ)
codec_attention_kernel(const float* __restrict__ qkv, float* __restrict__ out, int T, int heads, int context, int chunk,
                       float scale) {
  __shared__ float sk[kAttK][kHD];
  __shared__ float sv[kAttK][kHD];
  const int q0 = blockIdx.x * kAttQ, h = blockIdx.y, b = blockIdx.z;
  const int d = heads * kHD;
  const int t = q0 + threadIdx.x;
  const bool valid = t < T;
  const long long tok0 = static_cast<long long>(b) * T;
  float q[kHD], o[kHD];
  if (valid) {
    const float4* qp = reinterpret_cast<const float4*>(qkv + (tok0 + t) * 3 * d + h * kHD);
#pragma unroll
    for (int i = 0; i < kHD / 4; ++i) {
      const float4 v = qp[i];
      q[4 * i] = v.x * scale; q[4 * i + 1] = v.y * scale; q[4 * i + 2] = v.z * scale; q[4 * i + 3] = v.w * scale;
    }
  }
#pragma unroll
  for (int i = 0; i < kHD; ++i) o[i] = 0.f;
  float mx = -INFINITY, l = 0.f;
  const int my_lo = valid ? vis_lo(t, T, context, chunk) : 0x7fffffff;
  // key range of the block: from the smallest lo over its queries (lo is non-decreasing in t except
  // at chunk starts, where it can only jump up) to its last query
  const int last_q = min(q0 + kAttQ, T) - 1;
  int k_begin = vis_lo(q0, T, context, chunk);
  k_begin = (k_begin / kAttK) * kAttK;
  for (int k0 = k_begin; k0 <= last_q; k0 += kAttK) {
    __syncthreads();
    for (int i = threadIdx.x; i < kAttK * kHD / 4; i += kAttQ) {
      const int kk = i / (kHD / 4), c = i % (kHD / 4);
      const int kt = k0 + kk;
      float4 kv4 = make_float4(0.f, 0.f, 0.f, 0.f), vv4 = kv4;
      if (kt < T) {
        const float* base = qkv + (tok0 + kt) * 3 * d + h * kHD;
        kv4 = reinterpret_cast<const float4*>(base + d)[c];
        vv4 = reinterpret_cast<const float4*>(base + 2 * d)[c];
      }
      reinterpret_cast<float4*>(&sk[kk][0])[c] = kv4;
      reinterpret_cast<float4*>(&sv[kk][0])[c] = vv4;
    }
    __syncthreads();
    if (!valid) continue;
    const int ka = max(k0, my_lo), kb = min(k0 + kAttK - 1, t);
    for (int kt = ka; kt <= kb; ++kt) {
      const int kk = kt - k0;
      float s = 0.f;
#pragma unroll
      for (int i = 0; i < kHD; ++i) s = fmaf(q[i], sk[kk][i], s);
      if (s > mx) {
        const float corr = __expf(mx - s);
        l *= corr;
#pragma unroll
        for (int i = 0; i < kHD; ++i) o[i] *= corr;
        mx = s;
      }
      const float p = __expf(s - mx);
      l += p;
#pragma unroll
      for (int i = 0; i < kHD; ++i) o[i] = fmaf(p, sv[kk][i], o[i]);
    }
  }
  if (valid) {
    const float inv = l > 0.f ? 1.f / l : 0.f;
    float4* op = reinterpret_cast<float4*>(out + (tok0 + t) * d + h * kHD);
#pragma unroll
    for (int i = 0; i < kHD / 4; ++i)
      op[i] = make_float4(o[4 * i] * inv, o[4 * i + 1] * inv, o[4 * i + 2] * inv, o[4 * i + 3] * inv);
  }
}

int codec_attention(const float* qkv, float* out, int B, int T, int heads, int hd, int context, int chunk,
                    cudaStream_t s) {
  if (hd != kHD) return fail(MT_ERR_UNSUPPORTED, "codec attention supports head_dim 64 only (got %d)", hd);
  dim3 grid(cdiv(T, kAttQ), heads, B);
  codec_attention_kernel<<<grid, kAttQ, 0, s>>>(qkv, out, T, heads, context, chunk, 1.0f / sqrtf(static_cast<float>(hd)));
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// debug probe (mt_codec_debug_set_absmax): max |x| over an fp16 operand buffer, atomically folded into *slot
// (non-negative floats order like their bit patterns; a NaN lands above +inf and stays visible)
// ---------------------------------------------------------------------------------------------
__global__ void absmax_f16_kernel(const __half2* __restrict__ x, long long n2, unsigned int* __restrict__ slot) {
  float m = 0.f;
  bool nan = false;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n2;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float2 v = __half22float2(x[i]);
    nan |= (v.x != v.x) || (v.y != v.y);
    m = fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y)));
  }
  m = warp_max(m);
  nan = __any_sync(0xffffffffu, nan);
  if ((threadIdx.x & 31) == 0) atomicMax(slot, nan ? 0x7fc00000u : __float_as_uint(m));
}
int absmax_f16(const void* x, long long n, float* slot, cudaStream_t s) {
  int blocks = static_cast<int>(cdivll(n / 2, 256 * 8));
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  if (blocks < 1) blocks = 1;
  absmax_f16_kernel<<<blocks, 256, 0, s>>>(reinterpret_cast<const __half2*>(x), n / 2, reinterpret_cast<unsigned int*>(slot));
  MT_LAUNCH_CHECK();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// patch upsample in token-major form: x'[b, l*h + j, c] = y[b, l, c*h + j]
// ---------------------------------------------------------------------------------------------
__global__ void unpatch_kernel(const float* __restrict__ y, float* __restrict__ x, long long tokens, int dh, int h) {
  const int d = dh / h;
  const long long total = tokens * dh;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long tok = idx / dh;
    const int r = static_cast<int>(idx % dh);     // r = j*d + c in the output
    const int j = r / d, c = r % d;
    x[idx] = y[tok * dh + c * h + j];
  }
}
int unpatch(const float* y, float* x, long long tokens, int dh, int h, cudaStream_t s) {
  const long long total = tokens * dh;
  int blocks = static_cast<int>(cdivll(total, 256));
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  unpatch_kernel<<<blocks, 256, 0, s>>>(y, x, tokens, dh, h);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

}  // namespace mt
