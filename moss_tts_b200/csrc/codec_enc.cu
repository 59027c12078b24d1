// MOSS-Audio-Tokenizer encode-side kernels (fp32), reference
// moss_audio_tokenizer/modeling_moss_audio_tokenizer.py:
//   patch downsample      PatchedPretransform.encode :936-944
//   residual LFQ quantise ResidualLFQ.forward :1238-1274, LFQ.forward / decode_latents :1055-1092
// Activations are token-major [B*T, C] like the decode side (codec_ops.cu).
#include "codec_ops.cuh"

namespace mt {

// ---------------------------------------------------------------------------------------------
// patch downsample in token-major form: y[b, l, c*h + j] = x[b, l*h + j, c]   (inverse of unpatch_kernel)
// ---------------------------------------------------------------------------------------------
__global__ void patchify_kernel(const float* __restrict__ x, float* __restrict__ y, long long tokens_out, int d, int h) {
  const int dh = d * h;
  const long long total = tokens_out * dh;
  pdl_trigger();
  pdl_wait();
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long tok = idx / dh;
    const int r = static_cast<int>(idx % dh);     // r = j*d + c in the input
    const int j = r / d, c = r % d;
    y[tok * dh + c * h + j] = x[idx];             // coalesced reads, stride-h writes inside one output row
  }
}
int patchify(const float* x, float* y, long long tokens_out, int d, int h, cudaStream_t s) {
  const long long total = tokens_out * d * h;
  if (total == 0) return MT_OK;
  int blocks = static_cast<int>(cdivll(total, 256));
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  MT_CUDA(launch(patchify_kernel, dim3(blocks), dim3(256), 0, s, x, y, tokens_out, d, h));
  count_launch();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// Residual LFQ quantise.  One block = kQF consecutive frames; the frames' residuals [kQF][R] live in shared memory for
// all nq quantisers, so the only global traffic is z once (R*4 B per frame), the codes (nq*8 B per frame) and the
// quantiser parameters (64 KB per quantiser at the released shape: L1/L2 hits shared by every block).
// Per quantiser q, per frame (reference :1258-1266 with :1058-1062 and :1076-1092):
//   z_e   = W_in[q] (residual * mask) + b_in[q]                      [CD]
//   e     = z_e / max(||z_e||, 1e-12)                                 F.normalize
//   dist_c= (sum e^2 - 2 e.c^_c) + sum c^_c^2,  c^ = F.normalize(codebook)   -> index = first argmax of -dist
//   z_q   = W_out[q] codebook[index] + b_out[q];  residual -= z_q * mask
// Frames past the sample's length (mask = 0) quantise the bias alone and leave the residual untouched, as the
// reference does.  Ties resolve to the lowest index.
// ---------------------------------------------------------------------------------------------
constexpr int kQF = 4;
constexpr int kQThreads = 256;

__device__ __forceinline__ unsigned int ordered_key(float v) {
  const unsigned int u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

template <int CD>
__global__ void __launch_bounds__(kQThreads)
rlfq_quantize_kernel(const float* __restrict__ z, long long M, int T, const int64_t* __restrict__ lengths, int nq, int R,
                     int CS, QuantWeights w, int64_t* __restrict__ codes, long long q_stride) {
  extern __shared__ float s_r[];                       // [kQF][R]
  __shared__ float s_ze[kQF][CD];
  __shared__ float s_zq[kQF][CD];
  __shared__ unsigned long long s_best[kQF][kQThreads / 32];
  __shared__ int s_idx[kQF];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long m0 = blockIdx.x * static_cast<long long>(kQF);
  pdl_trigger();
  pdl_wait();
  bool valid[kQF];
#pragma unroll
  for (int j = 0; j < kQF; ++j) {
    const long long m = m0 + j;
    valid[j] = m < M && (lengths == nullptr || (m % T) < lengths[m / T]);
  }
  for (int i = tid; i < kQF * R; i += kQThreads) {
    const long long m = m0 + i / R;
    s_r[i] = m < M ? z[m * R + (i % R)] : 0.f;
  }
  __syncthreads();
  for (int q = 0; q < nq; ++q) {
    // ---- in_proj: warp jo owns output jo of all kQF frames
    const float* Win = w.q_in_w + static_cast<long long>(q) * CD * R;
    for (int jo = warp; jo < CD; jo += kQThreads / 32) {
      float acc[kQF];
#pragma unroll
      for (int j = 0; j < kQF; ++j) acc[j] = 0.f;
      for (int c = lane; c < R; c += 32) {
        const float wv = __ldg(Win + jo * R + c);
#pragma unroll
        for (int j = 0; j < kQF; ++j) acc[j] = fmaf(wv, valid[j] ? s_r[j * R + c] : 0.f, acc[j]);
      }
#pragma unroll
      for (int j = 0; j < kQF; ++j) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
      }
      if (lane == 0) {
        const float b = __ldg(w.q_in_b + q * CD + jo);
#pragma unroll
        for (int j = 0; j < kQF; ++j) s_ze[j][jo] = acc[j] + b;
      }
    }
    __syncthreads();
    // ---- normalise (every thread, all frames) and scan the codebook
    float en[kQF][CD], esq[kQF];
#pragma unroll
    for (int j = 0; j < kQF; ++j) {
      float n2 = 0.f;
#pragma unroll
      for (int k = 0; k < CD; ++k) n2 = fmaf(s_ze[j][k], s_ze[j][k], n2);
      const float dn = fmaxf(sqrtf(n2), 1e-12f);
      esq[j] = 0.f;
#pragma unroll
      for (int k = 0; k < CD; ++k) { en[j][k] = s_ze[j][k] / dn; esq[j] = fmaf(en[j][k], en[j][k], esq[j]); }
    }
    unsigned long long best[kQF];
#pragma unroll
    for (int j = 0; j < kQF; ++j) best[j] = 0ull;
    const float* cbn = w.codebook_n + static_cast<long long>(q) * CS * CD;
    for (int c = tid; c < CS; c += kQThreads) {
      float cv[CD];
#pragma unroll
      for (int k = 0; k < CD; k += 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(cbn + static_cast<long long>(c) * CD + k));
        cv[k] = t.x; cv[k + 1] = t.y; cv[k + 2] = t.z; cv[k + 3] = t.w;
      }
      const float csq = __ldg(w.codebook_nsq + static_cast<long long>(q) * CS + c);
#pragma unroll
      for (int j = 0; j < kQF; ++j) {
        float dot = 0.f;
#pragma unroll
        for (int k = 0; k < CD; ++k) dot = fmaf(en[j][k], cv[k], dot);
        const float dist = (esq[j] - 2.f * dot) + csq;
        const unsigned long long p = (static_cast<unsigned long long>(ordered_key(-dist)) << 32) |
                                     (0xFFFFFFFFu - static_cast<unsigned int>(c));
        best[j] = p > best[j] ? p : best[j];
      }
    }
#pragma unroll
    for (int j = 0; j < kQF; ++j) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long t = __shfl_xor_sync(0xffffffffu, best[j], o);
        best[j] = t > best[j] ? t : best[j];
      }
      if (lane == 0) s_best[j][warp] = best[j];
    }
    __syncthreads();
    if (tid < kQF) {
      unsigned long long b = s_best[tid][0];
#pragma unroll
      for (int i = 1; i < kQThreads / 32; ++i) b = s_best[tid][i] > b ? s_best[tid][i] : b;
      const int idx = static_cast<int>(0xFFFFFFFFu - static_cast<unsigned int>(b & 0xFFFFFFFFu));
      s_idx[tid] = idx;
      if (m0 + tid < M) codes[q * q_stride + m0 + tid] = idx;
    }
    __syncthreads();
    if (tid < kQF * CD) {
      const int j = tid / CD, k = tid % CD;
      s_zq[j][k] = __ldg(w.codebook + (static_cast<long long>(q) * CS + s_idx[j]) * CD + k);
    }
    __syncthreads();
    // ---- out_proj and residual update
    for (int c = tid; c < R; c += kQThreads) {
      float wv[CD];
      const float* wo = w.q_out_w + (static_cast<long long>(q) * R + c) * CD;
#pragma unroll
      for (int k = 0; k < CD; k += 4) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(wo + k));
        wv[k] = t.x; wv[k + 1] = t.y; wv[k + 2] = t.z; wv[k + 3] = t.w;
      }
      const float bo = __ldg(w.q_out_b + static_cast<long long>(q) * R + c);
#pragma unroll
      for (int j = 0; j < kQF; ++j) {
        float o = 0.f;
#pragma unroll
        for (int k = 0; k < CD; ++k) o = fmaf(wv[k], s_zq[j][k], o);
        if (valid[j]) s_r[j * R + c] -= o + bo;
      }
    }
    __syncthreads();
  }
}

int rlfq_quantize(const float* z, long long M, int T, const int64_t* lengths, int nq, int R, int CS, int CD,
                  const QuantWeights& w, int64_t* codes, long long q_stride, cudaStream_t s) {
  MT_CHECK(CD == 8 || CD == 4 || CD == 16, "rlfq_quantize: codebook_dim %d not supported (4, 8, 16)", CD);
  MT_CHECK(R > 0 && kQF * R * 4 <= 160 * 1024, "rlfq_quantize: rvq_dim %d too large", R);
  if (M == 0 || nq == 0) return MT_OK;
  const unsigned blocks = static_cast<unsigned>(cdivll(M, kQF));
  const size_t smem = static_cast<size_t>(kQF) * R * sizeof(float);
#define MT_Q(CDV)                                                                                                  \
  case CDV:                                                                                                        \
    if (smem > 48 * 1024)                                                                                          \
      MT_CUDA(cudaFuncSetAttribute(rlfq_quantize_kernel<CDV>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem))); \
    MT_CUDA(launch(rlfq_quantize_kernel<CDV>, dim3(blocks), dim3(kQThreads), smem, s, z, M, T, lengths, nq, R, CS, w, codes, q_stride)); \
    break;
  switch (CD) { MT_Q(4) MT_Q(8) MT_Q(16) }
#undef MT_Q
  count_launch();
  return MT_OK;
}

}  // namespace mt
