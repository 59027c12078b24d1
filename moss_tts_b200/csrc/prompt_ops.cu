// Prompt-side integer / elementwise kernels of MossTTSDelayProcessor, reference
// moss_tts_delay/processing_moss_tts.py:
//   loudness_normalize        :736-748  (before audio_tokenizer.batch_encode, :799-806)
//   _get_unified_codes        :588-641  (text ids + delayed audio codes -> [L, 1+n_vq])
//   _pad                      :410-431  (left padding + attention mask)
#include "common.cuh"
#include "gemm.cuh"   // count_launch

namespace mt {

// one block per waveform row: mean square -> dBFS -> clamped gain -> in-place scale
__global__ void __launch_bounds__(1024)
loudness_normalize_kernel(float* __restrict__ wav, long long ld, const int64_t* __restrict__ lengths, float target_dbfs,
                          float gain_lo, float gain_hi) {
  __shared__ float s_part[32];
  __shared__ float s_factor;
  float* row = wav + blockIdx.x * ld;
  const long long n = lengths[blockIdx.x];
  if (n <= 0) return;
  float acc = 0.f;
  for (long long i = threadIdx.x; i < n; i += blockDim.x) acc = fmaf(row[i], row[i], acc);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? s_part[threadIdx.x] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) {
      const float dbfs = 10.0f * log10f(v / static_cast<float>(n) + 1e-9f);      // fp32 tensor math in the reference
      double gain = static_cast<double>(target_dbfs) - static_cast<double>(dbfs); // python floats from here on
      gain = fmax(static_cast<double>(gain_lo), fmin(gain, static_cast<double>(gain_hi)));
      s_factor = static_cast<float>(pow(10.0, gain / 20.0));
    }
  }
  __syncthreads();
  const float f = s_factor;
  for (long long i = threadIdx.x; i < n; i += blockDim.x) row[i] *= f;
}

int loudness_normalize(float* wav, long long ld, const int64_t* lengths, int B, float target_dbfs, float gain_lo,
                       float gain_hi, cudaStream_t s) {
  if (B == 0) return MT_OK;
  loudness_normalize_kernel<<<B, 1024, 0, s>>>(wav, ld, lengths, target_dbfs, gain_lo, gain_hi);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

// blocks[j] = {first output row, rows, T (>= 0: delayed codes block, -1: pad block), first row of its codes}
__global__ void pack_unified_kernel(const int64_t* __restrict__ text, const int4* __restrict__ blocks, int n_blocks,
                                    const int64_t* __restrict__ codes, int n_vq, long long pad_code, int rows,
                                    int64_t* __restrict__ out) {
  const int C = 1 + n_vq;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < static_cast<long long>(rows) * C;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(idx / C), ch = static_cast<int>(idx % C);
    if (ch == 0) { out[idx] = text[r]; continue; }
    const int i = ch - 1;
    long long v = pad_code;
    for (int j = 0; j < n_blocks; ++j) {
      const int4 b = blocks[j];
      if (r >= b.x && r < b.x + b.y) {
        const int t = r - b.x - i;                      // delay pattern: channel i lags i rows (:516-525)
        if (b.z >= 0 && t >= 0 && t < b.z) v = codes[(static_cast<long long>(b.w) + t) * n_vq + i];
        break;
      }
    }
    out[idx] = v;
  }
}

int pack_unified_codes(const int64_t* text, const int* blocks, int n_blocks, const int64_t* codes, int n_vq,
                       long long pad_code, int rows, int64_t* out, cudaStream_t s) {
  if (rows == 0) return MT_OK;
  const long long total = static_cast<long long>(rows) * (1 + n_vq);
  int grid = static_cast<int>(cdivll(total, 256));
  if (grid > kNumSMs * 8) grid = kNumSMs * 8;
  pack_unified_kernel<<<grid, 256, 0, s>>>(text, reinterpret_cast<const int4*>(blocks), n_blocks, codes, n_vq, pad_code,
                                            rows, out);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

// out[b, Lmax - len_b + r, :] = rows_b[r, :]; the left padding is (pad_token, pad_code, ..., pad_code); mask = not padding
__global__ void pad_left_kernel(const int64_t* const* __restrict__ rows, const int* __restrict__ lens, int Lmax, int C,
                                long long pad_token, long long pad_code, int64_t* __restrict__ out,
                                uint8_t* __restrict__ mask) {
  const int b = blockIdx.y;
  const int len = lens[b], shift = Lmax - len;
  const int64_t* src = rows[b];
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < static_cast<long long>(Lmax) * C;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(idx / C), ch = static_cast<int>(idx % C);
    const bool real = r >= shift;
    out[(static_cast<long long>(b) * Lmax) * C + idx] = real ? src[static_cast<long long>(r - shift) * C + ch]
                                                             : (ch == 0 ? pad_token : pad_code);
    if (ch == 0) mask[static_cast<long long>(b) * Lmax + r] = real ? 1 : 0;
  }
}

int pad_left(const int64_t* const* rows, const int* lens, int B, int Lmax, int C, long long pad_token, long long pad_code,
             int64_t* out, uint8_t* mask, cudaStream_t s) {
  if (B == 0 || Lmax == 0) return MT_OK;
  int gx = static_cast<int>(cdivll(static_cast<long long>(Lmax) * C, 256));
  if (gx > 64) gx = 64;
  pad_left_kernel<<<dim3(gx, B), 256, 0, s>>>(rows, lens, Lmax, C, pad_token, pad_code, out, mask);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

// AudioStreamDecoder._apply_crossfade (streaming_mossttsrealtime.py:776-799): linear fade between the previous chunk's
// tail and the head of the new chunk; fade_out = torch.linspace(1, 0, n) (start + i*step below the middle, end - (n-1-i)*step
// above it, fp32).
__global__ void crossfade_kernel(const float* __restrict__ prev_tail, const float* __restrict__ wav, int n,
                                 float* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  const float step = n > 1 ? (0.f - 1.f) / static_cast<float>(n - 1) : 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float fo = n == 1 ? 1.f : (i < n / 2 ? 1.f + step * static_cast<float>(i) : 0.f - step * static_cast<float>(n - 1 - i));
    out[i] = prev_tail[i] * fo + wav[i] * (1.f - fo);
  }
}

// Streaming synthesis: collect the frames whose n_vq codes are final out of the delayed rows of gen_ids.
// rows[b, f] = generated-row index (0 = first generated row) of the gen-slot row that opens frame f of row b, or -1;
// out[q, b, f] = gen_ids[b, S0 + rows[b, f] + q, 1 + q]  (the de-delay rule, processing_moss_tts.py:528-537), 0 for -1.
__global__ void gather_frames_kernel(const int64_t* __restrict__ gen, int B, int max_len, int C, int S0,
                                     const int* __restrict__ rows, int T, int64_t* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  const int nq = C - 1;
  const long long total = static_cast<long long>(nq) * B * T;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int f = static_cast<int>(i % T), b = static_cast<int>((i / T) % B), q = static_cast<int>(i / (static_cast<long long>(T) * B));
    const int r = rows[b * T + f];
    int64_t v = 0;
    if (r >= 0 && S0 + r + q < max_len) v = gen[(static_cast<long long>(b) * max_len + S0 + r + q) * C + 1 + q];
    out[i] = v;
  }
}

}  // namespace mt

using namespace mt;

extern "C" int mt_crossfade(const float* prev_tail, const float* wav, int overlap, float* out, mt_stream s) {
  MT_CHECK(overlap >= 0 && (overlap == 0 || (prev_tail && wav && out)), "mt_crossfade: null argument");
  if (overlap == 0) return MT_OK;
  const int blocks = overlap / 256 + 1 < 4 * kNumSMs ? overlap / 256 + 1 : 4 * kNumSMs;
  MT_CUDA(launch(crossfade_kernel, dim3(blocks), dim3(256), 0, reinterpret_cast<cudaStream_t>(s), prev_tail, wav, overlap, out));
  count_launch();
  return MT_OK;
}
extern "C" int mt_stream_gather_frames(const int64_t* gen_ids, int B, int max_len, int C, int S0, const int* rows, int T,
                                       int64_t* out, mt_stream s) {
  MT_CHECK(B >= 0 && T >= 0 && C >= 2 && S0 >= 0 && max_len >= S0, "mt_stream_gather_frames: bad sizes");
  if (B * T == 0) return MT_OK;
  MT_CHECK(gen_ids && rows && out, "mt_stream_gather_frames: null argument");
  const long long total = static_cast<long long>(C - 1) * B * T;
  const int blocks = static_cast<int>(total / 256 + 1 < 4 * kNumSMs ? total / 256 + 1 : 4 * kNumSMs);
  MT_CUDA(launch(gather_frames_kernel, dim3(blocks), dim3(256), 0, reinterpret_cast<cudaStream_t>(s), gen_ids, B, max_len, C,
                 S0, rows, T, out));
  count_launch();
  return MT_OK;
}

extern "C" int mt_loudness_normalize(float* wav, long long ld, const int64_t* lengths, int B, float target_dbfs,
                                     float gain_lo, float gain_hi, mt_stream s) {
  MT_CHECK(B >= 0 && (B == 0 || (wav && lengths)), "mt_loudness_normalize: null argument");
  return loudness_normalize(wav, ld, lengths, B, target_dbfs, gain_lo, gain_hi, reinterpret_cast<cudaStream_t>(s));
}
extern "C" int mt_pack_unified_codes(const int64_t* text_codes, const int* blocks, int n_blocks, const int64_t* codes,
                                     int n_vq, long long pad_code, int rows, int64_t* out, mt_stream s) {
  MT_CHECK(rows >= 0 && n_vq >= 1 && n_blocks >= 0, "mt_pack_unified_codes: bad sizes");
  MT_CHECK(rows == 0 || (text_codes && out && (n_blocks == 0 || blocks)), "mt_pack_unified_codes: null argument");
  return pack_unified_codes(text_codes, blocks, n_blocks, codes, n_vq, pad_code, rows, out, reinterpret_cast<cudaStream_t>(s));
}
extern "C" int mt_pad_left(const int64_t* const* rows, const int* lens, int B, int Lmax, int C, long long pad_token,
                           long long pad_code, int64_t* out, uint8_t* mask, mt_stream s) {
  MT_CHECK(B >= 0 && Lmax >= 0 && C >= 1, "mt_pad_left: bad sizes");
  MT_CHECK(B == 0 || Lmax == 0 || (rows && lens && out && mask), "mt_pad_left: null argument");
  return pad_left(rows, lens, B, Lmax, C, pad_token, pad_code, out, mask, reinterpret_cast<cudaStream_t>(s));
}
