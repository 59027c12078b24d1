// Non-GEMM kernels of the MossTTSDelay step (lm_ops.cu) and the sampler (sampling.cu).
#pragma once
#include "common.cuh"
#include "gemm.cuh"

namespace mt {

// ids rows are `row_stride` int64 wide (0 = 1 + n_vq); only channel 0 and the first n_vq audio channels are summed
int embed_sum(const int64_t* ids, const bf16* emb_text, const bf16* emb_ext, bf16* out, int M, int H, int n_vq,
              int audio_rows, int vocab, cudaStream_t s, int row_stride = 0);
int add_rmsnorm(const bf16* residual, const bf16* delta, const float* partials, int nsplit, long long split_stride,
                bf16* x_out, const bf16* gain, bf16* normed, int M, int H, float eps, cudaStream_t s);
int qkv_post(const bf16* qkv, const float* partials, int nsplit, long long split_stride, const bf16* q_gain,
             const bf16* k_gain, const bf16* rope_cos, const bf16* rope_sin, bf16* q_out, bf16* k_cache,
             bf16* v_cache, int B, int T, int n_heads, int n_kv, int D, int past_len, const int* past_len_dev,
             int max_seq, float eps, cudaStream_t s);
int attention(const bf16* q, const bf16* k_cache, const bf16* v_cache, const uint8_t* mask, long long mask_ld,
              bf16* out, int B, int T, int n_heads, int n_kv, int D, int past_len, const int* past_len_dev,
              int max_keys, int max_seq, cudaStream_t s);
// tensor-core two-pass form of `attention` for any T and any number of keys (lm_attn_prefill.cu); `attention`
// dispatches to it when the head layout is supported
bool attention_tc_supported(int n_heads, int n_kv, int D);
int attention_tc(const bf16* q, const bf16* k_cache, const bf16* v_cache, const uint8_t* mask, long long mask_ld, bf16* out,
                 int B, int T, int n_heads, int n_kv, int D, int past_len, const int* past_len_dev, int max_seq,
                 cudaStream_t s);
// decode (T = 1): qkv_post + attention fused, one warp per (row, kv head), mma.sync tensor cores (lm_attn.cu)
bool attn_decode_fits(int B, int n_heads, int n_kv, int D, int max_keys);
int attn_decode(const bf16* qkv, const float* partials, int nsplit, long long split_stride, const bf16* q_gain,
                const bf16* k_gain, const bf16* rope_cos, const bf16* rope_sin, bf16* k_cache, bf16* v_cache,
                const uint8_t* mask, long long mask_ld, bf16* out, int B, int n_heads, int n_kv, int D,
                const int* pos_dev, int pos_host, int max_keys, int max_seq, float eps, cudaStream_t s, int early_pos = 0);
// early_pos = 1: the kernel may read the position, the rotary row and the K rows [0, pos) BEFORE its programmatic dependency
// wait.  Safe only when their writers have certainly finished by the time this kernel's blocks run, which the launch order
// alone does not give (a chain of waiting kernels can be resident).  The engines' rule: this attention follows a decode GEMM
// G1 (the qkv projection); a decode GEMM CTA owns > half an SM's shared memory, so once all c1 CTAs of G1 are resident any
// earlier GEMM X has at most 148 - c1 CTAs left; if X launched more than that, one of its CTAs has exited, i.e. X passed
// its own dependency wait, i.e. everything launched before X has completed.  early_pos is set when such an X (the gate|up
// GEMM of the previous layer / previous depth step) lies between the writers and this kernel: c(X) + c(G1) > 148.
bool attn_early_reads_safe(int ctas_qkv, int ctas_between);
int silu_mul(const bf16* gu, const float* partials, int nsplit, long long split_stride, bf16* act, int M, int I,
             cudaStream_t s);
int fill_audio_pad(bf16* logits, long long rows, int head_size, cudaStream_t s);
int gather_last(const bf16* src, bf16* dst, int B, int T, int H, cudaStream_t s);
// MossTTSLocal: MossTTSRMSNorm (all ops in bf16) and the per-channel embedding lookup of a sampled token
int moss_rmsnorm(const bf16* x, const float* partials, int nsplit, long long split_stride, const bf16* gain, bf16* out,
                 int M, int H, float eps, cudaStream_t s);
int embed_rows(const int64_t* ids, int stride, int col, const bf16* table, long long rows, bf16* out, int B, int H,
               cudaStream_t s);

}  // namespace mt
