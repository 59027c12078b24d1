// MOSS-Audio-Tokenizer decode-side kernels (codec_ops.cu).
#pragma once
#include "common.cuh"
#include "gemm.cuh"

namespace mt {
int rvq_fold_tables(const float* codebook, const float* q_g, const float* q_v, const float* q_b, const float* o_g,
                    const float* o_v, int nq, int cs, int cd, int rvq, int code_dim, float* table, cudaStream_t s);
int rvq_dequant(const int64_t* codes, const float* table, const float* bias, int nq, int B, int T, int cs, int code_dim,
                int channel_major, float* out, cudaStream_t s);
int rvq_dequant_strided(const int64_t* codes, const float* table, const float* bias, int nq, long long q_stride,
                        long long frames, int cs, int code_dim, float* out, cudaStream_t s);
// out_f16 = 1: `out` receives IEEE half values (the fp16 operand of the next GEMM) instead of fp32
int layernorm(const float* x, const float* w, const float* b, float* out, long long M, int d, cudaStream_t s, int out_f16 = 0);
int codec_rope(float* qkv, long long M, int T, int d, int hd, float max_period, cudaStream_t s);
int codec_attention(const float* qkv, float* out, int B, int T, int heads, int hd, int context, int chunk,
                    cudaStream_t s);
// tensor-core (TF32 mma) version of codec_attention: the product path (codec_attn.cu)
// (cos, sin) of every (position, pair) as codec_rope computes them: table[t * hd/2 + j] = (cos, sin)(t * freq_j)
// pos0: absolute position of table row 0 (streaming pushes)
int codec_rope_table(float* table, int T, int hd, float max_period, cudaStream_t s, long long pos0 = 0);
// rope_table != nullptr: q and k are rotated while they are staged (no separate codec_rope pass over qkv)
int codec_attention_tc(const float* qkv, float* out, int B, int T, int heads, int hd, int context, int chunk,
                       cudaStream_t s, const float* rope_table = nullptr, int out_f16 = 0);
// fp16 flavour (codec_attn.cu): qkv [B*T, 3*heads*64] and out [B*T, heads*64] are IEEE half; RoPE table required
int codec_attention_f16(const void* qkv, void* out, int B, int T, int heads, int hd, int context, int chunk,
                        cudaStream_t s, const float* rope_table);
// tcgen05 / TMEM / TMA flavour (codec_attn_tc5.cu): q and k of `qkv` are rotated IN PLACE first, then S = Q.K^T and P.V run
// on the 5th-generation tensor cores.  The product path of the one-pass decode / encode; codec_attention_f16 stays the
// kernel of the incremental ring-cache streaming path and the A/B cross-check (mt_codec_debug_set_attention(1)).
int codec_attention_tc5(void* qkv, void* out, int B, int T, int heads, int hd, int context, int chunk, cudaStream_t s,
                        const float* rope_table, bool pre_rotated = false);
// Streaming decode (RingKVCache :413-480, MHAState :488-499).  ring: one layer's cache [B][cap][2][heads*64] half, slot =
// position % cap, K stored AFTER RoPE (as the reference caches it).  kv_append rotates the new tokens' keys (table row t =
// position pos0 + t) and writes k, v into their slots; the attention then reads q from qkv (rotated on the way in) and
// K / V from the ring: visible keys of every query of the push are [max(pos0 + T - cap, 0), its own position].
int codec_kv_append_f16(const void* qkv, void* ring, int B, int T, int heads, int cap, long long pos0, const float* rope_table,
                        cudaStream_t s);
int codec_attention_ring_f16(const void* qkv, const void* ring, void* out, int B, int T, int heads, int cap, long long pos0,
                             cudaStream_t s, const float* rope_table);
int unpatch(const float* y, float* x, long long tokens, int dh, int h, cudaStream_t s);
// debug probe: *slot = max(*slot, max |x[i]|) over n IEEE-half values (n even); NaN if any element is NaN
int absmax_f16(const void* x, long long n, float* slot, cudaStream_t s);
// encode side (codec_enc.cu)
using QuantWeights = mt_codec_quant_weights;
int patchify(const float* x, float* y, long long tokens_out, int d, int h, cudaStream_t s);
int rlfq_quantize(const float* z, long long M, int T, const int64_t* lengths, int nq, int R, int CS, int CD,
                  const QuantWeights& w, int64_t* codes, long long q_stride, cudaStream_t s);
}  // namespace mt
