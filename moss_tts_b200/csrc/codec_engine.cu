// Host-side sequencing of MossAudioTokenizerModel._decode_frame (reference
// moss_audio_tokenizer/modeling_moss_audio_tokenizer.py:1513-1534): RVQ dequantise -> per stage
// [input_proj -> N x (LN, in_proj, RoPE, windowed attention, out_proj*scale+res, LN, linear1+GELU,
// linear2*scale+res) -> output_proj -> patch upsample].  Linear layers run on the tcgen05 GEMM with
// fp32 storage / TF32 math; LayerScale + residual and GELU are GEMM epilogues.
#include "codec_ops.cuh"
#include <cuda_fp16.h>

#include <stdlib.h>
#include <vector>

struct mt_codec {
  mt_codec_config cfg;
  mt_codec_weights w;
  std::vector<std::vector<mt_codec_layer>> layers;
  int upsample_total = 1;      // product of the patch sizes (samples per frame)
  mt_codec_quant_weights quant{};
  bool has_quant = false;
};

namespace mt {

// mt_codec_debug_set_absmax: device floats [MT_CODEC_MAX_STAGES][4] = max |value| of the fp16 operands
// {LayerNorm output, q|k|v, attention output, GELU output} of every stage, over all layers and launches since it was set
static float* g_absmax = nullptr;
// mt_codec_debug_set_attention: 0 = tcgen05 attention (codec_attn_tc5.cu, the product path), 1 = the mma.sync FlashAttention-2
// kernel (codec_attn.cu) for A/B measurements and as a cross-check of the tensor-memory kernel
static int g_attn_mode = 0;

struct StageDims { long long x, qkv, ff, y; };

// per-sample float counts of each activation buffer for T frames
static void per_sample_floats(const mt_codec_config& c, int T, long long* xs, long long* qkvs, long long* ffs,
                              long long* ys, long long* es) {
  long long tx = 0, tq = 0, tf = 0, ty = 0, te = static_cast<long long>(T) * c.code_dim;
  long long Ts = T;
  if (c.encoder) {
    // encoder: Ts shrinks; `e` holds a stage's patched input, `y` its projected output, x also the quantiser's z
    for (int s = 0; s < c.num_stages; ++s) Ts *= c.patch[s];
    tx = static_cast<long long>(T) * c.rvq_dim;
    for (int s = 0; s < c.num_stages; ++s) {
      Ts /= c.patch[s];
      tx = std::max(tx, Ts * c.d_model[s]);
      tq = std::max(tq, Ts * 3 * c.d_model[s]);
      tf = std::max(tf, Ts * c.dim_ff[s]);
      ty = std::max(ty, std::max(Ts * c.output_dim[s], Ts * 64));
      te = std::max(te, Ts * c.input_dim[s]);
    }
    *xs = tx; *qkvs = tq; *ffs = tf; *ys = ty; *es = te;
    return;
  }
  for (int s = 0; s < c.num_stages; ++s) {
    tx = std::max(tx, Ts * c.d_model[s]);
    tq = std::max(tq, Ts * 3 * c.d_model[s]);
    tf = std::max(tf, Ts * c.dim_ff[s]);
    ty = std::max(ty, Ts * c.output_dim[s]);
    te = std::max(te, Ts * c.output_dim[s]);       // unpatched input of the next stage has the same size
    Ts *= c.patch[s];
  }
  *xs = tx; *qkvs = tq; *ffs = tf; *ys = ty; *es = te;
}

static int codec_linear(const void* x, long long M, int K, const void* w, int N, void* out, long long ldo,
                        int act, const float* residual, const float* scale, cudaStream_t s, int din = MT_DTYPE_F32,
                        int dout = MT_DTYPE_F32, const float* rope = nullptr, int rope_T = 0, int rope_cols = 0) {
  LinearPlan p;
  p.dtype_in = din; p.dtype_out = dout;
  p.M = static_cast<int>(M); p.N = N; p.K = K; p.x = x; p.ldx = K; p.w = w; p.ldw = K; p.out = out; p.ldo = ldo;
  p.act = act; p.residual = residual; p.ldr = ldo; p.scale = scale;
  p.w_static = true;                       // codec weights never change while the handle is alive
  p.rope = rope; p.rope_T = rope_T; p.rope_cols = rope_cols;
  return linear(p, s);
}

}  // namespace mt

using namespace mt;

extern "C" int mt_codec_create(const mt_codec_config* cfg, const mt_codec_weights* w, mt_codec** out) {
  MT_CHECK(cfg && w && out, "mt_codec_create: null argument");
  MT_CHECK(cfg->num_stages > 0 && cfg->num_stages <= MT_CODEC_MAX_STAGES, "bad num_stages");
  mt_codec* h = new mt_codec();
  h->cfg = *cfg; h->w = *w;
  h->layers.resize(cfg->num_stages);
  for (int s = 0; s < cfg->num_stages; ++s) {
    if (cfg->d_model[s] % cfg->num_heads[s] != 0 || cfg->d_model[s] / cfg->num_heads[s] != 64) {
      delete h;
      return fail(MT_ERR_UNSUPPORTED, "codec stage %d: head_dim must be 64", s);
    }
    // K dims of the TF32 GEMMs: rows must be 16-byte aligned for TMA; a ragged last 32-float k-block is zero-filled
    const bool dims_ok = cfg->d_model[s] % 32 == 0 && cfg->dim_ff[s] % 32 == 0 && cfg->input_dim[s] % 4 == 0 &&
                         (cfg->encoder ? cfg->input_dim[s] % cfg->patch[s] == 0
                                       : (cfg->input_dim[s] % 32 == 0 && cfg->output_dim[s] % cfg->patch[s] == 0));
    if (!dims_ok) {
      delete h;
      return fail(MT_ERR_ARG, "codec stage %d: d_model/dim_ff must be multiples of 32, input_dim of %d, and the patched "
                              "width divisible by patch", s, cfg->encoder ? 4 : 32);
    }
    h->layers[s].assign(w->layers[s], w->layers[s] + cfg->num_layers[s]);
    h->upsample_total *= cfg->patch[s];
  }
  if (w->quant) { h->quant = *w->quant; h->has_quant = true; h->w.quant = &h->quant; }
  if (cfg->encoder && !h->has_quant) { delete h; return fail(MT_ERR_ARG, "mt_codec_create: an encoder handle needs quant weights"); }
  *out = h;
  return MT_OK;
}
extern "C" int mt_codec_destroy(mt_codec* h) { delete h; return MT_OK; }
extern "C" int mt_codec_debug_set_absmax(float* slots) { mt::g_absmax = slots; return MT_OK; }
extern "C" int mt_codec_debug_set_attention(int mode) { mt::g_attn_mode = mode; return MT_OK; }

extern "C" int mt_codec_workspace_bytes(const mt_codec* h, int B, int T, long long* bytes) {
  MT_CHECK(h && bytes, "null argument");
  long long xs, qs, fs, ys, es;
  per_sample_floats(h->cfg, T, &xs, &qs, &fs, &ys, &es);
  *bytes = (xs * 2 + qs + fs + ys + es) * 4 * B + 4096;
  return MT_OK;
}

extern "C" int mt_rvq_fold_tables(const float* codebook, const float* q_g, const float* q_v, const float* q_b,
                                  const float* o_g, const float* o_v, int nq, int cs, int cd, int rvq, int code_dim,
                                  float* table, mt_stream s) {
  return rvq_fold_tables(codebook, q_g, q_v, q_b, o_g, o_v, nq, cs, cd, rvq, code_dim, table,
                         reinterpret_cast<cudaStream_t>(s));
}

extern "C" int mt_rvq_dequant(const mt_codec* h, const int64_t* codes, int nq, int B, int T, int channel_major,
                              float* out, mt_stream s) {
  MT_CHECK(h && codes && out, "null argument");
  MT_CHECK(h->w.rvq_table, "mt_rvq_dequant: this handle has no folded RVQ table (encoder handle?)");
  MT_CHECK(nq >= 0 && nq <= h->cfg.num_quantizers, "nq %d exceeds the %d quantizers of the codec", nq, h->cfg.num_quantizers);
  return rvq_dequant(codes, h->w.rvq_table, h->w.rvq_out_bias, nq, B, T, h->cfg.codebook_size, h->cfg.code_dim,
                     channel_major, out, reinterpret_cast<cudaStream_t>(s));
}

// input_proj + the transformer layers of stage `st` for nb samples x Ts tokens (M = nb*Ts rows of `in`); the stage's
// activations are left in x [M, d_model].  `y` (y_floats floats) parks the stage's RoPE table while the layers run.
static int run_stage_layers(mt_codec* h, int st, const float* in, long long M, int nb, long long Ts, int chunk,
                          float* x, float* hbuf, float* qkv, float* ff, float* y, long long y_floats, cudaStream_t s,
                          __half* ring = nullptr, long long ring_layer_halves = 0, long long pos0 = 0) {
  const mt_codec_config& c = h->cfg;
  const int d = c.d_model[st], heads = c.num_heads[st], ffd = c.dim_ff[st];
  MT_CHECK(M < (1ll << 31), "codec: too many tokens in one group");
  if (h->w.input_proj[st]) MT_TRY(codec_linear(in, M, c.input_dim[st], h->w.input_proj[st], d, x, d, 0, nullptr, nullptr, s));
  else MT_CUDA(cudaMemcpyAsync(x, in, M * d * sizeof(float), cudaMemcpyDeviceToDevice, s));
  // MT_CODEC_ATTN_SIMT=1 keeps the fp32 CUDA-core kernel (+ the separate RoPE pass) as a cross-check of the TF32 one
  static const bool simt_attn = exp_env("MT_CODEC_ATTN_SIMT", 0) != 0;
  // (cos, sin) table of this stage's positions, parked in `y` (only written again after the last layer)
  const bool table_fits = static_cast<long long>(Ts) * 64 <= y_floats;
  if (!simt_attn && table_fits) MT_TRY(codec_rope_table(y, static_cast<int>(Ts), 64, c.max_period, s, pos0));
  // layer_gemm_f16: hbuf, qkv and ff hold IEEE half values (GEMM / attention operands); the residual stream x stays fp32
  const int h16 = c.layer_gemm_f16 ? 1 : 0;
  const int din = h16 ? MT_DTYPE_F16 : MT_DTYPE_F32;
  MT_CHECK(!(h16 && simt_attn), "MT_CODEC_ATTN_SIMT=1 needs a codec built with layer_gemm_f16 = 0");
  MT_CHECK(!h16 || table_fits, "codec: no room for the RoPE table of stage %d", st);
  // MT_CODEC_ATTN_TF32=1: keep q/k/v in fp32 and run the TF32 attention kernel inside an fp16-GEMM codec (A/B switch)
  static const bool attn_tf32 = exp_env("MT_CODEC_ATTN_TF32", 0) != 0;
  const bool a16 = h16 && !attn_tf32;
  float* probe = (g_absmax && h16 && a16) ? g_absmax + 4 * st : nullptr;
  for (int l = 0; l < c.num_layers[st]; ++l) {
    const mt_codec_layer& L = h->layers[st][l];
    MT_TRY(layernorm(x, L.norm1_w, L.norm1_b, hbuf, M, d, s, h16));
    if (probe) MT_TRY(absmax_f16(hbuf, M * d, probe + 0, s));
    // the tcgen05 attention takes q and k already rotated.  Since the token-major GEMMs' epilogue walks its tile row-major
    // (gemm_epi.cuh) the rotation is part of the in_proj epilogue: coalesced table reads, the fp32 accumulators rotated
    // before their one rounding to half (with the per-lane-row epilogue of round 1 the same fusion cost +80 us per launch).
    // Otherwise codec_attention_tc5 rotates q and k in place first.
    const bool tc5 = a16 && !ring && g_attn_mode == 0;
    const bool rope_in_gemm = tc5 && M > 256 && !g_force_simt && !(g_gemm_pair_mode & (2 | 16)) && d % 64 == 0;
    MT_TRY(codec_linear(hbuf, M, d, L.in_proj, 3 * d, qkv, 3 * d, 0, nullptr, nullptr, s, din, a16 ? MT_DTYPE_F16 : MT_DTYPE_F32,
                        rope_in_gemm ? y : nullptr, static_cast<int>(Ts), 2 * d));
    if (probe) MT_TRY(absmax_f16(qkv, M * 3 * d, probe + 1, s));
    if (ring) {
      // streaming push: the layer's ring cache takes the new keys / values, the new queries attend to the ring
      MT_CHECK(a16, "codec streaming needs layer_gemm_f16 = 1 (half ring caches)");
      __half* lr = ring + l * ring_layer_halves;
      MT_TRY(codec_kv_append_f16(qkv, lr, nb, static_cast<int>(Ts), heads, c.context[st], pos0, y, s));
      MT_TRY(codec_attention_ring_f16(qkv, lr, hbuf, nb, static_cast<int>(Ts), heads, c.context[st], pos0, s, y));
    } else if (tc5) {
      MT_TRY(codec_attention_tc5(qkv, hbuf, nb, static_cast<int>(Ts), heads, 64, c.context[st], chunk, s, y, rope_in_gemm));
    } else if (a16) {
      MT_TRY(codec_attention_f16(qkv, hbuf, nb, static_cast<int>(Ts), heads, 64, c.context[st], chunk, s, y));
    } else if (simt_attn || !table_fits) {
      MT_TRY(codec_rope(qkv, M, static_cast<int>(Ts), d, 64, c.max_period, s));
      if (simt_attn) MT_TRY(codec_attention(qkv, hbuf, nb, static_cast<int>(Ts), heads, 64, c.context[st], chunk, s));
      else MT_TRY(codec_attention_tc(qkv, hbuf, nb, static_cast<int>(Ts), heads, 64, c.context[st], chunk, s, nullptr, h16));
    } else {
      MT_TRY(codec_attention_tc(qkv, hbuf, nb, static_cast<int>(Ts), heads, 64, c.context[st], chunk, s, y, h16));
    }
    if (probe) MT_TRY(absmax_f16(hbuf, M * d, probe + 2, s));
    MT_TRY(codec_linear(hbuf, M, d, L.out_proj, d, x, d, 0, x, L.scale1, s, din));
    MT_TRY(layernorm(x, L.norm2_w, L.norm2_b, hbuf, M, d, s, h16));
    if (probe) MT_TRY(absmax_f16(hbuf, M * d, probe + 0, s));
    MT_TRY(codec_linear(hbuf, M, d, L.linear1, ffd, ff, ffd, 1, nullptr, nullptr, s, din, din));
    if (probe) MT_TRY(absmax_f16(ff, M * ffd, probe + 3, s));
    MT_TRY(codec_linear(ff, M, ffd, L.linear2, d, x, d, 0, x, L.scale2, s, din));
  }
  return MT_OK;
}

// decode `nb` samples starting at batch offset b0 (codes are [nq, B, T]: sample stride T, quantizer stride B*T)
struct StreamRef { __half* cache; const long long* stage_off; const long long* layer_halves; long long frames_done; };

static int decode_group(mt_codec* h, const int64_t* codes, int nq, int B, int T, int b0, int nb, int chunk_frames,
                        float* wav, float* ws, cudaStream_t s, const StreamRef* sr = nullptr) {
  const mt_codec_config& c = h->cfg;
  long long xs, qs, fs, ys, es;
  per_sample_floats(c, T, &xs, &qs, &fs, &ys, &es);
  float* x = ws;
  float* hbuf = x + xs * nb;
  float* qkv = hbuf + xs * nb;
  float* ff = qkv + qs * nb;
  float* y = ff + fs * nb;
  float* e = y + ys * nb;
  // RVQ: codes for this group are strided by B*T per quantizer -> dequantise with a pointer offset
  // (the kernel indexes codes[q*frames_total + f]; give it the full frame count and shift f by b0*T)
  {
    const long long frames_total = static_cast<long long>(B) * T;
    // launch over nb*T frames with an offset view: codes + b0*T, quantizer stride stays frames_total
    MT_TRY(rvq_dequant_strided(codes + static_cast<long long>(b0) * T, h->w.rvq_table, h->w.rvq_out_bias, nq,
                               frames_total, static_cast<long long>(nb) * T, c.codebook_size, c.code_dim, e, s));
  }
  long long Ts = T;
  const float* in = e;
  for (int st = 0; st < c.num_stages; ++st) {
    const int d = c.d_model[st];
    const long long M = Ts * nb;
    if (sr)
      MT_TRY(run_stage_layers(h, st, in, M, nb, Ts, 0, x, hbuf, qkv, ff, y, ys * nb, s, sr->cache + sr->stage_off[st],
                              sr->layer_halves[st], sr->frames_done * (Ts / T)));
    else
      MT_TRY(run_stage_layers(h, st, in, M, nb, Ts, chunk_frames > 0 ? chunk_frames * static_cast<int>(Ts / T) : 0,
                              x, hbuf, qkv, ff, y, ys * nb, s));
    const bool last = st + 1 == c.num_stages;
    const int od = c.output_dim[st], p = c.patch[st];
    const bool identity_patch = od == p;           // one output channel: the reshape is the identity
    float* dst = last && identity_patch ? wav + static_cast<long long>(b0) * T * h->upsample_total : y;
    const float* yo;
    if (h->w.output_proj[st]) { MT_TRY(codec_linear(x, M, d, h->w.output_proj[st], od, dst, od, 0, nullptr, nullptr, s)); yo = dst; }
    else yo = x;
    if (last) {
      if (!identity_patch) MT_TRY(unpatch(yo, wav + static_cast<long long>(b0) * T * h->upsample_total, M, od, p, s));
      else if (yo != dst) MT_CUDA(cudaMemcpyAsync(dst, yo, M * od * sizeof(float), cudaMemcpyDeviceToDevice, s));
    } else {
      if (p == 1) { MT_CUDA(cudaMemcpyAsync(e, yo, M * od * sizeof(float), cudaMemcpyDeviceToDevice, s)); }
      else MT_TRY(unpatch(yo, e, M, od, p, s));
      in = e;
    }
    Ts *= p;
  }
  return MT_OK;
}

extern "C" int mt_codec_decode(mt_codec* h, const int64_t* codes, int nq, int B, int T, int chunk_frames, float* wav,
                               void* workspace, long long workspace_bytes, mt_stream stream) {
  MT_CHECK(h && codes && wav && workspace, "mt_codec_decode: null argument");
  MT_CHECK(!h->cfg.encoder, "mt_codec_decode: this handle is an encoder");
  MT_CHECK(nq > 0 && nq <= h->cfg.num_quantizers, "nq %d out of range (codec has %d quantizers)", nq, h->cfg.num_quantizers);
  MT_CHECK(B > 0 && T > 0, "mt_codec_decode: empty input B=%d T=%d", B, T);
  long long per = 0;
  MT_TRY(mt_codec_workspace_bytes(h, 1, T, &per));
  const long long group = (workspace_bytes - 4096) / (per - 4096);
  MT_CHECK(group >= 1, "codec workspace too small: need %lld bytes per sample, have %lld", per, workspace_bytes);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  for (int b0 = 0; b0 < B; b0 += static_cast<int>(group)) {
    const int nb = static_cast<int>(std::min<long long>(group, B - b0));
    MT_TRY(decode_group(h, codes, nq, B, T, b0, nb, chunk_frames, wav, reinterpret_cast<float*>(workspace), s));
  }
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// encode side
// ---------------------------------------------------------------------------------------------
// hidden [M, code_dim] token-major (M = nb*T rows) -> codes[q*q_stride + m]; z: scratch of M*rvq_dim floats
static int quantize_rows(mt_codec* h, const float* hidden, long long M, int T, const int64_t* lengths, int nq,
                         int64_t* codes, long long q_stride, float* z, cudaStream_t s) {
  const mt_codec_config& c = h->cfg;
  MT_CHECK(M < (1ll << 31), "quantise: too many frames in one group");
  const float* zin = hidden;
  if (h->quant.in_w) {
    // fp32 FMA arithmetic on purpose (CUDA-core GEMM): the code index is an argmax over distances of this projection;
    // TF32 inputs would flip near-ties that the reference's fp32 conv1d resolves
    LinearPlan p;
    p.dtype_in = MT_DTYPE_F32; p.dtype_out = MT_DTYPE_F32;
    p.M = static_cast<int>(M); p.N = c.rvq_dim; p.K = c.code_dim;
    p.x = hidden; p.ldx = c.code_dim; p.w = h->quant.in_w; p.ldw = c.code_dim; p.out = z; p.ldo = c.rvq_dim;
    p.bias = h->quant.in_b; p.impl = MT_GEMM_SIMT;
    MT_TRY(linear(p, s));
    zin = z;
  } else {
    MT_CHECK(c.rvq_dim == c.code_dim, "quantise: no input_proj but rvq_dim != code_dim");
  }
  return rlfq_quantize(zin, M, T, lengths, nq, c.rvq_dim, c.codebook_size, c.codebook_dim, h->quant, codes, q_stride, s);
}

extern "C" int mt_rlfq_quantize(mt_codec* h, const float* hidden, int B, int T, const int64_t* code_lengths, int nq,
                                int64_t* codes, void* workspace, long long workspace_bytes, mt_stream stream) {
  MT_CHECK(h && hidden && codes && workspace, "mt_rlfq_quantize: null argument");
  MT_CHECK(h->has_quant, "mt_rlfq_quantize: this handle has no quantiser weights");
  MT_CHECK(nq > 0 && nq <= h->cfg.num_quantizers, "nq %d out of range (codec has %d quantizers)", nq, h->cfg.num_quantizers);
  MT_CHECK(B > 0 && T > 0, "mt_rlfq_quantize: empty input B=%d T=%d", B, T);
  const long long M = static_cast<long long>(B) * T;
  MT_CHECK(workspace_bytes >= M * h->cfg.rvq_dim * 4, "mt_rlfq_quantize: workspace too small (%lld bytes needed)",
           M * h->cfg.rvq_dim * 4);
  return quantize_rows(h, hidden, M, T, code_lengths, nq, codes, M, reinterpret_cast<float*>(workspace),
                       reinterpret_cast<cudaStream_t>(stream));
}

// encode `nb` samples starting at batch offset b0: wav [B, T*samples_per_frame] -> codes [nq, B, T]
static int encode_group(mt_codec* h, const float* wav, int B, int T, int b0, int nb, const int64_t* lengths, int nq,
                        int chunk_frames, int64_t* codes, float* hidden_out, float* ws, cudaStream_t s,
                        const StreamRef* sr = nullptr) {
  const mt_codec_config& c = h->cfg;
  long long xs, qs, fs, ys, es;
  per_sample_floats(c, T, &xs, &qs, &fs, &ys, &es);
  float* x = ws;
  float* hbuf = x + xs * nb;
  float* qkv = hbuf + xs * nb;
  float* ff = qkv + qs * nb;
  float* y = ff + fs * nb;
  float* e = y + ys * nb;
  long long Ts = static_cast<long long>(T) * h->upsample_total;
  const float* in = wav + static_cast<long long>(b0) * Ts;     // [nb*Ts, 1]
  int width = 1;
  for (int st = 0; st < c.num_stages; ++st) {
    const int p = c.patch[st], d = c.d_model[st], od = c.output_dim[st];
    Ts /= p;
    const long long M = Ts * nb;
    MT_CHECK(width * p == c.input_dim[st], "encoder stage %d: input_dim %d != %d channels x patch %d", st, c.input_dim[st], width, p);
    const float* stage_in = in;
    if (width != 1 && p != 1) { MT_TRY(patchify(in, e, M, width, p, s)); stage_in = e; }   // one channel: the reshape is the identity
    if (sr)
      MT_TRY(run_stage_layers(h, st, stage_in, M, nb, Ts, 0, x, hbuf, qkv, ff, y, ys * nb, s, sr->cache + sr->stage_off[st],
                              sr->layer_halves[st], sr->frames_done * (Ts / T)));
    else
      MT_TRY(run_stage_layers(h, st, stage_in, M, nb, Ts, chunk_frames > 0 ? chunk_frames * static_cast<int>(Ts / T) : 0,
                              x, hbuf, qkv, ff, y, ys * nb, s));
    const bool last = st + 1 == c.num_stages;
    float* direct = last && hidden_out ? hidden_out + static_cast<long long>(b0) * T * od : nullptr;
    if (h->w.output_proj[st]) {
      float* dst = direct ? direct : y;
      MT_TRY(codec_linear(x, M, d, h->w.output_proj[st], od, dst, od, 0, nullptr, nullptr, s));
      in = dst;
    } else {
      if (direct) MT_CUDA(cudaMemcpyAsync(direct, x, M * od * sizeof(float), cudaMemcpyDeviceToDevice, s));
      in = x;
    }
    width = od;
  }
  MT_CHECK(width == c.code_dim && Ts == T, "encoder: last stage must produce code_dim channels at the frame rate");
  return quantize_rows(h, in, static_cast<long long>(nb) * T, T, lengths ? lengths + b0 : nullptr, nq,
                       codes + static_cast<long long>(b0) * T, static_cast<long long>(B) * T, hbuf, s);
}

extern "C" int mt_codec_encode(mt_codec* h, const float* wav, int B, int T, const int64_t* code_lengths, int nq,
                               int chunk_frames, int64_t* codes, float* hidden, void* workspace, long long workspace_bytes,
                               mt_stream stream) {
  MT_CHECK(h && wav && codes && workspace, "mt_codec_encode: null argument");
  MT_CHECK(h->cfg.encoder && h->has_quant, "mt_codec_encode: this handle is not an encoder");
  MT_CHECK(nq > 0 && nq <= h->cfg.num_quantizers, "nq %d out of range (codec has %d quantizers)", nq, h->cfg.num_quantizers);
  MT_CHECK(B > 0 && T > 0, "mt_codec_encode: empty input B=%d T=%d", B, T);
  long long per = 0;
  MT_TRY(mt_codec_workspace_bytes(h, 1, T, &per));
  const long long group = (workspace_bytes - 4096) / (per - 4096);
  MT_CHECK(group >= 1, "codec workspace too small: need %lld bytes per sample, have %lld", per, workspace_bytes);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  for (int b0 = 0; b0 < B; b0 += static_cast<int>(group)) {
    const int nb = static_cast<int>(std::min<long long>(group, B - b0));
    MT_TRY(encode_group(h, wav, B, T, b0, nb, code_lengths, nq, chunk_frames, codes, hidden,
                        reinterpret_cast<float*>(workspace), s));
  }
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// incremental streaming (model.streaming(batch_size) + _decode_frame / _encode_frame per push)
// ---------------------------------------------------------------------------------------------
struct mt_codec_stream {
  mt_codec* h;
  int B;
  __half* cache;
  long long stage_off[MT_CODEC_MAX_STAGES], layer_halves[MT_CODEC_MAX_STAGES];
  long long frames_done;
};

static long long stream_layout(const mt_codec* h, int B, long long* stage_off, long long* layer_halves) {
  const mt_codec_config& c = h->cfg;
  long long off = 0;
  for (int s = 0; s < c.num_stages; ++s) {
    const long long lh = static_cast<long long>(B) * c.context[s] * 2 * c.d_model[s];
    if (stage_off) { stage_off[s] = off; layer_halves[s] = lh; }
    off += lh * c.num_layers[s];
  }
  return off;
}

extern "C" int mt_codec_stream_bytes(const mt_codec* h, int B, long long* bytes) {
  MT_CHECK(h && bytes && B > 0, "mt_codec_stream_bytes: bad argument");
  *bytes = stream_layout(h, B, nullptr, nullptr) * 2 + 256;
  return MT_OK;
}

extern "C" int mt_codec_stream_begin(mt_codec* h, int B, void* cache, long long cache_bytes, mt_codec_stream** out) {
  MT_CHECK(h && cache && out && B > 0, "mt_codec_stream_begin: bad argument");
  if (!h->cfg.layer_gemm_f16) return fail(MT_ERR_UNSUPPORTED, "codec streaming needs a codec built with layer_gemm_f16 = 1");
  long long need = 0;
  MT_TRY(mt_codec_stream_bytes(h, B, &need));
  MT_CHECK(cache_bytes >= need, "mt_codec_stream_begin: cache of %lld bytes, %lld needed", cache_bytes, need);
  mt_codec_stream* st = new mt_codec_stream();
  st->h = h; st->B = B; st->cache = reinterpret_cast<__half*>(cache); st->frames_done = 0;
  stream_layout(h, B, st->stage_off, st->layer_halves);
  *out = st;
  return MT_OK;
}

// every stage's ring holds `context` tokens: a push of T frames is T * (stage rate / frame rate) tokens there
static int stream_push_fits(const mt_codec* h, int T, const char* who) {
  const mt_codec_config& c = h->cfg;
  long long Ts = c.encoder ? static_cast<long long>(T) * h->upsample_total : T;
  for (int st = 0; st < c.num_stages; ++st) {
    if (c.encoder) Ts /= c.patch[st];
    MT_CHECK(Ts <= c.context[st], "%s: %d frames are %lld tokens at stage %d, more than its ring of %d tokens of context",
             who, T, Ts, st, c.context[st]);
    if (!c.encoder) Ts *= c.patch[st];
  }
  return MT_OK;
}

extern "C" int mt_codec_stream_push(mt_codec_stream* st, const int64_t* codes, int nq, int T, float* wav, void* workspace,
                                    long long workspace_bytes, mt_stream stream) {
  MT_CHECK(st && codes && wav && workspace, "mt_codec_stream_push: null argument");
  mt_codec* h = st->h;
  MT_CHECK(!h->cfg.encoder, "mt_codec_stream_push: this stream belongs to an encoder (mt_codec_stream_push_encode)");
  MT_CHECK(nq > 0 && nq <= h->cfg.num_quantizers, "nq %d out of range (codec has %d quantizers)", nq, h->cfg.num_quantizers);
  MT_CHECK(T > 0, "mt_codec_stream_push: empty push");
  MT_TRY(stream_push_fits(h, T, "mt_codec_stream_push"));
  long long need = 0;
  MT_TRY(mt_codec_workspace_bytes(h, st->B, T, &need));
  MT_CHECK(workspace_bytes >= need, "mt_codec_stream_push: workspace of %lld bytes, %lld needed", workspace_bytes, need);
  StreamRef sr{st->cache, st->stage_off, st->layer_halves, st->frames_done};
  MT_TRY(decode_group(h, codes, nq, st->B, T, 0, st->B, 0, wav, reinterpret_cast<float*>(workspace),
                      reinterpret_cast<cudaStream_t>(stream), &sr));
  st->frames_done += T;
  return MT_OK;
}

extern "C" int mt_codec_stream_push_encode(mt_codec_stream* st, const float* wav, int T, const int64_t* code_lengths, int nq,
                                           int64_t* codes, float* hidden, void* workspace, long long workspace_bytes,
                                           mt_stream stream) {
  MT_CHECK(st && wav && codes && workspace, "mt_codec_stream_push_encode: null argument");
  mt_codec* h = st->h;
  MT_CHECK(h->cfg.encoder && h->has_quant, "mt_codec_stream_push_encode: this stream belongs to a decoder (mt_codec_stream_push)");
  MT_CHECK(nq > 0 && nq <= h->cfg.num_quantizers, "nq %d out of range (codec has %d quantizers)", nq, h->cfg.num_quantizers);
  MT_CHECK(T > 0, "mt_codec_stream_push_encode: empty push");
  MT_TRY(stream_push_fits(h, T, "mt_codec_stream_push_encode"));
  long long need = 0;
  MT_TRY(mt_codec_workspace_bytes(h, st->B, T, &need));
  MT_CHECK(workspace_bytes >= need, "mt_codec_stream_push_encode: workspace of %lld bytes, %lld needed", workspace_bytes, need);
  StreamRef sr{st->cache, st->stage_off, st->layer_halves, st->frames_done};
  MT_TRY(encode_group(h, wav, st->B, T, 0, st->B, code_lengths, nq, 0, codes, hidden, reinterpret_cast<float*>(workspace),
                      reinterpret_cast<cudaStream_t>(stream), &sr));
  st->frames_done += T;
  return MT_OK;
}

extern "C" int mt_codec_stream_end(mt_codec_stream* st) { delete st; return MT_OK; }
