// Host-side sequencing of MossAudioTokenizerModel._decode_frame (reference
// moss_audio_tokenizer/modeling_moss_audio_tokenizer.py:1513-1534): RVQ dequantise -> per stage
// [input_proj -> N x (LN, in_proj, RoPE, windowed attention, out_proj*scale+res, LN, linear1+GELU,
// linear2*scale+res) -> output_proj -> patch upsample].  Linear layers run on the tcgen05 GEMM with
// fp32 storage / TF32 math; LayerScale + residual and GELU are GEMM epilogues.
#include "codec_ops.cuh"

#include <stdlib.h>
#include <vector>

struct mt_codec {
  mt_codec_config cfg;
  mt_codec_weights w;
  std::vector<std::vector<mt_codec_layer>> layers;
  int upsample_total = 1;
};

namespace mt {

struct StageDims { long long x, qkv, ff, y; };

// per-sample float counts of each activation buffer for T frames
static void per_sample_floats(const mt_codec_config& c, int T, long long* xs, long long* qkvs, long long* ffs,
                              long long* ys, long long* es) {
  long long tx = 0, tq = 0, tf = 0, ty = 0, te = static_cast<long long>(T) * c.code_dim;
  long long Ts = T;
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

static int codec_linear(const float* x, long long M, int K, const float* w, int N, float* out, long long ldo,
                        int act, const float* residual, const float* scale, cudaStream_t s) {
  LinearPlan p;
  p.dtype_in = MT_DTYPE_F32; p.dtype_out = MT_DTYPE_F32;
  p.M = static_cast<int>(M); p.N = N; p.K = K; p.x = x; p.ldx = K; p.w = w; p.ldw = K; p.out = out; p.ldo = ldo;
  p.act = act; p.residual = residual; p.ldr = ldo; p.scale = scale;
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
    if (cfg->output_dim[s] % cfg->patch[s] != 0 || cfg->d_model[s] % 32 != 0 || cfg->input_dim[s] % 32 != 0 ||
        cfg->dim_ff[s] % 32 != 0) {
      delete h;
      return fail(MT_ERR_ARG, "codec stage %d: dims must be multiples of 32 and output_dim divisible by patch", s);
    }
    h->layers[s].assign(w->layers[s], w->layers[s] + cfg->num_layers[s]);
    h->upsample_total *= cfg->patch[s];
  }
  *out = h;
  return MT_OK;
}
extern "C" int mt_codec_destroy(mt_codec* h) { delete h; return MT_OK; }

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
  MT_CHECK(nq >= 0 && nq <= h->cfg.num_quantizers, "nq %d exceeds the %d quantizers of the codec", nq, h->cfg.num_quantizers);
  return rvq_dequant(codes, h->w.rvq_table, h->w.rvq_out_bias, nq, B, T, h->cfg.codebook_size, h->cfg.code_dim,
                     channel_major, out, reinterpret_cast<cudaStream_t>(s));
}

// decode `nb` samples starting at batch offset b0 (codes are [nq, B, T]: sample stride T, quantizer stride B*T)
static int decode_group(mt_codec* h, const int64_t* codes, int nq, int B, int T, int b0, int nb, int chunk_frames,
                        float* wav, float* ws, cudaStream_t s) {
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
    const int d = c.d_model[st], heads = c.num_heads[st], ffd = c.dim_ff[st];
    const long long M = Ts * nb;
    MT_CHECK(M < (1ll << 31), "codec: too many tokens in one group");
    if (h->w.input_proj[st]) MT_TRY(codec_linear(in, M, c.input_dim[st], h->w.input_proj[st], d, x, d, 0, nullptr, nullptr, s));
    else MT_CUDA(cudaMemcpyAsync(x, in, M * d * sizeof(float), cudaMemcpyDeviceToDevice, s));
    const int chunk = chunk_frames > 0 ? chunk_frames * static_cast<int>(Ts / T) : 0;
    // MT_CODEC_ATTN_SIMT=1 keeps the fp32 CUDA-core kernel (+ the separate RoPE pass) as a cross-check of the TF32 one
    static const bool simt_attn = [] { const char* e = getenv("MT_CODEC_ATTN_SIMT"); return e && e[0] == '1'; }();
    // (cos, sin) table of this stage's positions, parked in `y` (only written again after the last layer)
    const bool table_fits = static_cast<long long>(Ts) * 64 <= ys * nb;
    if (!simt_attn && table_fits) MT_TRY(codec_rope_table(y, static_cast<int>(Ts), 64, c.max_period, s));
    for (int l = 0; l < c.num_layers[st]; ++l) {
      const mt_codec_layer& L = h->layers[st][l];
      MT_TRY(layernorm(x, L.norm1_w, L.norm1_b, hbuf, M, d, s));
      MT_TRY(codec_linear(hbuf, M, d, L.in_proj, 3 * d, qkv, 3 * d, 0, nullptr, nullptr, s));
      if (simt_attn || !table_fits) {
        MT_TRY(codec_rope(qkv, M, static_cast<int>(Ts), d, 64, c.max_period, s));
        if (simt_attn) MT_TRY(codec_attention(qkv, hbuf, nb, static_cast<int>(Ts), heads, 64, c.context[st], chunk, s));
        else MT_TRY(codec_attention_tc(qkv, hbuf, nb, static_cast<int>(Ts), heads, 64, c.context[st], chunk, s));
      } else {
        MT_TRY(codec_attention_tc(qkv, hbuf, nb, static_cast<int>(Ts), heads, 64, c.context[st], chunk, s, y));
      }
      MT_TRY(codec_linear(hbuf, M, d, L.out_proj, d, x, d, 0, x, L.scale1, s));
      MT_TRY(layernorm(x, L.norm2_w, L.norm2_b, hbuf, M, d, s));
      MT_TRY(codec_linear(hbuf, M, d, L.linear1, ffd, ff, ffd, 1, nullptr, nullptr, s));
      MT_TRY(codec_linear(ff, M, ffd, L.linear2, d, x, d, 0, x, L.scale2, s));
    }
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
