// MossTTSLocal engine: a Qwen3 "global" backbone step followed by the depth loop that
// emits the 1+n_vq tokens of one frame sequentially through a small causal transformer without rotary embedding.
// Reference: moss_tts_local/modeling_moss_tts.py  CustomMixin._sample :316-477 (frame loop :379-450),
//            MosiTTSModel :495-569, MossTTSLocalTransformer :178-292, module wiring :573-611.
//
// The global backbone is an mt_lm (lm_engine.cu) created without heads; the depth transformer reuses the same
// kernels (tcgen05 decode GEMM with split-K partials, fused norm/KV-append/attention with rope disabled) over its
// own hidden size and a KV cache of 1+n_vq positions per row.  The reference recomputes the causal prefix at every
// depth step (use_cache=False, :205); a causal stack makes that equal to keeping the prefix K/V, which is what
// happens here.  One frame = 1 backbone step + (1+n_inf) depth steps, replayed from a CUDA graph.
#include "lm_engine.cuh"

#include <algorithm>
#include <vector>

struct mt_local {
  mt_local_config cfg;
  mt_local_weights w;
  mt_lm* g = nullptr;                 // global backbone (owns the tensor maps of its layers, the split-K buffer)
  int C = 0;                          // channels = 1 + n_vq
  std::vector<const void*> w_qkv, w_o, w_gu, w_down, ln_in, ln_post, q_norm, k_norm;   // local layers
  std::vector<const void*> out_gu, out_down, head_norm, heads;                         // per channel
  std::vector<CUtensorMap> tm_qkv, tm_o, tm_gu, tm_down, tm_out_gu, tm_out_down, tm_heads;
  CUtensorMap tm_in_gu, tm_in_down;
  // workspace views
  mt::bf16 *x, *xn, *qkv, *attn, *gu, *act, *delta, *hn, *emb, *logits, *k_cache, *v_cache;
  void* state_buf;
};

namespace mt {

namespace {

struct Carver {
  uint8_t* base; size_t off = 0;
  explicit Carver(void* b) : base(reinterpret_cast<uint8_t*>(b)) {}
  template <typename T> T* take(size_t n) {
    off = (off + 255) & ~size_t(255);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += n * sizeof(T);
    return p;
  }
};

int check_local_cfg(const mt_local_config* c) {
  MT_CHECK(c, "null config");
  MT_CHECK(c->local_hidden_size % 64 == 0 && c->local_ffn_hidden_size % 64 == 0 && c->additional_mlp_ffn_hidden_size % 64 == 0,
           "local hidden / ffn / adapter sizes must be multiples of 64");
  MT_CHECK(c->local_num_layers >= 1, "local_num_layers must be >= 1");
  return MT_OK;
}

// local workspace follows the backbone's inside the caller's buffer
size_t carve_local(mt_local* h, const mt_local_config& c, void* ws) {
  Carver cv(ws);
  const size_t B = c.lm.max_batch, H = c.lm.hidden_size, LH = c.local_hidden_size, LF = c.local_ffn_hidden_size;
  const size_t AF = c.additional_mlp_ffn_hidden_size, D = c.lm.head_dim, C = 1 + c.lm.n_vq;
  const size_t qs = static_cast<size_t>(c.lm.num_heads) * D, kvs = static_cast<size_t>(c.lm.num_kv_heads) * D;
  const size_t wide = std::max(LH, H), ff = std::max(LF, AF);
  const size_t V = std::max<size_t>(c.lm.vocab_size, c.lm.audio_head_size);
  auto x = cv.take<bf16>(B * LH);
  auto xn = cv.take<bf16>(B * wide);
  auto qkv = cv.take<bf16>(B * (qs + 2 * kvs));
  auto attn = cv.take<bf16>(B * qs);
  auto gu = cv.take<bf16>(B * 2 * ff);
  auto act = cv.take<bf16>(B * ff);
  auto delta = cv.take<bf16>(B * wide);
  auto hn = cv.take<bf16>(B * H);
  auto emb = cv.take<bf16>(B * H);
  auto logits = cv.take<bf16>(B * V);
  const size_t kv_half = static_cast<size_t>(c.local_num_layers) * B * c.lm.num_kv_heads * C * D;
  auto kc = cv.take<bf16>(kv_half);
  auto vc = cv.take<bf16>(kv_half);
  auto sb = cv.take<uint8_t>(local_state_bytes(c.lm.max_batch, static_cast<int>(C), c.lm.max_seq, c.lm.audio_head_size));
  if (h) {
    h->x = x; h->xn = xn; h->qkv = qkv; h->attn = attn; h->gu = gu; h->act = act; h->delta = delta; h->hn = hn;
    h->emb = emb; h->logits = logits; h->k_cache = kc; h->v_cache = vc; h->state_buf = sb;
  }
  return cv.off + 256;
}

}  // namespace

// swiglu adapter (MossTTSMLP :47-95): out partials/bf16 of down( silu(gate(x)) * up(x) ); gate|up rows are stacked
static int local_mlp(mt_local* h, const bf16* x, int B, int K, int F, int N, const void* w_gu, const CUtensorMap* tm_gu,
                     const void* w_down, const CUtensorMap* tm_down, bf16* out, int* sp, cudaStream_t s) {
  int s1 = 0;
  MT_TRY(lm_linear(h->g, x, B, K, w_gu, tm_gu, 2 * F, h->gu, &s1, s));
  MT_TRY(silu_mul(h->gu, s1 ? h->g->splitk : nullptr, s1, static_cast<long long>(B) * 2 * F, h->act, B, F, s));
  return lm_linear(h->g, h->act, B, F, w_down, tm_down, N, out, sp, s);
}

// one depth step: `x_in` (bf16 [B, LH]) or split-K partials of it -> logits of channel c in h->logits
static int local_depth_step(mt_local* h, int B, int c, int sp_in, const uint8_t* ones_mask, long long mask_ld, cudaStream_t s) {
  const mt_local_config& cf = h->cfg;
  const int LH = cf.local_hidden_size, LF = cf.local_ffn_hidden_size, AF = cf.additional_mlp_ffn_hidden_size;
  const int H = cf.lm.hidden_size, D = cf.lm.head_dim, C = h->C;
  const int qs = cf.lm.num_heads * D, kvs = cf.lm.num_kv_heads * D, nqkv = qs + 2 * kvs;
  const float eps = cf.lm.rms_norm_eps;
  const size_t layer_stride = static_cast<size_t>(cf.lm.max_batch) * cf.lm.num_kv_heads * C * D;
  const long long BL = static_cast<long long>(B) * LH;
  float* pk = h->g->splitk;
  // K rows of the earlier channels were written a whole depth step ago, behind at least one gate|up GEMM (lm_ops.cuh)
  const int early = attn_early_reads_safe(dec_grid_ctas(B, nqkv, LH, kSplitKBytes), dec_grid_ctas(B, 2 * LF, LH, kSplitKBytes));
  // x = input ; xn = norm(x)
  MT_TRY(add_rmsnorm(nullptr, h->delta, sp_in ? pk : nullptr, sp_in, BL, h->x, reinterpret_cast<const bf16*>(h->ln_in[0]),
                     h->xn, B, LH, eps, s));
  for (int l = 0; l < cf.local_num_layers; ++l) {
    int sp = 0;
    MT_TRY(lm_linear(h->g, h->xn, B, LH, h->w_qkv[l], &h->tm_qkv[l], nqkv, h->qkv, &sp, s));
    MT_TRY(attn_decode(h->qkv, sp ? pk : nullptr, sp, static_cast<long long>(B) * nqkv,
                       reinterpret_cast<const bf16*>(h->q_norm[l]), reinterpret_cast<const bf16*>(h->k_norm[l]), nullptr,
                       nullptr, h->k_cache + l * layer_stride, h->v_cache + l * layer_stride, ones_mask, mask_ld, h->attn, B,
                       cf.lm.num_heads, cf.lm.num_kv_heads, D, nullptr, c, c + 1, C, eps, s, early));
    MT_TRY(lm_linear(h->g, h->attn, B, qs, h->w_o[l], &h->tm_o[l], LH, h->delta, &sp, s));
    MT_TRY(add_rmsnorm(h->x, h->delta, sp ? pk : nullptr, sp, BL, h->x, reinterpret_cast<const bf16*>(h->ln_post[l]), h->xn,
                       B, LH, eps, s));
    MT_TRY(lm_linear(h->g, h->xn, B, LH, h->w_gu[l], &h->tm_gu[l], 2 * LF, h->gu, &sp, s));
    MT_TRY(silu_mul(h->gu, sp ? pk : nullptr, sp, static_cast<long long>(B) * 2 * LF, h->act, B, LF, s));
    MT_TRY(lm_linear(h->g, h->act, B, LF, h->w_down[l], &h->tm_down[l], LH, h->delta, &sp, s));
    const void* next_gain = l + 1 < cf.local_num_layers ? h->ln_in[l + 1] : h->w.l_final_norm;
    MT_TRY(add_rmsnorm(h->x, h->delta, sp ? pk : nullptr, sp, BL, h->x, reinterpret_cast<const bf16*>(next_gain), h->xn, B,
                       LH, eps, s));
  }
  // local_to_speech_embedding_mlps[c] -> layer_norm_before_lm_heads[c] -> lm_heads[c]   (:403-409)
  int sp = 0;
  MT_TRY(local_mlp(h, h->xn, B, LH, AF, H, h->out_gu[c], &h->tm_out_gu[c], h->out_down[c], &h->tm_out_down[c], h->delta, &sp, s));
  MT_TRY(moss_rmsnorm(h->delta, sp ? pk : nullptr, sp, static_cast<long long>(B) * H, reinterpret_cast<const bf16*>(h->head_norm[c]),
                      h->hn, B, H, 1e-6f, s));
  const int V = c == 0 ? cf.lm.vocab_size : cf.lm.audio_head_size;
  return lm_linear(h->g, h->hn, B, H, h->heads[c], &h->tm_heads[c], V, h->logits, nullptr, s);
}

struct LocalSampling {
  std::vector<int> do_sample, top_k;
  std::vector<float> temperature, top_p, penalty;
  unsigned seed;
};

// everything of one frame after the backbone step: `hs` = final-norm hidden state of the last position [B, H]
// Teacher forcing (mt_local_forward): `forced` [B, C] replaces the sampled tokens and every channel's logits
// are copied to logits_out [n_ch][B][ldv] (pad code = -inf on audio channels, :411-412); no commit.
static int local_frame(mt_local* h, const LocalState& st, const bf16* hs, int B, int n_ch, const LocalSampling& sm, int eos,
                       cudaStream_t s, const int64_t* forced = nullptr, bf16* logits_out = nullptr, long long ldv = 0) {
  const mt_local_config& cf = h->cfg;
  const int H = cf.lm.hidden_size, LH = cf.local_hidden_size, AF = cf.additional_mlp_ffn_hidden_size;
  const bf16* x_in = hs;
  for (int c = 0; c < n_ch; ++c) {
    int sp = 0;
    MT_TRY(local_mlp(h, x_in, B, H, AF, LH, h->w.in_gate_up, &h->tm_in_gu, h->w.in_down, &h->tm_in_down, h->delta, &sp, s));
    MT_TRY(local_depth_step(h, B, c, sp, st.ones, 0, s));
    const int V = c == 0 ? cf.lm.vocab_size : cf.lm.audio_head_size;
    if (forced) {
      if (c > 0) MT_TRY(fill_audio_pad(h->logits, B, V, s));
      MT_CUDA(cudaMemcpy2DAsync(logits_out + static_cast<long long>(c) * B * ldv, ldv * sizeof(bf16), h->logits,
                                static_cast<size_t>(V) * sizeof(bf16), static_cast<size_t>(V) * sizeof(bf16), B,
                                cudaMemcpyDeviceToDevice, s));
    } else {
      MT_TRY(local_sample(st, h->logits, V, c, sm.do_sample[c], sm.temperature[c], sm.penalty[c], sm.top_p[c], sm.top_k[c],
                          sm.seed, s, h->g->full_hist));
    }
    if (c + 1 < n_ch) {
      const bf16* table = c == 0 ? reinterpret_cast<const bf16*>(h->w.lm.embed_tokens)
                                 : reinterpret_cast<const bf16*>(h->w.lm.emb_ext) + static_cast<size_t>(c - 1) * cf.lm.audio_head_size * H;
      MT_TRY(embed_rows(forced ? forced : st.next_ids, h->C, c, table, V, h->emb, B, H, s));
      x_in = h->emb;
    }
  }
  if (forced) return MT_OK;
  return local_commit(st, eos, cf.lm.audio_pad_code, n_ch, s);
}

}  // namespace mt

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
using namespace mt;

extern "C" int mt_local_query_sizes(const mt_local_config* cfg, long long* kv_bytes, long long* workspace_bytes) {
  MT_TRY(check_local_cfg(cfg));
  long long kv = 0, ws = 0;
  MT_TRY(mt_lm_query_sizes(&cfg->lm, &kv, &ws));
  if (kv_bytes) *kv_bytes = kv;
  if (workspace_bytes) *workspace_bytes = ((ws + 255) & ~255ll) + static_cast<long long>(carve_local(nullptr, *cfg, nullptr));
  return MT_OK;
}

extern "C" int mt_local_create(const mt_local_config* cfg, const mt_local_weights* w, void* kv_cache, void* workspace,
                               mt_local** out) {
  MT_TRY(check_local_cfg(cfg));
  MT_CHECK(w && kv_cache && workspace && out, "mt_local_create: null argument");
  long long kv = 0, ws = 0;
  MT_TRY(mt_lm_query_sizes(&cfg->lm, &kv, &ws));
  mt_local* h = new mt_local();
  h->cfg = *cfg; h->w = *w; h->C = 1 + cfg->lm.n_vq;
  mt_lm_weights gw = w->lm;
  gw.text_head = nullptr; gw.audio_heads = nullptr;
  int st = mt_lm_create(&cfg->lm, &gw, kv_cache, workspace, &h->g);
  if (st != MT_OK) { delete h; return st; }
  carve_local(h, *cfg, reinterpret_cast<uint8_t*>(workspace) + ((ws + 255) & ~255ll));
  const int LL = cfg->local_num_layers, C = h->C;
  auto cp = [](std::vector<const void*>& v, const void* const* src, int n) { v.assign(src, src + n); };
  cp(h->w_qkv, w->l_w_qkv, LL); cp(h->w_o, w->l_w_o, LL); cp(h->w_gu, w->l_w_gate_up, LL); cp(h->w_down, w->l_w_down, LL);
  cp(h->ln_in, w->l_ln_in, LL); cp(h->ln_post, w->l_ln_post, LL); cp(h->q_norm, w->l_q_norm, LL); cp(h->k_norm, w->l_k_norm, LL);
  cp(h->out_gu, w->out_gate_up, C); cp(h->out_down, w->out_down, C); cp(h->head_norm, w->head_norm, C); cp(h->heads, w->heads, C);
  const long long H = cfg->lm.hidden_size, LH = cfg->local_hidden_size, LF = cfg->local_ffn_hidden_size;
  const long long AF = cfg->additional_mlp_ffn_hidden_size, D = cfg->lm.head_dim;
  const long long qs = cfg->lm.num_heads * D, kvs = cfg->lm.num_kv_heads * D;
  auto wmap = [&](CUtensorMap* tm, const void* base, long long N, long long K) {
    if (st != MT_OK) return;
    st = cfg->lm.weights_tiled ? make_tmap_tiled(tm, base, MT_DTYPE_BF16, N, K)
                               : make_tmap_2d(tm, base, MT_DTYPE_BF16, N, K, K, 128);
  };
  h->tm_qkv.resize(LL); h->tm_o.resize(LL); h->tm_gu.resize(LL); h->tm_down.resize(LL);
  h->tm_out_gu.resize(C); h->tm_out_down.resize(C); h->tm_heads.resize(C);
  for (int l = 0; l < LL; ++l) {
    wmap(&h->tm_qkv[l], h->w_qkv[l], qs + 2 * kvs, LH);
    wmap(&h->tm_o[l], h->w_o[l], LH, qs);
    wmap(&h->tm_gu[l], h->w_gu[l], 2 * LF, LH);
    wmap(&h->tm_down[l], h->w_down[l], LH, LF);
  }
  wmap(&h->tm_in_gu, w->in_gate_up, 2 * AF, H);
  wmap(&h->tm_in_down, w->in_down, LH, AF);
  for (int c = 0; c < C; ++c) {
    wmap(&h->tm_out_gu[c], h->out_gu[c], 2 * AF, LH);
    wmap(&h->tm_out_down[c], h->out_down[c], H, AF);
    wmap(&h->tm_heads[c], h->heads[c], c == 0 ? cfg->lm.vocab_size : cfg->lm.audio_head_size, H);
  }
  if (st != MT_OK) { mt_lm_destroy(h->g); delete h; return st; }
  *out = h;
  return MT_OK;
}

extern "C" int mt_local_destroy(mt_local* h) {
  if (!h) return MT_OK;
  mt_lm_destroy(h->g);
  delete h;
  return MT_OK;
}


extern "C" int mt_local_forward(mt_local* h, const int64_t* input_ids, const uint8_t* attention_mask, int B, int S0,
                                int n_vq_for_inference, const int64_t* frame_ids, void* logits_out, long long ldv,
                                mt_stream stream) {
  MT_CHECK(h && input_ids && frame_ids && logits_out, "mt_local_forward: null argument");
  const mt_lm_config& c = h->cfg.lm;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  MT_CHECK(B > 0 && B <= c.max_batch && S0 > 0 && S0 <= c.max_seq && B * S0 <= c.max_tokens,
           "mt_local_forward: B=%d S0=%d exceed the engine capacity", B, S0);
  MT_CHECK(n_vq_for_inference >= 0 && n_vq_for_inference <= c.n_vq, "mt_local_forward: bad n_vq_for_inference %d", n_vq_for_inference);
  MT_CHECK(ldv >= c.vocab_size && ldv >= c.audio_head_size, "mt_local_forward: ldv %lld is smaller than the widest head", ldv);
  const int n_ch = std::min(h->C, 1 + n_vq_for_inference);
  LocalState st = local_state_carve(h->state_buf, B, h->C, S0, c.audio_head_size, nullptr);
  MT_TRY(local_init(st, input_ids, S0, s, attention_mask));
  MT_TRY(lm_forward_impl(h->g, input_ids, st.mask, S0, B, S0, 0, nullptr, S0, 1, nullptr, nullptr, s, nullptr, n_vq_for_inference));
  LocalSampling none{};
  return local_frame(h, st, S0 > 1 ? h->g->last_h : h->g->xn, B, n_ch, none, -1, s, frame_ids, reinterpret_cast<bf16*>(logits_out), ldv);
}

extern "C" int mt_local_generate(mt_local* h, const int64_t* input_ids, int B, int S0, const mt_local_gen_params* p,
                                 int64_t* gen_ids, int* n_steps, float* step_ms, mt_stream stream) {
  MT_CHECK(h && input_ids && p && gen_ids && n_steps, "mt_local_generate: null argument");
  const mt_lm_config& c = h->cfg.lm;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int C = h->C, max_len = S0 + p->max_new_tokens;
  MT_CHECK(B > 0 && B <= c.max_batch, "generate: batch %d exceeds max_batch %d", B, c.max_batch);
  MT_CHECK(S0 > 0 && p->max_new_tokens >= 0 && max_len <= c.max_seq, "generate: %d positions exceed max_seq %d", max_len, c.max_seq);
  MT_CHECK(B * S0 <= c.max_tokens, "generate: prompt of %d tokens exceeds max_tokens %d", B * S0, c.max_tokens);
  MT_CHECK(p->n_vq_for_inference >= 0 && p->n_vq_for_inference <= c.n_vq, "generate: n_vq_for_inference %d outside [0, %d]",
           p->n_vq_for_inference, c.n_vq);
  MT_CHECK(p->do_sample && p->temperature && p->top_p && p->top_k && p->repetition_penalty, "generate: null per-channel array");
  const int n_inf = p->n_vq_for_inference, n_ch = std::min(C, 1 + n_inf);
  LocalSampling sm;
  sm.seed = p->seed;
  for (int ch = 0; ch < C; ++ch) {
    sm.do_sample.push_back(p->do_sample[ch] != 0);
    sm.temperature.push_back(p->temperature[ch]); sm.top_p.push_back(p->top_p[ch]); sm.top_k.push_back(p->top_k[ch]);
    sm.penalty.push_back(p->repetition_penalty[ch]);
    if (ch < n_ch && sm.do_sample[ch]) {
      MT_CHECK(sm.temperature[ch] > 0.f, "generate: channel %d temperature must be > 0", ch);
    }
  }
  *n_steps = 0;
  LocalState st = local_state_carve(h->state_buf, B, C, max_len, c.audio_head_size, gen_ids);
  MT_TRY(local_init(st, input_ids, S0, s, p->attention_mask));
  if (p->max_new_tokens == 0) return MT_OK;

  std::vector<cudaEvent_t> ev;
  if (step_ms) {
    ev.resize(p->max_new_tokens + 1);
    for (auto& e : ev) MT_CUDA(cudaEventCreate(&e));
    MT_CUDA(cudaEventRecord(ev[0], s));
  }
  // frame 0: backbone prefill over the prompt, then the depth loop
  MT_TRY(lm_forward_impl(h->g, input_ids, st.mask, max_len, B, S0, 0, nullptr, S0, 1, nullptr, nullptr, s, nullptr, n_inf));
  MT_TRY(local_frame(h, st, S0 > 1 ? h->g->last_h : h->g->xn, B, n_ch, sm, p->eos_token_id, s));
  if (step_ms) MT_CUDA(cudaEventRecord(ev[1], s));

  auto frame = [&](cudaStream_t cs) -> int {
    MT_TRY(lm_forward_impl(h->g, st.next_ids, st.mask, max_len, B, 1, 0, st.pos, max_len, 1, nullptr, nullptr, cs, nullptr, n_inf));
    return local_frame(h, st, h->g->xn, B, n_ch, sm, p->eos_token_id, cs);
  };

  int frames_launched = 1;
  int* done_host = nullptr;
  MT_CUDA(cudaMallocHost(&done_host, 4 * sizeof(int)));
  done_host[0] = done_host[1] = 0;
  // pipelined stop polling (see mt_lm_generate): the host waits for poll c-1 only after poll c is enqueued, so the GPU
  // never drains while a ~1,700-node frame graph is being launched
  cudaEvent_t poll_ev[2] = {nullptr, nullptr};
  for (auto& e : poll_ev) MT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  int polls = 0;
  const int check_every = p->check_stop_every > 0 ? p->check_stop_every : 4;
  cudaGraphExec_t exec = nullptr;
  long long per_replay = 0;
  int status = MT_OK;
  for (int step = 1; step < p->max_new_tokens && status == MT_OK; ++step) {
    if (p->use_cuda_graph && step >= 2) {
      if (!exec) {
        cudaGraph_t graph = nullptr;
        if (!h->g->cap_stream && cudaStreamCreateWithFlags(&h->g->cap_stream, cudaStreamNonBlocking) != cudaSuccess) { status = fail(MT_ERR_CUDA, "capture stream creation failed"); break; }
        cudaError_t be = cudaStreamBeginCapture(h->g->cap_stream, cudaStreamCaptureModeThreadLocal);
        if (be != cudaSuccess) { status = fail(MT_ERR_CUDA, "begin capture failed: %s", cudaGetErrorString(be)); break; }
        const long long before = g_launches;
        status = frame(h->g->cap_stream);
        cudaError_t ce = cudaStreamEndCapture(h->g->cap_stream, &graph);
        if (status == MT_OK && ce != cudaSuccess) status = fail(MT_ERR_CUDA, "end capture: %s", cudaGetErrorString(ce));
        if (status == MT_OK && cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) status = fail(MT_ERR_CUDA, "graph instantiate failed");
        if (graph) cudaGraphDestroy(graph);
        per_replay = g_launches - before;
        g_launches = before;
        if (status != MT_OK) break;
      }
      if (cudaGraphLaunch(exec, s) != cudaSuccess) { status = fail(MT_ERR_CUDA, "graph launch failed"); break; }
      g_launches += per_replay;
    } else {
      status = frame(s);
      if (status != MT_OK) break;
    }
    ++frames_launched;
    if (step_ms) cudaEventRecord(ev[frames_launched], s);
    if (step % check_every == 0) {
      const int slot = polls & 1;
      cudaMemcpyAsync(done_host + slot, st.done, sizeof(int), cudaMemcpyDeviceToHost, s);
      cudaEventRecord(poll_ev[slot], s);
      if (polls > 0) {
        if (cudaEventSynchronize(poll_ev[slot ^ 1]) != cudaSuccess) { status = fail(MT_ERR_CUDA, "frame failed: %s", cudaGetErrorString(cudaGetLastError())); break; }
        if (done_host[slot ^ 1]) break;
      }
      ++polls;
    }
  }
  if (status == MT_OK) {
    cudaMemcpyAsync(done_host + 2, st.step, sizeof(int), cudaMemcpyDeviceToHost, s);
    cudaError_t e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) status = fail(MT_ERR_CUDA, "generate failed: %s", cudaGetErrorString(e));
    else *n_steps = done_host[2];
  }
  if (step_ms && status == MT_OK) {
    for (int i = 0; i < p->max_new_tokens; ++i) step_ms[i] = -1.f;
    for (int i = 0; i < frames_launched; ++i) cudaEventElapsedTime(&step_ms[i], ev[i], ev[i + 1]);
  }
  for (auto& e : ev) cudaEventDestroy(e);
  if (exec) cudaGraphExecDestroy(exec);
  for (auto& e : poll_ev) if (e) cudaEventDestroy(e);
  cudaFreeHost(done_host);
  return status;
}
