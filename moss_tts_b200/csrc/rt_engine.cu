// MOSS-TTS-Realtime engine: the step-wise decode of MossTTSRealtimeInference.
// Reference: moss_tts_realtime/mossttsrealtime/streaming_mossttsrealtime.py  prefill :121-218, step :220-272,
//            finish :274-299, generate_local_transformer :301-357 (the one function the reference wraps in torch.compile),
//            apply_repetition_penalty :359-375, sample_token :377-421;
//            modeling_mossttsrealtime.py :101-142 (backbone), modeling_mossttsrealtime_local.py :123-448 (local transformer).
//
// One frame = one backbone decode step (an mt_lm created without heads: embedding sum of the text token and the rvq codes
// of the previous frame, Qwen3 layers, final norm) + rvq depth steps of the local transformer.  The local transformer is
// the backbone's layer body over its own head layout with rotary position = channel index and a KV cache of rvq slots
// (the reference's StaticCache(max_cache_len=channels), rebuilt every frame: slots past the current channel are masked
// there and simply not read here).  Session state (attention mask, history, stopping flags, counters) lives on the device,
// so a step is enqueued without a host round trip and replayed from a CUDA graph.
#include "lm_engine.cuh"

#include <algorithm>
#include <cstring>
#include <vector>

struct mt_rt {
  mt_rt_config cfg;
  mt_rt_weights w;
  mt_lm* g = nullptr;                 // backbone (owns its tensor maps and the split-K buffer)
  int R = 0;                          // rvq channels
  std::vector<const void*> w_qkv, w_o, w_gu, w_down, ln_in, ln_post, q_norm, k_norm, heads;
  std::vector<CUtensorMap> tm_qkv, tm_o, tm_gu, tm_down, tm_heads;
  mt::bf16 *x, *xn, *qkv, *attn, *gu, *act, *delta, *emb, *logits, *k_cache, *v_cache;
  void *state_buf, *aux_buf;
  uint8_t* ones;                      // [R] all ones: a depth step sees its whole causal prefix
  mt::RtState st{};
  int B = 0;                          // batch of the open session (0: none)
  int host_len = 0, host_frames = 0;  // upper bounds of the device counters (capacity checks without a sync)
  const mt::bf16* hid = nullptr;      // backbone hidden state of the newest position
  cudaGraphExec_t exec = nullptr;
  mt_rt_sample_params exec_p{};
  int exec_B = 0;
  long long exec_launches = 0;
  int* poll_host = nullptr;           // pinned: {done, done, len, step, done}
  cudaEvent_t poll_ev[2] = {nullptr, nullptr};
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

int check_rt_cfg(const mt_rt_config* c) {
  MT_CHECK(c, "null config");
  MT_CHECK(c->local_num_layers >= 1 && c->local_intermediate_size % 64 == 0, "bad local transformer shape");
  MT_CHECK(c->local_head_dim == 64 || c->local_head_dim == 128, "local_head_dim must be 64 or 128");
  MT_CHECK(c->local_num_kv_heads > 0 && c->local_num_heads % c->local_num_kv_heads == 0 &&
           c->local_num_heads / c->local_num_kv_heads <= 8, "local heads must be a multiple (<= 8x) of local kv heads");
  MT_CHECK(c->max_frames > 0, "max_frames must be positive");
  return MT_OK;
}

size_t carve_rt(mt_rt* h, const mt_rt_config& c, void* ws) {
  Carver cv(ws);
  const size_t B = c.lm.max_batch, H = c.lm.hidden_size, LI = c.local_intermediate_size, D = c.local_head_dim;
  const size_t qs = static_cast<size_t>(c.local_num_heads) * D, kvs = static_cast<size_t>(c.local_num_kv_heads) * D;
  const size_t R = c.lm.n_vq, V = c.lm.audio_head_size;
  auto x = cv.take<bf16>(B * H);
  auto xn = cv.take<bf16>(B * H);
  auto qkv = cv.take<bf16>(B * (qs + 2 * kvs));
  auto attn = cv.take<bf16>(B * qs);
  auto gu = cv.take<bf16>(B * 2 * LI);
  auto act = cv.take<bf16>(B * LI);
  auto delta = cv.take<bf16>(B * H);
  auto emb = cv.take<bf16>(B * H);
  auto logits = cv.take<bf16>(B * V);
  const size_t kv_half = static_cast<size_t>(c.local_num_layers) * B * c.local_num_kv_heads * R * D;
  auto kc = cv.take<bf16>(kv_half);
  auto vc = cv.take<bf16>(kv_half);
  auto sb = cv.take<uint8_t>(rt_state_bytes(c.lm.max_batch, static_cast<int>(R), c.max_frames, c.lm.max_seq, c.lm.audio_head_size));
  auto ab = cv.take<uint8_t>(rt_state_bytes(c.lm.max_batch, static_cast<int>(R), 1, 1, c.lm.audio_head_size));   // mt_rt_local_generate
  auto ones = cv.take<uint8_t>(R);
  if (h) {
    h->aux_buf = ab; h->ones = ones;
    h->x = x; h->xn = xn; h->qkv = qkv; h->attn = attn; h->gu = gu; h->act = act; h->delta = delta; h->emb = emb;
    h->logits = logits; h->k_cache = kc; h->v_cache = vc; h->state_buf = sb;
  }
  return cv.off + 256;
}

int check_params(const mt_rt_sample_params* p) {
  MT_CHECK(p, "null sampling parameters");
  if (p->do_sample && p->temperature != 0.f) {
    MT_CHECK(p->temperature > 0.f, "temperature must be >= 0");
    MT_CHECK(p->top_p >= 0.f && p->top_p <= 1.f, "`top_p` has to be a float > 0 and < 1, but is %f", p->top_p);
    if (p->top_k <= 0) return fail(MT_ERR_ARG, "`top_k` has to be a strictly positive integer, but is %d", p->top_k);
    if (p->top_k > kMaxTopK) return fail(MT_ERR_UNSUPPORTED, "sampling requires top_k <= %d (got %d)", kMaxTopK, p->top_k);
  }
  return MT_OK;
}

inline bool greedy(const mt_rt_sample_params& p) { return !p.do_sample || p.temperature == 0.f; }
inline float penalty_of(const mt_rt_sample_params& p) { return p.repetition_penalty == 0.f ? 1.f : p.repetition_penalty; }

// one depth step (modeling_mossttsrealtime_local.py:351-368 + :423): x_in [B,H] -> raw logits of channel c in h->logits
int depth_step(mt_rt* h, const bf16* x_in, int B, int c, cudaStream_t s) {
  const mt_rt_config& cf = h->cfg;
  const int H = cf.lm.hidden_size, LI = cf.local_intermediate_size, D = cf.local_head_dim, R = h->R;
  const int qs = cf.local_num_heads * D, kvs = cf.local_num_kv_heads * D, nqkv = qs + 2 * kvs;
  const float eps = cf.local_rms_norm_eps;
  const size_t layer_stride = static_cast<size_t>(cf.lm.max_batch) * cf.local_num_kv_heads * R * D;
  const long long BH = static_cast<long long>(B) * H;
  float* pk = h->g->splitk;
  // K rows of the earlier channels were written a whole depth step ago, behind at least one gate|up GEMM (lm_ops.cuh)
  const int early = attn_early_reads_safe(dec_grid_ctas(B, nqkv, H, kSplitKBytes), dec_grid_ctas(B, 2 * LI, H, kSplitKBytes));
  MT_TRY(add_rmsnorm(nullptr, x_in, nullptr, 0, 0, h->x, reinterpret_cast<const bf16*>(h->ln_in[0]), h->xn, B, H, eps, s));
  for (int l = 0; l < cf.local_num_layers; ++l) {
    int sp = 0;
    MT_TRY(lm_linear(h->g, h->xn, B, H, h->w_qkv[l], &h->tm_qkv[l], nqkv, h->qkv, &sp, s));
    MT_TRY(attn_decode(h->qkv, sp ? pk : nullptr, sp, static_cast<long long>(B) * nqkv,
                       reinterpret_cast<const bf16*>(h->q_norm[l]), reinterpret_cast<const bf16*>(h->k_norm[l]),
                       reinterpret_cast<const bf16*>(h->w.l_rope_cos), reinterpret_cast<const bf16*>(h->w.l_rope_sin),
                       h->k_cache + l * layer_stride, h->v_cache + l * layer_stride, h->ones, 0, h->attn, B,
                       cf.local_num_heads, cf.local_num_kv_heads, D, nullptr, c, c + 1, R, eps, s, early));
    MT_TRY(lm_linear(h->g, h->attn, B, qs, h->w_o[l], &h->tm_o[l], H, h->delta, &sp, s));
    MT_TRY(add_rmsnorm(h->x, h->delta, sp ? pk : nullptr, sp, BH, h->x, reinterpret_cast<const bf16*>(h->ln_post[l]), h->xn,
                       B, H, eps, s));
    MT_TRY(lm_linear(h->g, h->xn, B, H, h->w_gu[l], &h->tm_gu[l], 2 * LI, h->gu, &sp, s));
    MT_TRY(silu_mul(h->gu, sp ? pk : nullptr, sp, static_cast<long long>(B) * 2 * LI, h->act, B, LI, s));
    MT_TRY(lm_linear(h->g, h->act, B, LI, h->w_down[l], &h->tm_down[l], H, h->delta, &sp, s));
    const void* next_gain = l + 1 < cf.local_num_layers ? h->ln_in[l + 1] : h->w.l_final_norm;
    MT_TRY(add_rmsnorm(h->x, h->delta, sp ? pk : nullptr, sp, BH, h->x, reinterpret_cast<const bf16*>(next_gain), h->xn, B,
                       H, eps, s));
  }
  return lm_linear(h->g, h->xn, B, H, h->heads[c], &h->tm_heads[c], cf.lm.audio_head_size, h->logits, nullptr, s);
}

// generate_local_transformer :301-357.  Tokens go to `tok` ([B,R], st.last in a session); history/step select the
// repetition window.  Teacher forcing: `forced` replaces the sampled tokens, logits_out receives the raw logits.
int depth_loop(mt_rt* h, const RtState& st, const bf16* hidden, int B, const mt_rt_sample_params& p, const int64_t* history,
               long long ld_frames, const int* step_dev, int step_host, int64_t* tok, const int64_t* forced, bf16* logits_out,
               cudaStream_t s) {
  const mt_rt_config& cf = h->cfg;
  const int H = cf.lm.hidden_size, V = cf.lm.audio_head_size, R = h->R;
  const float pen = penalty_of(p);
  if (!forced && pen != 1.f) MT_TRY(rt_window_sets(st, history, ld_frames, step_dev, step_host, p.repetition_window, s));
  const bf16* x_in = hidden;
  for (int c = 0; c < R; ++c) {
    MT_TRY(depth_step(h, x_in, B, c, s));
    if (logits_out)
      MT_CUDA(cudaMemcpyAsync(logits_out + static_cast<long long>(c) * B * V, h->logits, static_cast<size_t>(B) * V * sizeof(bf16),
                              cudaMemcpyDeviceToDevice, s));
    if (!forced)
      MT_TRY(rt_sample(st, h->logits, V, c, !greedy(p), p.temperature, pen, p.top_p, p.top_k, p.seed, step_dev, step_host, tok, s));
    if (c + 1 < R) {
      const bf16* table = reinterpret_cast<const bf16*>(h->w.l_embed) + static_cast<size_t>(c) * V * H;
      MT_TRY(embed_rows(forced ? forced : tok, R, c, table, V, h->emb, B, H, s));
      x_in = h->emb;
    }
  }
  return MT_OK;
}

// the graph-capturable body of step(): :244-267
int step_body(mt_rt* h, const mt_rt_sample_params& p, cudaStream_t s) {
  const RtState& st = h->st;
  const int B = h->B, max_seq = h->cfg.lm.max_seq;
  MT_TRY(rt_begin_step(st, s));
  MT_TRY(lm_forward_impl(h->g, st.step_ids, st.mask, max_seq, B, 1, 0, st.pos, max_seq, 1, nullptr, nullptr, s));
  MT_TRY(depth_loop(h, st, h->g->xn, B, p, st.history, st.max_frames, st.step, 0, st.last, nullptr, nullptr, s));
  return rt_commit(st, h->cfg.audio_eos_token, 1, nullptr, s);
}

int enqueue_step(mt_rt* h, const mt_rt_sample_params& p, int use_graph, cudaStream_t s) {
  if (!use_graph) return step_body(h, p, s);
  if (h->exec && (h->exec_B != h->B || std::memcmp(&h->exec_p, &p, sizeof(p)) != 0)) {
    cudaGraphExecDestroy(h->exec);
    h->exec = nullptr;
  }
  if (!h->exec) {
    cudaGraph_t graph = nullptr;
    if (!h->g->cap_stream && cudaStreamCreateWithFlags(&h->g->cap_stream, cudaStreamNonBlocking) != cudaSuccess)
      return fail(MT_ERR_CUDA, "capture stream creation failed");
    cudaError_t be = cudaStreamBeginCapture(h->g->cap_stream, cudaStreamCaptureModeThreadLocal);
    if (be != cudaSuccess) return fail(MT_ERR_CUDA, "begin capture failed: %s", cudaGetErrorString(be));
    const long long before = g_launches;
    int status = step_body(h, p, h->g->cap_stream);
    cudaError_t ce = cudaStreamEndCapture(h->g->cap_stream, &graph);       // always ends the capture, also on failure
    if (status == MT_OK && ce != cudaSuccess) status = fail(MT_ERR_CUDA, "end capture: %s", cudaGetErrorString(ce));
    if (status == MT_OK && cudaGraphInstantiate(&h->exec, graph, 0) != cudaSuccess) status = fail(MT_ERR_CUDA, "graph instantiate failed");
    if (graph) cudaGraphDestroy(graph);
    h->exec_launches = g_launches - before;
    g_launches = before;
    if (status != MT_OK) { h->exec = nullptr; return status; }
    h->exec_p = p; h->exec_B = h->B;
  }
  MT_CUDA(cudaGraphLaunch(h->exec, s));
  g_launches += h->exec_launches;
  return MT_OK;
}

int check_step_capacity(mt_rt* h) {
  MT_CHECK(h->B > 0, "You must call prefill() before step().");
  MT_CHECK(h->host_len + 1 <= h->cfg.lm.max_seq, "step: %d positions exceed max_seq %d", h->host_len + 1, h->cfg.lm.max_seq);
  MT_CHECK(h->host_frames + 1 <= h->cfg.max_frames, "step: %d frames exceed max_frames %d", h->host_frames + 1, h->cfg.max_frames);
  return MT_OK;
}

}  // namespace

}  // namespace mt

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
using namespace mt;

extern "C" int mt_rt_query_sizes(const mt_rt_config* cfg, long long* kv_bytes, long long* workspace_bytes) {
  MT_TRY(check_rt_cfg(cfg));
  long long kv = 0, ws = 0;
  MT_TRY(mt_lm_query_sizes(&cfg->lm, &kv, &ws));
  if (kv_bytes) *kv_bytes = kv;
  if (workspace_bytes) *workspace_bytes = ((ws + 255) & ~255ll) + static_cast<long long>(carve_rt(nullptr, *cfg, nullptr));
  return MT_OK;
}

extern "C" int mt_rt_destroy(mt_rt* h) {
  if (!h) return MT_OK;
  if (h->exec) cudaGraphExecDestroy(h->exec);
  for (auto& e : h->poll_ev) if (e) cudaEventDestroy(e);
  if (h->poll_host) cudaFreeHost(h->poll_host);
  mt_lm_destroy(h->g);
  delete h;
  return MT_OK;
}

extern "C" int mt_rt_create(const mt_rt_config* cfg, const mt_rt_weights* w, void* kv_cache, void* workspace, mt_rt** out) {
  MT_TRY(check_rt_cfg(cfg));
  MT_CHECK(w && kv_cache && workspace && out, "mt_rt_create: null argument");
  MT_CHECK(w->l_embed && w->l_heads && w->l_rope_cos && w->l_rope_sin && w->l_final_norm, "mt_rt_create: null local weight");
  long long kv = 0, ws = 0;
  MT_TRY(mt_lm_query_sizes(&cfg->lm, &kv, &ws));
  mt_rt* h = new mt_rt();
  h->cfg = *cfg; h->w = *w; h->R = cfg->lm.n_vq;
  mt_lm_weights gw = w->lm;
  gw.text_head = nullptr; gw.audio_heads = nullptr;
  int st = mt_lm_create(&cfg->lm, &gw, kv_cache, workspace, &h->g);
  if (st != MT_OK) { delete h; return st; }
  carve_rt(h, *cfg, reinterpret_cast<uint8_t*>(workspace) + ((ws + 255) & ~255ll));
  h->st = rt_state_carve(h->state_buf, cfg->lm.max_batch, h->R, cfg->max_frames, cfg->lm.max_seq, cfg->lm.audio_head_size);
  const int LL = cfg->local_num_layers, R = h->R;
  auto cp = [](std::vector<const void*>& v, const void* const* src, int n) { v.assign(src, src + n); };
  cp(h->w_qkv, w->l_w_qkv, LL); cp(h->w_o, w->l_w_o, LL); cp(h->w_gu, w->l_w_gate_up, LL); cp(h->w_down, w->l_w_down, LL);
  cp(h->ln_in, w->l_ln_in, LL); cp(h->ln_post, w->l_ln_post, LL); cp(h->q_norm, w->l_q_norm, LL); cp(h->k_norm, w->l_k_norm, LL);
  cp(h->heads, w->l_heads, R);
  const long long H = cfg->lm.hidden_size, LI = cfg->local_intermediate_size, D = cfg->local_head_dim;
  const long long qs = cfg->local_num_heads * D, kvs = cfg->local_num_kv_heads * D;
  auto wmap = [&](CUtensorMap* tm, const void* base, long long N, long long K) {
    if (st != MT_OK) return;
    st = cfg->lm.weights_tiled ? make_tmap_tiled(tm, base, MT_DTYPE_BF16, N, K)
                               : make_tmap_2d(tm, base, MT_DTYPE_BF16, N, K, K, 128);
  };
  h->tm_qkv.resize(LL); h->tm_o.resize(LL); h->tm_gu.resize(LL); h->tm_down.resize(LL); h->tm_heads.resize(R);
  for (int l = 0; l < LL; ++l) {
    wmap(&h->tm_qkv[l], h->w_qkv[l], qs + 2 * kvs, H);
    wmap(&h->tm_o[l], h->w_o[l], H, qs);
    wmap(&h->tm_gu[l], h->w_gu[l], 2 * LI, H);
    wmap(&h->tm_down[l], h->w_down[l], H, LI);
  }
  for (int c = 0; c < R; ++c) wmap(&h->tm_heads[c], h->heads[c], cfg->lm.audio_head_size, H);
  if (st == MT_OK && cudaMemset(h->ones, 1, R) != cudaSuccess) st = fail(MT_ERR_CUDA, "mt_rt_create: workspace is not device memory");
  if (st == MT_OK && cudaMallocHost(&h->poll_host, 8 * sizeof(int)) != cudaSuccess) st = fail(MT_ERR_CUDA, "mt_rt_create: pinned poll slots");
  for (auto& e : h->poll_ev)
    if (st == MT_OK && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) st = fail(MT_ERR_CUDA, "mt_rt_create: poll event");
  if (st != MT_OK) { mt_rt_destroy(h); return st; }
  *out = h;
  return MT_OK;
}

extern "C" int mt_rt_prefill(mt_rt* h, const int64_t* input_ids, const uint8_t* attention_mask, int B, int S0, int keep_cache,
                             const mt_rt_sample_params* p, int64_t* audio_tokens, mt_stream stream) {
  MT_CHECK(h && input_ids && attention_mask && audio_tokens, "mt_rt_prefill: null argument");
  MT_TRY(check_params(p));
  const mt_lm_config& c = h->cfg.lm;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  MT_CHECK(B > 0 && B <= c.max_batch, "prefill: batch %d exceeds max_batch %d", B, c.max_batch);
  MT_CHECK(S0 > 0, "Prefill requires at least one text token.");
  int past = 0;
  if (keep_cache && h->B > 0) {
    MT_CHECK(B == h->B, "prefill: batch %d differs from the cached conversation's %d", B, h->B);
    MT_CUDA(cudaMemcpyAsync(h->poll_host + 2, h->st.len, sizeof(int), cudaMemcpyDeviceToHost, s));
    MT_CUDA(cudaStreamSynchronize(s));
    past = h->poll_host[2];
  }
  MT_CHECK(past + S0 <= c.max_seq, "prefill: %d positions exceed max_seq %d", past + S0, c.max_seq);
  MT_CHECK(static_cast<long long>(B) * S0 <= c.max_tokens, "prefill: %d tokens exceed max_tokens %d", B * S0, c.max_tokens);
  if (h->B != B) h->st = rt_state_carve(h->state_buf, B, h->R, h->cfg.max_frames, c.max_seq, c.audio_head_size);
  h->B = B;
  const RtState& st = h->st;
  MT_TRY(rt_prefill_init(st, attention_mask, S0, past, s));
  MT_TRY(lm_forward_impl(h->g, input_ids, st.mask, c.max_seq, B, S0, past, nullptr, past + S0, 1, nullptr, nullptr, s));
  h->hid = S0 > 1 ? h->g->last_h : h->g->xn;
  MT_TRY(depth_loop(h, st, h->hid, B, *p, nullptr, st.max_frames, nullptr, 0, st.last, nullptr, nullptr, s));
  MT_TRY(rt_commit(st, h->cfg.audio_eos_token, 0, audio_tokens, s));
  h->host_len = past + S0;
  h->host_frames = 1;
  return MT_OK;
}

extern "C" int mt_rt_step(mt_rt* h, const int64_t* text_tokens, const mt_rt_sample_params* p, int use_cuda_graph,
                          int64_t* audio_tokens, mt_stream stream) {
  MT_CHECK(h && audio_tokens, "mt_rt_step: null argument");
  MT_TRY(check_params(p));
  MT_TRY(check_step_capacity(h));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  MT_TRY(rt_set_text(h->st, text_tokens, h->cfg.text_pad_id, s));
  MT_TRY(enqueue_step(h, *p, use_cuda_graph, s));
  h->hid = h->g->xn;
  ++h->host_len; ++h->host_frames;
  MT_CUDA(cudaMemcpyAsync(audio_tokens, h->st.last, static_cast<size_t>(h->B) * h->R * sizeof(int64_t), cudaMemcpyDeviceToDevice, s));
  return MT_OK;
}

extern "C" int mt_rt_finish(mt_rt* h, int max_steps, const mt_rt_sample_params* p, int use_cuda_graph, int64_t* frames,
                            int* n_steps, mt_stream stream) {
  MT_CHECK(h && frames && n_steps, "mt_rt_finish: null argument");
  MT_TRY(check_params(p));
  MT_CHECK(h->B > 0, "You must call prefill() before step().");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const size_t frame_elems = static_cast<size_t>(h->B) * h->R;
  *n_steps = 0;
  // the reference tests is_finished before every step (:286); the flag of poll k is read after poll k+1 is enqueued
  // so the GPU never drains (frames enqueued past the stop change nothing and are not counted)
  int* ph = h->poll_host;
  MT_CUDA(cudaMemcpyAsync(ph + 3, h->st.step, sizeof(int), cudaMemcpyDeviceToHost, s));
  MT_CUDA(cudaMemcpyAsync(ph + 4, h->st.done, sizeof(int), cudaMemcpyDeviceToHost, s));
  MT_CUDA(cudaStreamSynchronize(s));
  const int step0 = ph[3];
  if (ph[4] || max_steps <= 0) return MT_OK;
  MT_TRY(rt_set_text(h->st, nullptr, h->cfg.text_pad_id, s));
  int launched = 0, polls = 0;
  bool stop = false;
  while (launched < max_steps && !stop) {
    MT_TRY(check_step_capacity(h));
    MT_TRY(enqueue_step(h, *p, use_cuda_graph, s));
    ++h->host_len; ++h->host_frames;
    MT_CUDA(cudaMemcpyAsync(frames + launched * frame_elems, h->st.last, frame_elems * sizeof(int64_t), cudaMemcpyDeviceToDevice, s));
    ++launched;
    if (launched % 4 == 0) {
      const int slot = polls & 1;
      MT_CUDA(cudaMemcpyAsync(ph + slot, h->st.done, sizeof(int), cudaMemcpyDeviceToHost, s));
      MT_CUDA(cudaEventRecord(h->poll_ev[slot], s));
      if (polls > 0) {
        MT_CUDA(cudaEventSynchronize(h->poll_ev[slot ^ 1]));
        if (ph[slot ^ 1]) stop = true;
      }
      ++polls;
    }
  }
  MT_CUDA(cudaMemcpyAsync(ph + 3, h->st.step, sizeof(int), cudaMemcpyDeviceToHost, s));
  MT_CUDA(cudaMemcpyAsync(ph + 2, h->st.len, sizeof(int), cudaMemcpyDeviceToHost, s));
  MT_CUDA(cudaStreamSynchronize(s));
  *n_steps = std::min(ph[3] - step0, launched);
  h->hid = h->g->xn;
  h->host_len = ph[2]; h->host_frames = ph[3];
  return MT_OK;
}

extern "C" int mt_rt_state(mt_rt* h, uint8_t* stopping, uint8_t* mask, long long mask_ld, int* len, int* step_idx, int* finished,
                           mt_stream stream) {
  MT_CHECK(h && h->B > 0, "mt_rt_state: no session (call prefill first)");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int* ph = h->poll_host;
  MT_CUDA(cudaMemcpyAsync(ph + 2, h->st.len, sizeof(int), cudaMemcpyDeviceToHost, s));
  MT_CUDA(cudaMemcpyAsync(ph + 3, h->st.step, sizeof(int), cudaMemcpyDeviceToHost, s));
  MT_CUDA(cudaMemcpyAsync(ph + 4, h->st.done, sizeof(int), cudaMemcpyDeviceToHost, s));
  MT_CUDA(cudaStreamSynchronize(s));
  if (len) *len = ph[2];
  if (step_idx) *step_idx = ph[3];
  if (finished) *finished = ph[4];
  if (stopping) MT_CUDA(cudaMemcpyAsync(stopping, h->st.stopping, h->B, cudaMemcpyDeviceToDevice, s));
  if (mask) {
    MT_CHECK(mask_ld >= ph[2], "mt_rt_state: mask_ld %lld is smaller than the %d filled positions", mask_ld, ph[2]);
    MT_CUDA(cudaMemcpy2DAsync(mask, static_cast<size_t>(mask_ld), h->st.mask, static_cast<size_t>(h->st.max_seq),
                              static_cast<size_t>(ph[2]), h->B, cudaMemcpyDeviceToDevice, s));
  }
  return MT_OK;
}

extern "C" int mt_rt_local_generate(mt_rt* h, const void* hidden, int B, const int64_t* history, int gen_step,
                                    const mt_rt_sample_params* p, const int64_t* forced, void* logits_out, int64_t* audio_tokens,
                                    mt_stream stream) {
  MT_CHECK(h && hidden && (audio_tokens || forced), "mt_rt_local_generate: null argument");
  MT_TRY(check_params(p));
  MT_CHECK(B > 0 && B <= h->cfg.lm.max_batch, "local_generate: batch %d exceeds max_batch %d", B, h->cfg.lm.max_batch);
  MT_CHECK(gen_step >= 0 && (gen_step == 0 || history || forced || penalty_of(*p) == 1.f), "local_generate: gen_step %d needs a history", gen_step);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  // its own small state (id-set bitmaps, a cleared done flag, a token row): an open session is left untouched
  RtState st = rt_state_carve(h->aux_buf, B, h->R, 1, 1, h->cfg.lm.audio_head_size);
  MT_CUDA(cudaMemsetAsync(st.done, 0, sizeof(int), s));
  int64_t* tok = audio_tokens ? audio_tokens : st.last;
  return depth_loop(h, st, reinterpret_cast<const bf16*>(hidden), B, *p, history, gen_step, nullptr, gen_step, tok, forced,
                    reinterpret_cast<bf16*>(logits_out), s);
}

extern "C" int mt_rt_last_hidden(mt_rt* h, int B, void* out, mt_stream stream) {
  MT_CHECK(h && out && h->hid && B == h->B, "mt_rt_last_hidden: no step has run for batch %d", B);
  MT_CUDA(cudaMemcpyAsync(out, h->hid, static_cast<size_t>(B) * h->cfg.lm.hidden_size * sizeof(bf16), cudaMemcpyDeviceToDevice,
                          reinterpret_cast<cudaStream_t>(stream)));
  return MT_OK;
}
