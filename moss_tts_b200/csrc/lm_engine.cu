// Host-side engine of the MossTTSDelay hot path: owns tensor maps, carves the caller's workspace,
// sequences one forward (prefill or single-token decode) and the generate() loop with a CUDA graph.
// Reference: moss_tts_delay/modeling_moss_tts.py forward :225-300, generate :392-525.
#include "lm_engine.cuh"

#include <stdlib.h>
#include <algorithm>
#include <vector>

namespace mt {

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


}  // namespace mt

namespace mt {

static size_t carve_workspace(mt_lm* h, const mt_lm_config& c, void* ws) {
  Carver cv(ws);
  const size_t M = c.max_tokens, H = c.hidden_size, I = c.intermediate_size;
  const size_t qs = static_cast<size_t>(c.num_heads) * c.head_dim, kvs = static_cast<size_t>(c.num_kv_heads) * c.head_dim;
  auto x = cv.take<bf16>(M * H);
  auto xn = cv.take<bf16>(M * H);
  auto qkv = cv.take<bf16>(M * (qs + 2 * kvs));
  auto q = cv.take<bf16>(M * qs);
  auto attn = cv.take<bf16>(M * qs);
  auto gu = cv.take<bf16>(M * 2 * I);
  auto act = cv.take<bf16>(M * I);
  auto delta = cv.take<bf16>(M * H);
  auto last_h = cv.take<bf16>(static_cast<size_t>(c.max_batch) * H);
  auto tl = cv.take<bf16>(static_cast<size_t>(c.max_batch) * c.vocab_size);
  auto al = cv.take<bf16>(static_cast<size_t>(c.max_batch) * c.n_vq * c.audio_head_size);
  auto sk = cv.take<float>(kSplitKBytes / 4);
  auto gs = cv.take<uint8_t>(gen_state_bytes(c.max_batch, 1 + c.n_vq, c.max_seq, c.audio_head_size));
  auto fh = cv.take<uint32_t>(static_cast<size_t>(c.max_batch) * (sample_full_scratch_bytes() / 4));   // text_top_k = None
  if (h) {
    h->full_hist = fh;
    h->x = x; h->xn = xn; h->qkv = qkv; h->q = q; h->attn = attn; h->gu = gu; h->act = act; h->delta = delta;
    h->last_h = last_h; h->text_logits = tl; h->audio_logits = al; h->splitk = sk; h->gen_state_buf = gs;
  }
  return cv.off + 256;
}

static size_t kv_bytes_of(const mt_lm_config& c) {
  return static_cast<size_t>(c.num_layers) * 2 * c.max_batch * c.num_kv_heads * c.max_seq * c.head_dim * sizeof(bf16);
}

static int check_cfg(const mt_lm_config* c) {
  MT_CHECK(c, "null config");
  MT_CHECK(c->hidden_size % 64 == 0 && c->intermediate_size % 64 == 0, "hidden/intermediate size must be multiples of 64");
  MT_CHECK(c->head_dim == 64 || c->head_dim == 128, "head_dim must be 64 or 128");
  MT_CHECK(c->num_heads % c->num_kv_heads == 0, "num_heads must be a multiple of num_kv_heads");
  MT_CHECK(c->max_batch > 0 && c->max_seq > 0 && c->max_tokens >= c->max_batch, "bad capacity (max_batch/max_seq/max_tokens)");
  MT_CHECK(c->n_vq >= 1 && c->audio_head_size >= 2, "bad n_vq/audio_head_size");
  return MT_OK;
}

// One Linear on the LM path.  Small M: split-K partials handed to the consumer kernel (returns the
// split count through *splits, partials in h->splitk); large M: finished bf16 output in `out`.
int lm_linear(mt_lm* h, const bf16* x, int M, int K, const void* w, const CUtensorMap* tm, int N, bf16* out,
              int* splits, cudaStream_t s, const int* tile_filter_flag, int tile_lo, int tile_hi) {
  LinearPlan p;
  p.M = M; p.N = N; p.K = K; p.x = x; p.ldx = K; p.w = w; p.ldw = K; p.out = out; p.ldo = N;
  p.round_bf16 = 0; p.w_tmap = tm;
  p.w_static = true;                       // model weights never change while the engine is alive
  p.w_tiled = h->cfg.weights_tiled != 0;
  p.workspace = h->splitk; p.workspace_bytes = kSplitKBytes;
  if (tile_filter_flag && M <= 256) {
    p.tile_filter_flag = tile_filter_flag; p.tile_lo = tile_lo; p.tile_hi = tile_hi;
    p.workspace = nullptr; p.workspace_bytes = 0;        // the filter needs an unsplit GEMM
  }
  if (splits) {
    if (M <= 256 && static_cast<long long>(M) * N * 4 <= kSplitKBytes) p.out_splits = splits;
    else { *splits = 0; p.workspace = nullptr; p.workspace_bytes = 0; }
  }
  return linear(p, s);
}

// forward over B*T tokens; ids may live in the generation state (decode) or come from the caller.
int lm_forward_impl(mt_lm* h, const int64_t* ids, const uint8_t* mask, long long mask_ld, int B, int T,
                    int past_len, const int* past_len_dev, int max_keys, int last_only, bf16* text_logits,
                    bf16* audio_logits, cudaStream_t s, const int* text_full_flag, int n_vq_embed) {
  const mt_lm_config& c = h->cfg;
  const int M = B * T, H = c.hidden_size, I = c.intermediate_size, D = c.head_dim;
  const int qs = c.num_heads * D, kvs = c.num_kv_heads * D, nqkv = qs + 2 * kvs;
  MT_CHECK(M <= c.max_tokens, "forward: %d tokens exceed max_tokens %d", M, c.max_tokens);
  MT_CHECK(B <= c.max_batch && past_len + T <= c.max_seq, "forward: batch %d / length %d exceed KV capacity", B, past_len + T);
  const size_t layer_stride = static_cast<size_t>(c.max_batch) * c.num_kv_heads * c.max_seq * D;
  const float eps = c.rms_norm_eps;
  const long long MH = static_cast<long long>(M) * H;

  const bool fused_attn = T == 1 && attn_decode_fits(B, c.num_heads, c.num_kv_heads, D, max_keys);
  // layers >= 1 may read the position / rotary row / old K rows before their dependency wait (lm_ops.cuh: attn_decode)
  const bool early_reads = fused_attn && attn_early_reads_safe(dec_grid_ctas(M, nqkv, H, kSplitKBytes), dec_grid_ctas(M, 2 * I, H, kSplitKBytes));
  MT_TRY(embed_sum(ids, reinterpret_cast<const bf16*>(h->w.embed_tokens), reinterpret_cast<const bf16*>(h->w.emb_ext),
                   h->x, M, H, n_vq_embed < 0 ? c.n_vq : n_vq_embed, c.audio_head_size, c.vocab_size, s, 1 + c.n_vq));
  MT_TRY(add_rmsnorm(nullptr, h->x, nullptr, 0, 0, nullptr, reinterpret_cast<const bf16*>(h->ln_in[0]), h->xn, M, H, eps, s));
  for (int l = 0; l < c.num_layers; ++l) {
    bf16* kc = h->k_cache + l * layer_stride;
    bf16* vc = h->v_cache + l * layer_stride;
    int sp = 0;
    MT_TRY(lm_linear(h, h->xn, M, H, h->w_qkv[l], &h->tm_qkv[l], nqkv, h->qkv, &sp, s));
    if (fused_attn) {
      // decode: q/k norm + RoPE + KV append + attention in one warp-per-(row, kv head) kernel
      MT_TRY(attn_decode(h->qkv, sp ? h->splitk : nullptr, sp, static_cast<long long>(M) * nqkv,
                           reinterpret_cast<const bf16*>(h->q_norm[l]), reinterpret_cast<const bf16*>(h->k_norm[l]),
                           reinterpret_cast<const bf16*>(h->w.rope_cos), reinterpret_cast<const bf16*>(h->w.rope_sin), kc, vc,
                           mask, mask_ld, h->attn, B, c.num_heads, c.num_kv_heads, D, past_len_dev, past_len, max_keys,
                           c.max_seq, eps, s, l > 0 && early_reads));     // layer 0 follows the kernel that writes the position too closely
    } else {
      MT_TRY(qkv_post(h->qkv, sp ? h->splitk : nullptr, sp, static_cast<long long>(M) * nqkv,
                      reinterpret_cast<const bf16*>(h->q_norm[l]), reinterpret_cast<const bf16*>(h->k_norm[l]),
                      reinterpret_cast<const bf16*>(h->w.rope_cos), reinterpret_cast<const bf16*>(h->w.rope_sin), h->q, kc,
                      vc, B, T, c.num_heads, c.num_kv_heads, D, past_len, past_len_dev, c.max_seq, eps, s));
      MT_TRY(attention(h->q, kc, vc, mask, mask_ld, h->attn, B, T, c.num_heads, c.num_kv_heads, D, past_len,
                         past_len_dev, max_keys, c.max_seq, s));
    }
    MT_TRY(lm_linear(h, h->attn, M, qs, h->w_o[l], &h->tm_o[l], H, h->delta, &sp, s));
    MT_TRY(add_rmsnorm(h->x, h->delta, sp ? h->splitk : nullptr, sp, MH, h->x,
                       reinterpret_cast<const bf16*>(h->ln_post[l]), h->xn, M, H, eps, s));
    MT_TRY(lm_linear(h, h->xn, M, H, h->w_gu[l], &h->tm_gu[l], 2 * I, h->gu, &sp, s));
    MT_TRY(silu_mul(h->gu, sp ? h->splitk : nullptr, sp, static_cast<long long>(M) * 2 * I, h->act, M, I, s));
    MT_TRY(lm_linear(h, h->act, M, I, h->w_down[l], &h->tm_down[l], H, h->delta, &sp, s));
    const void* next_gain = l + 1 < c.num_layers ? h->ln_in[l + 1] : h->w.final_norm;
    MT_TRY(add_rmsnorm(h->x, h->delta, sp ? h->splitk : nullptr, sp, MH, h->x,
                       reinterpret_cast<const bf16*>(next_gain), h->xn, M, H, eps, s));
  }
  // heads (modeling_moss_tts.py:292-300)
  const bf16* hs = h->xn;
  int Mh = M;
  if (last_only && T > 1) {
    MT_TRY(gather_last(h->xn, h->last_h, B, T, H, s));
    hs = h->last_h;
  }
  if (last_only) Mh = B;
  MT_CHECK((!text_logits || h->w.text_head) && (!audio_logits || h->w.audio_heads), "forward: this engine was created without heads");
  if (text_logits) {
    // generate(): while every row is forced or inside an audio span the sampler reads only the gen/delay-slot
    // logits, so the 1.2 GB text head shrinks to the weight tile(s) holding those two rows
    const int lo = std::min(c.audio_assistant_gen_slot_token_id, c.audio_assistant_delay_slot_token_id);
    const int hi = std::max(c.audio_assistant_gen_slot_token_id, c.audio_assistant_delay_slot_token_id);
    MT_TRY(lm_linear(h, hs, Mh, H, h->w.text_head, &h->tm_text, c.vocab_size, text_logits, nullptr, s,
                     text_full_flag, lo, hi));
  }
  if (audio_logits) {
    const int NA = c.n_vq * c.audio_head_size;
    MT_TRY(lm_linear(h, hs, Mh, H, h->w.audio_heads, &h->tm_audio, NA, audio_logits, nullptr, s));
    MT_TRY(fill_audio_pad(audio_logits, static_cast<long long>(Mh) * c.n_vq, c.audio_head_size, s));
  }
  return MT_OK;
}

}  // namespace mt

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
using namespace mt;

extern "C" int mt_lm_query_sizes(const mt_lm_config* cfg, long long* kv_bytes, long long* workspace_bytes) {
  MT_TRY(check_cfg(cfg));
  if (kv_bytes) *kv_bytes = static_cast<long long>(kv_bytes_of(*cfg));
  if (workspace_bytes) *workspace_bytes = static_cast<long long>(carve_workspace(nullptr, *cfg, nullptr));
  return MT_OK;
}

extern "C" int mt_lm_create(const mt_lm_config* cfg, const mt_lm_weights* w, void* kv_cache, void* workspace,
                            mt_lm** out) {
  MT_TRY(check_cfg(cfg));
  MT_CHECK(w && kv_cache && workspace && out, "mt_lm_create: null argument");
  int dev = 0;
  MT_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  MT_CUDA(cudaGetDeviceProperties(&prop, dev));
  if (prop.major != 10) return fail(MT_ERR_UNSUPPORTED, "libmosstts_sm100 needs an sm_100 device, found sm_%d%d", prop.major, prop.minor);
  mt_lm* h = new mt_lm();
  h->cfg = *cfg; h->w = *w;
  const int L = cfg->num_layers;
  auto cp = [&](std::vector<const void*>& v, const void* const* src) { v.assign(src, src + L); };
  cp(h->w_qkv, w->w_qkv); cp(h->w_o, w->w_o); cp(h->w_gu, w->w_gate_up); cp(h->w_down, w->w_down);
  cp(h->ln_in, w->ln_in); cp(h->ln_post, w->ln_post); cp(h->q_norm, w->q_norm); cp(h->k_norm, w->k_norm);
  h->ws_bytes = carve_workspace(h, *cfg, workspace);
  const size_t half = kv_bytes_of(*cfg) / 2 / sizeof(bf16);
  h->k_cache = reinterpret_cast<bf16*>(kv_cache);
  h->v_cache = h->k_cache + half;
  const int H = cfg->hidden_size, I = cfg->intermediate_size, D = cfg->head_dim;
  const int qs = cfg->num_heads * D, kvs = cfg->num_kv_heads * D;
  h->tm_qkv.resize(L); h->tm_o.resize(L); h->tm_gu.resize(L); h->tm_down.resize(L);
  int st = MT_OK;
  const long long NA = static_cast<long long>(cfg->n_vq) * cfg->audio_head_size;
  auto wmap = [&](CUtensorMap* tm, const void* base, long long N, long long K) {
    return cfg->weights_tiled ? make_tmap_tiled(tm, base, MT_DTYPE_BF16, N, K)
                              : make_tmap_2d(tm, base, MT_DTYPE_BF16, N, K, K, 128);
  };
  for (int l = 0; l < L && st == MT_OK; ++l) {
    st = wmap(&h->tm_qkv[l], h->w_qkv[l], qs + 2 * kvs, H);
    if (st == MT_OK) st = wmap(&h->tm_o[l], h->w_o[l], H, qs);
    if (st == MT_OK) st = wmap(&h->tm_gu[l], h->w_gu[l], 2 * I, H);
    if (st == MT_OK) st = wmap(&h->tm_down[l], h->w_down[l], H, I);
  }
  // a backbone without heads (MossTTSLocal's global transformer) passes NULL head weights
  if (st == MT_OK && w->text_head) st = wmap(&h->tm_text, w->text_head, cfg->vocab_size, H);
  if (st == MT_OK && w->audio_heads) st = wmap(&h->tm_audio, w->audio_heads, NA, H);
  if (st == MT_OK && cudaMallocHost(&h->done_host, 4 * sizeof(int)) != cudaSuccess) st = fail(MT_ERR_CUDA, "mt_lm_create: pinned poll slots");
  for (auto& e : h->poll_ev)
    if (st == MT_OK && cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) st = fail(MT_ERR_CUDA, "mt_lm_create: poll event");
  if (st != MT_OK) { mt_lm_destroy(h); return st; }
  *out = h;
  return MT_OK;
}

extern "C" int mt_lm_destroy(mt_lm* h) {
  if (!h) return MT_OK;
  if (h->graph_exec) cudaGraphExecDestroy(h->graph_exec);
  if (h->sess.exec) cudaGraphExecDestroy(h->sess.exec);
  if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
  for (auto& e : h->poll_ev) if (e) cudaEventDestroy(e);
  if (h->done_host) cudaFreeHost(h->done_host);
  delete h;
  return MT_OK;
}

extern "C" int mt_lm_forward(mt_lm* h, const int64_t* ids, const uint8_t* attn_mask, long long mask_ld, int B, int T,
                             int past_len, int last_only, void* text_logits, void* audio_logits, mt_stream s) {
  MT_CHECK(h && ids && attn_mask, "mt_lm_forward: null argument");
  MT_CHECK(!h->sess.active, "mt_lm_forward: a sliced generation (mt_lm_stream_begin) owns this handle's KV cache and workspace");
  MT_CHECK(B > 0 && T > 0 && past_len >= 0, "mt_lm_forward: bad shape B=%d T=%d past=%d", B, T, past_len);
  return lm_forward_impl(h, ids, attn_mask, mask_ld, B, T, past_len, nullptr, past_len + T, last_only,
                         reinterpret_cast<bf16*>(text_logits), reinterpret_cast<bf16*>(audio_logits),
                         reinterpret_cast<cudaStream_t>(s));
}

extern "C" int mt_lm_embed(mt_lm* h, const int64_t* ids, long long n_tokens, void* out, mt_stream s) {
  MT_CHECK(h && ids && out, "mt_lm_embed: null argument");
  MT_CHECK(n_tokens > 0 && n_tokens < (1ll << 31), "mt_lm_embed: bad token count %lld", n_tokens);
  const mt_lm_config& c = h->cfg;
  return embed_sum(ids, reinterpret_cast<const bf16*>(h->w.embed_tokens), reinterpret_cast<const bf16*>(h->w.emb_ext),
                   reinterpret_cast<bf16*>(out), static_cast<int>(n_tokens), c.hidden_size, c.n_vq, c.audio_head_size, c.vocab_size,
                   reinterpret_cast<cudaStream_t>(s), 1 + c.n_vq);
}

static TokenIds token_ids(const mt_lm_config& c) {
  TokenIds t;
  t.pad = c.pad_token_id; t.im_start = c.im_start_token_id; t.im_end = c.im_end_token_id;
  t.audio_start = c.audio_start_token_id; t.audio_end = c.audio_end_token_id;
  t.gen_slot = c.audio_assistant_gen_slot_token_id; t.delay_slot = c.audio_assistant_delay_slot_token_id;
  t.audio_pad_code = c.audio_pad_code;
  return t;
}

// mt_gen_params -> SampleCfg (shared by mt_lm_generate and the sliced mt_lm_stream_* form)
static int make_sample_cfg(const mt_lm_config& c, const mt_gen_params* p, SampleCfg& sc) {
  sc.text_do_sample = p->text_temperature > 0.f; sc.audio_do_sample = p->audio_temperature > 0.f;
  sc.text_temperature = sc.text_do_sample ? p->text_temperature : 1.f;
  sc.audio_temperature = sc.audio_do_sample ? p->audio_temperature : 1.f;
  sc.text_top_p = p->text_top_p; sc.text_top_k = p->text_top_k;
  sc.audio_top_p = p->audio_top_p; sc.audio_top_k = p->audio_top_k;
  sc.penalty = p->audio_repetition_penalty; sc.seed = p->seed;
  // audio_top_k = None / <= 0 (no top-k filter, inference_utils.py:135): the pad code of every audio head is -inf
  // (modeling_moss_tts.py:295-300, 486-487), so keeping the audio_head_size - 1 real codes IS "no filter"
  if (sc.audio_do_sample && sc.audio_top_k <= 0 && c.audio_head_size - 1 <= kMaxTopK) sc.audio_top_k = c.audio_head_size - 1;
  // text_top_k = None / <= 0 / > 1024: the full-vocabulary counting-sort sampler (sampling_full.cuh)
  if (sc.audio_do_sample && (sc.audio_top_k <= 0 || sc.audio_top_k > kMaxTopK))
    return fail(MT_ERR_UNSUPPORTED, "audio sampling requires 1 <= top_k <= %d", kMaxTopK);
  return MT_OK;
}

extern "C" int mt_lm_generate(mt_lm* h, const int64_t* input_ids, const uint8_t* attention_mask, int B, int S0,
                              const mt_gen_params* p, int64_t* gen_ids, int64_t* start_indices, int* n_steps,
                              float* step_ms, mt_stream stream) {
  MT_CHECK(h && input_ids && p && gen_ids && start_indices && n_steps, "mt_lm_generate: null argument");
  MT_CHECK(!h->sess.active, "mt_lm_generate: a sliced generation (mt_lm_stream_begin) owns this handle's KV cache and workspace");
  const mt_lm_config& c = h->cfg;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int max_len = S0 + p->max_new_tokens;
  MT_CHECK(B > 0 && B <= c.max_batch, "generate: batch %d exceeds max_batch %d", B, c.max_batch);
  MT_CHECK(S0 > 0 && p->max_new_tokens >= 0 && max_len <= c.max_seq, "generate: %d positions exceed max_seq %d", max_len, c.max_seq);
  MT_CHECK(B * S0 <= c.max_tokens, "generate: prompt of %d tokens exceeds max_tokens %d", B * S0, c.max_tokens);
  SampleCfg sc;
  MT_TRY(make_sample_cfg(c, p, sc));
  *n_steps = 0;
  if (p->max_new_tokens == 0) {
    MT_CUDA(cudaMemcpyAsync(gen_ids, input_ids, sizeof(int64_t) * B * S0 * (1 + c.n_vq), cudaMemcpyDeviceToDevice, s));
  }
  const TokenIds tk = token_ids(c);
  GenState st = gen_state_carve(h->gen_state_buf, B, 1 + c.n_vq, max_len, c.audio_head_size, gen_ids);
  st.full_hist = h->full_hist;
  MT_TRY(gen_init(st, input_ids, attention_mask, S0, tk, start_indices, s));
  if (p->max_new_tokens == 0) return MT_OK;

  // per-step timing events (record_step_times only); destroyed on every exit path
  struct Events {
    std::vector<cudaEvent_t> v;
    ~Events() { for (auto e : v) if (e) cudaEventDestroy(e); }
  } evs;
  std::vector<cudaEvent_t>& ev = evs.v;
  if (step_ms) {
    ev.assign(p->max_new_tokens + 1, nullptr);
    for (auto& e : ev) MT_CUDA(cudaEventCreate(&e));
    MT_CUDA(cudaEventRecord(ev[0], s));
  }
  // time_step 0: prefill over the whole prompt
  MT_TRY(lm_forward_impl(h, input_ids, st.mask, max_len, B, S0, 0, nullptr, S0, 1, h->text_logits, h->audio_logits, s));
  MT_TRY(gen_sample_step(st, h->text_logits, c.vocab_size, h->audio_logits, c.audio_head_size, sc, tk, s));
  if (step_ms) MT_CUDA(cudaEventRecord(ev[1], s));

  auto decode_step = [&](cudaStream_t cs) -> int {
    // position of the token appended by the previous commit = *len - 1; kernels read it from st.pos
    MT_TRY(lm_forward_impl(h, st.next_ids, st.mask, max_len, B, 1, 0, st.pos, max_len, 1, h->text_logits,
                           h->audio_logits, cs, st.text_full));
    return gen_sample_step(st, h->text_logits, c.vocab_size, h->audio_logits, c.audio_head_size, sc, tk, cs);
  };

  int steps_launched = 1;
  // Stop polling without draining the stream: the all-stopped flag of check c is copied to a pinned slot behind an
  // event, and the host waits for check c-1 only after check c (and the steps before it) have been enqueued, so the GPU
  // always has check_every .. 2*check_every steps queued.  (A blocking poll empties the queue and exposes the launch
  // latency of a ~300-node graph, ~0.4 ms, at every poll.)  Steps that run after every row has stopped are no-ops for
  // the state machine; the step count comes from the device at the end.
  int* done_host = h->done_host;
  done_host[0] = done_host[1] = 0;
  cudaEvent_t* poll_ev = h->poll_ev;
  int polls = 0;
  const int check_every = p->check_stop_every > 0 ? p->check_stop_every : 8;
  cudaGraphExec_t exec = nullptr;
  int status = MT_OK;
  for (int step = 1; step < p->max_new_tokens && status == MT_OK; ++step) {
    if (p->use_cuda_graph && step >= 2) {
      if (!exec) {
        // capture one decode step (step 1 ran eagerly so every lazy attribute/JIT path is warm)
        cudaGraph_t graph = nullptr;
        if (!h->cap_stream && cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking) != cudaSuccess) { status = fail(MT_ERR_CUDA, "capture stream creation failed"); break; }
        cudaError_t be = cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal);
        if (be != cudaSuccess) { status = fail(MT_ERR_CUDA, "begin capture failed: %s", cudaGetErrorString(be)); break; }
        const long long before = g_launches;
        status = decode_step(h->cap_stream);
        cudaError_t ce = cudaStreamEndCapture(h->cap_stream, &graph);
        if (status == MT_OK && ce != cudaSuccess) status = fail(MT_ERR_CUDA, "end capture: %s", cudaGetErrorString(ce));
        if (status == MT_OK && cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) status = fail(MT_ERR_CUDA, "graph instantiate failed");
        if (graph) cudaGraphDestroy(graph);
        h->graph_key = g_launches - before;    // kernels per replay
        g_launches = before;
        if (status != MT_OK) break;
      }
      if (cudaGraphLaunch(exec, s) != cudaSuccess) { status = fail(MT_ERR_CUDA, "graph launch failed"); break; }
      g_launches += h->graph_key;
    } else {
      status = decode_step(s);
      if (status != MT_OK) break;
    }
    ++steps_launched;
    if (step_ms) cudaEventRecord(ev[steps_launched], s);
    if (step % check_every == 0) {
      const int slot = polls & 1;
      cudaMemcpyAsync(done_host + slot, st.done, sizeof(int), cudaMemcpyDeviceToHost, s);
      cudaEventRecord(poll_ev[slot], s);
      if (polls > 0) {
        if (cudaEventSynchronize(poll_ev[slot ^ 1]) != cudaSuccess) { status = fail(MT_ERR_CUDA, "decode step failed: %s", cudaGetErrorString(cudaGetLastError())); break; }
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
    for (int i = 0; i < steps_launched; ++i) cudaEventElapsedTime(&step_ms[i], ev[i], ev[i + 1]);
  }
  if (exec) cudaGraphExecDestroy(exec);
  return status;
}

// ---------------------------------------------------------------------------------------------------------
// Sliced form of generate(): the same prefill / decode-step / sampler sequence as mt_lm_generate, enqueued by the
// caller a few steps at a time and never synchronised here, so that a second stream can de-delay and decode the
// frames that are already final (codec streaming, mt_codec_stream_push) while later steps run.
// ---------------------------------------------------------------------------------------------------------
extern "C" int mt_lm_stream_begin(mt_lm* h, const int64_t* input_ids, const uint8_t* attention_mask, int B, int S0,
                                  const mt_gen_params* p, int64_t* gen_ids, int64_t* start_indices, mt_stream stream) {
  MT_CHECK(h && input_ids && p && gen_ids && start_indices, "mt_lm_stream_begin: null argument");
  MT_CHECK(!h->sess.active, "mt_lm_stream_begin: a sliced generation is already open on this handle");
  const mt_lm_config& c = h->cfg;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int max_len = S0 + p->max_new_tokens;
  MT_CHECK(B > 0 && B <= c.max_batch, "generate: batch %d exceeds max_batch %d", B, c.max_batch);
  MT_CHECK(S0 > 0 && p->max_new_tokens >= 1 && max_len <= c.max_seq, "generate: %d positions exceed max_seq %d", max_len, c.max_seq);
  MT_CHECK(B * S0 <= c.max_tokens, "generate: prompt of %d tokens exceeds max_tokens %d", B * S0, c.max_tokens);
  mt_lm::Session& ss = h->sess;
  MT_TRY(make_sample_cfg(c, p, ss.sc));
  ss.tk = token_ids(c);
  ss.st = gen_state_carve(h->gen_state_buf, B, 1 + c.n_vq, max_len, c.audio_head_size, gen_ids);
  ss.st.full_hist = h->full_hist;
  ss.B = B; ss.S0 = S0; ss.max_len = max_len; ss.max_new = p->max_new_tokens; ss.use_graph = p->use_cuda_graph;
  MT_TRY(gen_init(ss.st, input_ids, attention_mask, S0, ss.tk, start_indices, s));
  MT_TRY(lm_forward_impl(h, input_ids, ss.st.mask, max_len, B, S0, 0, nullptr, S0, 1, h->text_logits, h->audio_logits, s));
  MT_TRY(gen_sample_step(ss.st, h->text_logits, c.vocab_size, h->audio_logits, c.audio_head_size, ss.sc, ss.tk, s));
  ss.launched = 1;
  ss.active = true;
  return MT_OK;
}

extern "C" int mt_lm_stream_steps(mt_lm* h, int n, int* status_dev, int* launched, mt_stream stream) {
  MT_CHECK(h && h->sess.active, "mt_lm_stream_steps: no open sliced generation");
  mt_lm::Session& ss = h->sess;
  const mt_lm_config& c = h->cfg;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  auto decode_step = [&](cudaStream_t cs) -> int {
    MT_TRY(lm_forward_impl(h, ss.st.next_ids, ss.st.mask, ss.max_len, ss.B, 1, 0, ss.st.pos, ss.max_len, 1, h->text_logits,
                           h->audio_logits, cs, ss.st.text_full));
    return gen_sample_step(ss.st, h->text_logits, c.vocab_size, h->audio_logits, c.audio_head_size, ss.sc, ss.tk, cs);
  };
  for (int i = 0; i < n && ss.launched < ss.max_new; ++i) {
    if (ss.use_graph && ss.launched >= 2) {
      if (!ss.exec) {
        cudaGraph_t graph = nullptr;
        if (!h->cap_stream) MT_CUDA(cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
        MT_CUDA(cudaStreamBeginCapture(h->cap_stream, cudaStreamCaptureModeThreadLocal));
        const long long before = g_launches;
        int status = decode_step(h->cap_stream);
        cudaError_t ce = cudaStreamEndCapture(h->cap_stream, &graph);
        if (status == MT_OK && ce != cudaSuccess) status = fail(MT_ERR_CUDA, "end capture: %s", cudaGetErrorString(ce));
        if (status == MT_OK && cudaGraphInstantiate(&ss.exec, graph, 0) != cudaSuccess) status = fail(MT_ERR_CUDA, "graph instantiate failed");
        if (graph) cudaGraphDestroy(graph);
        ss.graph_launches = g_launches - before;
        g_launches = before;
        MT_TRY(status);
      }
      MT_CUDA(cudaGraphLaunch(ss.exec, s));
      g_launches += ss.graph_launches;
    } else {
      MT_TRY(decode_step(s));
    }
    ++ss.launched;
  }
  if (status_dev) {
    MT_CUDA(cudaMemcpyAsync(status_dev, ss.st.step, sizeof(int), cudaMemcpyDeviceToDevice, s));
    MT_CUDA(cudaMemcpyAsync(status_dev + 1, ss.st.done, sizeof(int), cudaMemcpyDeviceToDevice, s));
  }
  if (launched) *launched = ss.launched;
  return MT_OK;
}

extern "C" int mt_lm_stream_end(mt_lm* h) {
  MT_CHECK(h, "mt_lm_stream_end: null handle");
  if (h->sess.exec) { cudaGraphExecDestroy(h->sess.exec); h->sess.exec = nullptr; }
  h->sess.active = false;
  return MT_OK;
}
