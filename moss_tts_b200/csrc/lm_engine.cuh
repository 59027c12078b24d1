// Internal interface of the LM engine (lm_engine.cu), shared with the MossTTSLocal engine (local_engine.cu).
#pragma once
#include "lm_ops.cuh"
#include "sampling.cuh"

#include <vector>

struct mt_lm {
  mt_lm_config cfg;
  mt_lm_weights w;
  std::vector<const void*> w_qkv, w_o, w_gu, w_down, ln_in, ln_post, q_norm, k_norm;
  std::vector<CUtensorMap> tm_qkv, tm_o, tm_gu, tm_down;
  CUtensorMap tm_text, tm_audio;
  // workspace views
  mt::bf16 *x, *xn, *qkv, *q, *attn, *gu, *act, *delta, *last_h, *text_logits, *audio_logits;
  float* splitk;
  uint32_t* full_hist;                 // [max_batch][65536] scratch of the full-vocabulary text sampler
  void* gen_state_buf;
  mt::bf16 *k_cache, *v_cache;   // [L][B][kv][S][D] each
  size_t ws_bytes;
  // graph
  cudaGraphExec_t graph_exec = nullptr;
  cudaStream_t cap_stream = nullptr;   // capture-only stream (the caller's stream may be the legacy default stream)
  int graph_B = 0; long long graph_key = 0;
  // stop polling (mt_lm_generate): pinned result slots and their events live as long as the handle
  int* done_host = nullptr;            // 4 ints, cudaMallocHost at create
  cudaEvent_t poll_ev[2] = {nullptr, nullptr};
  // sliced generation (mt_lm_stream_*): the loop state of one generate() call kept between calls
  struct Session {
    bool active = false;
    mt::SampleCfg sc; mt::TokenIds tk; mt::GenState st;
    int B = 0, S0 = 0, max_len = 0, max_new = 0, launched = 0, use_graph = 0;
    cudaGraphExec_t exec = nullptr;
    long long graph_launches = 0;
  } sess;
};

namespace mt {

constexpr long long kSplitKBytes = 96ll << 20;

// One Linear on the LM path.  Small M: split-K partials handed to the consumer kernel (returns the split
// count through *splits, partials in h->splitk); large M or splits == nullptr: finished bf16 output in `out`.
int lm_linear(mt_lm* h, const bf16* x, int M, int K, const void* w, const CUtensorMap* tm, int N, bf16* out,
              int* splits, cudaStream_t s, const int* tile_filter_flag = nullptr, int tile_lo = 0, int tile_hi = 0);

// Backbone forward over B*T tokens.  Leaves the final-norm hidden states in h->xn ([B*T, H]; with last_only
// and T > 1 the last position of every row is also gathered into h->last_h).  Heads are skipped when both
// logit pointers are null.  n_vq_embed: audio channels summed into the input embedding (-1 = all).
int lm_forward_impl(mt_lm* h, const int64_t* ids, const uint8_t* mask, long long mask_ld, int B, int T,
                    int past_len, const int* past_len_dev, int max_keys, int last_only, bf16* text_logits,
                    bf16* audio_logits, cudaStream_t s, const int* text_full_flag = nullptr, int n_vq_embed = -1);

}  // namespace mt
