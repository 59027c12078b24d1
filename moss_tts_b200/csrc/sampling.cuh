// Fused per-step sampler + delay-pattern state machine (sampling.cu).
#pragma once
#include "common.cuh"
#include "gemm.cuh"

namespace mt {

constexpr int kMaxTopK = 1024;

// Device-resident generation state of one generate() call (reference modeling_moss_tts.py:417-440).
struct GenState {
  // scalars (device)
  int* len;          // filled sequence positions (prompt + generated) == past_len of the next forward
  int* step;         // time_step
  int* done;         // all rows stopped
  int* pos;          // index of the row appended by the last commit (= KV position of the next decode input)
  int* text_full;    // 1 if some row of the NEXT step reads its whole text-logit row (not forced, not inside audio);
                     // 0 lets the text head compute only the gen/delay-slot logits (modeling_moss_tts.py:459-460)
  // per row
  uint8_t* stopping;
  uint8_t* is_audio;
  int* audio_len;
  int* delayed;      // -1 encodes the reference's INT64_MAX sentinel
  // token buffers
  int64_t* next_ids; // [B, C] tokens chosen this step == input of the next forward
  int64_t* gen_ids;  // [B, max_len, C]
  uint8_t* mask;     // [B, max_len]
  uint32_t* seen0;   // repetition-penalty bitmaps over audio ids: channel 1 / channels 2..n_vq
  uint32_t* seen_rest;
  uint32_t* full_hist;   // [B][65536] counting-sort scratch of the full-vocabulary text sampler (sampling_full.cuh), or null
  int B, C, max_len, bitmap_words;
};

struct SampleCfg {
  float text_temperature, text_top_p; int text_top_k, text_do_sample;
  float audio_temperature, audio_top_p; int audio_top_k, audio_do_sample;
  float penalty;
  unsigned seed;
};

struct TokenIds {
  int pad, im_start, im_end, audio_start, audio_end, gen_slot, delay_slot, audio_pad_code;
};

size_t gen_state_bytes(int B, int C, int max_len, int audio_head);
// carve `buf` (device) into a GenState; gen_ids is caller-provided
GenState gen_state_carve(void* buf, int B, int C, int max_len, int audio_head, int64_t* gen_ids);

int gen_init(const GenState& st, const int64_t* input_ids, const uint8_t* attn_mask, int S0, const TokenIds& tk,
             int64_t* start_indices, cudaStream_t s);
// text_logits [B, vocab] bf16, audio_logits [B, n_vq, audio_head] bf16 (last position of this step)
int gen_sample_step(const GenState& st, const bf16* text_logits, int vocab, const bf16* audio_logits, int audio_head,
                    const SampleCfg& sc, const TokenIds& tk, cudaStream_t s);

// Device-resident state of one MossTTSLocal generate() call (reference moss_tts_local/modeling_moss_tts.py:343-352).
struct LocalState {
  int *len, *step, *done, *pos;   // as in GenState
  uint8_t* unfinished;            // [B] unfinished_sequences (:350)
  int64_t* next_ids;              // [B, C] tokens of the frame being built == input of the next global forward
  int64_t* gen_ids;               // [B, max_len, C]
  uint8_t* mask;                  // [B, max_len] all ones (HF extends the attention mask with ones)
  uint8_t* ones;                  // [C] all ones: the depth transformer attends to its whole causal prefix
  uint32_t* seen;                 // [B, C-1, words] per-row, per-channel id sets for RepetitionPenaltyLogitsProcessor
  int B, C, max_len, bitmap_words;
};
size_t local_state_bytes(int B, int C, int max_len, int audio_head);
LocalState local_state_carve(void* buf, int B, int C, int max_len, int audio_head, int64_t* gen_ids);
int local_init(const LocalState& st, const int64_t* input_ids, int S0, cudaStream_t s, const uint8_t* attn_mask = nullptr);
int local_sample(const LocalState& st, const bf16* logits, int V, int c, int do_sample, float temperature, float penalty,
                 float top_p, int top_k, unsigned seed, cudaStream_t s, uint32_t* full_hist = nullptr);
int local_commit(const LocalState& st, int eos, int pad_code, int n_channels_sampled, cudaStream_t s);

// Device-resident state of one MOSS-TTS-Realtime session (reference streaming_mossttsrealtime.py:63-86: attention_mask,
// _generated_tokens, _is_stopping, _last_audio_tokens, _step_idx).
struct RtState {
  int *len, *step, *done, *pos;   // mask columns filled / frames generated (_step_idx) / all rows stopping / KV row of the step
  uint8_t* stopping;              // [B] _is_stopping
  int64_t* text_in;               // [B] text token of the next step (text_pad when the caller has none)
  int64_t* step_ids;              // [B, 1+R] backbone input of a step: text token | previous frame (:244-246)
  int64_t* last;                  // [B, R] newest frame (_last_audio_tokens); the depth loop writes it channel by channel
  int64_t* history;               // [B, max_frames, R] _generated_tokens
  uint8_t* mask;                  // [B, max_seq] attention_mask (a stopped row's later columns are 0, :247)
  uint32_t* seen;                 // [B, R, words] ids inside the repetition window of each (row, channel)
  int B, R, max_frames, max_seq, words;
};
size_t rt_state_bytes(int B, int R, int max_frames, int max_seq, int audio_head);
RtState rt_state_carve(void* buf, int B, int R, int max_frames, int max_seq, int audio_head);
// prefill: copies attn_mask [B,S0] behind the `past` columns already filled, resets the generation state
int rt_prefill_init(const RtState& st, const uint8_t* attn_mask, int S0, int past, cudaStream_t s);
int rt_set_text(const RtState& st, const int64_t* text_tokens, long long text_pad, cudaStream_t s);
int rt_begin_step(const RtState& st, cudaStream_t s);
// id sets of the last `window` frames (window <= 0: all) of history [B, ld_frames, R] before frame `step`
int rt_window_sets(const RtState& st, const int64_t* history, long long ld_frames, const int* step_dev, int step_host,
                   int window, cudaStream_t s);
// channel c: apply_repetition_penalty (:359-375, also when greedy) -> sample_token (:377-397) -> out[b*R + c]
int rt_sample(const RtState& st, const bf16* logits, int V, int c, int do_sample, float temperature, float penalty,
              float top_p, int top_k, unsigned seed, const int* step_dev, int step_host, int64_t* out, cudaStream_t s);
int rt_commit(const RtState& st, int eos, int advance_len, int64_t* tokens_out, cudaStream_t s);

// scratch: N * sample_full_scratch_bytes() bytes, needed only when do_sample and (top_k <= 0 or top_k > kMaxTopK)
size_t sample_full_scratch_bytes();
int sample_tokens(const bf16* logits, long long ld, int N, int V, float temperature, const uint32_t* prev_bitmap,
                  float penalty, float top_p, int top_k, int do_sample, const float* uniforms, int64_t* out,
                  cudaStream_t s, void* scratch = nullptr, long long scratch_bytes = 0);
int find_last_equal(const int64_t* t, int B, int S, long long ld, long long value, int64_t* out, cudaStream_t s);
int apply_delay_pattern(const int64_t* codes, int T, int n, long long pad, int64_t* out, cudaStream_t s);
int apply_de_delay_pattern(const int64_t* delayed, int Td, int n, int64_t* out, cudaStream_t s);

}  // namespace mt
