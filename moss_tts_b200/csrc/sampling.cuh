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
int local_init(const LocalState& st, const int64_t* input_ids, int S0, cudaStream_t s);
int local_sample(const LocalState& st, const bf16* logits, int V, int c, int do_sample, float temperature, float penalty,
                 float top_p, int top_k, unsigned seed, cudaStream_t s);
int local_commit(const LocalState& st, int eos, int pad_code, int n_channels_sampled, cudaStream_t s);

int sample_tokens(const bf16* logits, long long ld, int N, int V, float temperature, const uint32_t* prev_bitmap,
                  float penalty, float top_p, int top_k, int do_sample, const float* uniforms, int64_t* out,
                  cudaStream_t s);
int find_last_equal(const int64_t* t, int B, int S, long long ld, long long value, int64_t* out, cudaStream_t s);
int apply_delay_pattern(const int64_t* codes, int T, int n, long long pad, int64_t* out, cudaStream_t s);
int apply_de_delay_pattern(const int64_t* delayed, int Td, int n, int64_t* out, cudaStream_t s);

}  // namespace mt
