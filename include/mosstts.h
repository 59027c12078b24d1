/*
 * libmosstts_sm100 — C ABI of the B200-native MOSS-TTS synthesis hot path.
 *
 * The reference (xiami2019/MOSS-TTS) is pure Python/PyTorch and has no FFI of its own
 * (SURVEY.md §2.2, §8b): its boundary is the Python method surface
 *     MossTTSDelayModel.generate / forward      moss_tts_delay/modeling_moss_tts.py:392-525, :225-390
 *     inference_utils.sample_token, find_last_equal_C   moss_tts_delay/inference_utils.py:111-165
 *     MossTTSDelayProcessor.apply_(de_)delay_pattern, decode_audio_codes
 *                                                moss_tts_delay/processing_moss_tts.py:516-537, :880-930
 *     MossAudioTokenizerModel.decode / batch_decode / quantizer.decode_codes
 *                                                moss_audio_tokenizer/modeling_moss_audio_tokenizer.py:1650-1755,
 *                                                :1431-1475, :1276-1283
 * Each entry point below names the reference interface it stands behind.  The Python host in
 * moss_tts_b200/ keeps those method names/arguments and calls this library through ctypes.
 *
 * Conventions
 *   - plain C types only: device pointers as void*, sizes as int / long long, cudaStream_t as void*.
 *   - every buffer is allocated by the caller (PyTorch); the library borrows and never frees it.
 *     Opaque handles own only small host-side plans (TMA descriptors, CUDA graphs).
 *   - return value: MT_OK (0) or a negative MT_ERR_*; mt_last_error() gives a message
 *     (thread-local).  No exceptions cross this boundary.
 *   - all entry points are asynchronous with respect to the host on the given stream unless noted.
 *   - there is NO CPU fallback: without an sm_100 device every compute entry point fails.
 */
#ifndef MOSSTTS_H_
#define MOSSTTS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define MT_OK 0
#define MT_ERR_ARG (-1)
#define MT_ERR_CUDA (-2)
#define MT_ERR_UNSUPPORTED (-3)

#define MT_DTYPE_BF16 0
#define MT_DTYPE_F32 1 /* as a GEMM input type: fp32 storage, TF32 tensor-core math */
#define MT_DTYPE_F16 2 /* IEEE half storage (x and w), fp32 accumulate: the TF32 mantissa at half the operand bytes */

#define MT_GEMM_AUTO 0   /* tcgen05/TMA kernel (the product path) */
#define MT_GEMM_SIMT 1   /* plain CUDA-core kernel, used by tests to cross-check the tcgen05 kernel */
/* OR-able flag for mt_gemm_args.impl: `w` is constant for the lifetime of the stream work in flight
 * (model weights), so the kernel may fetch weight tiles BEFORE the preceding kernel on the stream has
 * finished (programmatic dependent launch); x / residual / out are still ordered normally. */
#define MT_GEMM_W_STATIC 2
/* OR-able flag: `w` is stored tile-major, [ceil(N/128)][K/ke][128][ke] with ke = 64 (bf16) / 32 (fp32):
 * every 128-row x 128-byte operand tile is one contiguous 16 KB run (rows past N are zero padding).
 * This is the HBM layout the decode GEMMs stream fastest; ldw is ignored. */
#define MT_GEMM_W_TILED 4

typedef void* mt_stream; /* cudaStream_t */

int mt_version(void);
const char* mt_last_error(void);
/* Number of kernel launches issued by this library since load (bench.py's gpu_launches). */
long long mt_launch_count(void);
/* Debug knob for tests: force every internal GEMM to MT_GEMM_SIMT (1) or back to tcgen05 (0). */
int mt_debug_force_simt(int on);
/* Debug knob: launch kernels with (1, default) or without (0) the programmatic-dependent-launch
 * attribute.  Results are bit-identical either way; only kernel overlap changes. */
int mt_debug_set_pdl(int on);
/* Debug knob: CTAs per (batch row, kv head) of the fused decode attention: 0 = automatic (1 when B * n_kv fills the
 * GPU, else a thread-block cluster of 2/4/8 CTAs that split the keys), 1/2/4/8 = forced.  Any other value: MT_ERR_ARG. */
int mt_debug_set_attn_cluster(int cs);
/* Debug timeline: every decode-GEMM launch issued (or captured into a graph) after this call takes the next
 * 148*8 uint64 slots of the device buffer `buf` and records 8 %globaltimer stamps per CTA (scripts/trace_decode.py).
 * NULL switches it off.  Returns the number of slots handed out since the previous call. */
long long mt_debug_set_trace(void* buf, long long words);
/* Debug probe for the fp16-operand codec path (layer_gemm_f16): `slots` = MT_CODEC_MAX_STAGES x 4 device floats (zeroed by the
 * caller).  Every codec pass launched afterwards folds max |value| of each stage's half operands into
 * slots[stage][0..3] = {LayerNorm output, q|k|v, attention output, GELU output} (NaN if any NaN).  NULL switches it off.
 * Used by the tests to prove the operands stay inside the fp16 range (65504) through all 68 released layers. */
int mt_codec_debug_set_absmax(float* slots);
/* Attention kernel of the fp16-operand codec path: 0 (default) = tcgen05 / TMEM / TMA FlashAttention (codec_attn_tc5.cu),
 * 1 = the mma.sync FlashAttention-2 kernel (codec_attn.cu).  For A/B measurements and cross-checks; process-wide. */
int mt_codec_debug_set_attention(int mode);
/* Token-major GEMM kernels (M > 256: prefill, every codec Linear).  mode 0 (default) = chosen by shape: CTA pairs (256 x 256
 * tiles on tcgen05.mma.cta_group::2, gemm_pair.cu) for K >= 2048, the one-SM 128 x 256 kernel (gemm_big.cu) for the codec's
 * short-K Linears; 1 = pairs wherever M >= 256; 4 = never pairs; + 2 = the per-lane row-store epilogue in place of the
 * warp-transposed (coalesced) one; + 16 = the codec rotates q and k in a separate pass instead of the in_proj epilogue;
 * + 32 = 8 instead of 16 epilogue warps for the GELU epilogue.  The
 * GEMM kernels and epilogue forms give bit-identical results.  For A/B measurements and cross-checks;
 * process-wide. */
int mt_gemm_debug_set_pair(int mode);

/* ------------------------------------------------------------------------------------------
 * Linear layer:  out[M,N] = epilogue( x[M,K] * w[N,K]^T )       (every nn.Linear on the path:
 * Qwen3 q/k/v/o/gate/up/down, the 1+n_vq heads modeling_moss_tts.py:185-191,292-300, and the codec's
 * in_proj/out_proj/linear1/linear2/input_proj/output_proj modeling_moss_audio_tokenizer.py:563-568,731-732,907-914)
 * epilogue: v = acc (+bias[n]) ; act ; (round to bf16) ; out = residual[m,n] + (scale ? scale[n]*v : v)
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  int dtype_in;          /* MT_DTYPE_BF16 | MT_DTYPE_F32 | MT_DTYPE_F16 (x and w) */
  int dtype_out;         /* MT_DTYPE_BF16 | MT_DTYPE_F32 | MT_DTYPE_F16 (out and residual) */
  int M, N, K;
  const void* x;  long long ldx;   /* row stride in elements */
  const void* w;  long long ldw;
  void* out;      long long ldo;
  const float* bias;               /* [N] or NULL */
  int act;                         /* 0 none, 1 GELU(erf) */
  const void* residual; long long ldr;  /* [M,N] or NULL */
  const float* scale;              /* [N] or NULL (LayerScale) */
  int round_bf16;                  /* round v to bf16 before the residual add */
  float* workspace; long long workspace_bytes; /* split-K partials; may be NULL (no split-K) */
  int impl;                        /* MT_GEMM_AUTO | MT_GEMM_SIMT */
  int* out_splits;                 /* NULL: `out` receives the finished tensor (a split-K GEMM runs its own reduce kernel).
                                    * non-NULL (host int): the way the engines call it - the kernel leaves fp32 partials
                                    * [splits][M][N] in `workspace`, *out_splits = splits (>= 1), `out` is not written and
                                    * the consumer kernel (add_rmsnorm / silu_mul / attention) sums the slabs. */
} mt_gemm_args;
int mt_gemm(const mt_gemm_args* a, mt_stream s);

/* mt_gemm for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arithmetic, kernels and result as mt_gemm called with a workspace of MT_GEMM_STATIC_SCRATCH_BYTES, but the split-K
 * scratch (fp32 partial slabs) is a file-scope `static __device__` array inside the library (csrc/gemm_static.cu) that every
 * GPU loading the library holds until the process exits; this entry point calls no cudaMalloc / cudaFree / allocating routine.
 *   - a->workspace and a->out_splits must be NULL; `out` always receives the finished tensor.
 *   - limits: M <= 256 (only the decode-orientation GEMM uses scratch) and 2 * M * N * 4 <= MT_GEMM_STATIC_SCRATCH_BYTES
 *     (room for a two-way split); anything larger returns MT_ERR_ARG with a message naming the limit: use mt_gemm.
 *   - every call on a device shares the one array, which keeps the previous call's contents; nothing needs clearing
 *     (each partial is written before it is read within a call).  Calls that OVERLAP on the GPU (other streams,
 *     concurrently replayed graphs) RACE on it: serialise them on one stream or with events. */
#define MT_GEMM_STATIC_SCRATCH_BYTES (8 * 1024 * 1024)
int mt_gemm_static_scratch(const mt_gemm_args* a, mt_stream s);

/* ------------------------------------------------------------------------------------------
 * MossTTSDelay language model
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  int hidden_size, intermediate_size, num_layers, num_heads, num_kv_heads, head_dim;
  int vocab_size, n_vq, audio_head_size /* audio_vocab_size+1 */;
  float rms_norm_eps;
  int pad_token_id, im_start_token_id, im_end_token_id;
  int audio_start_token_id, audio_end_token_id;
  int audio_assistant_gen_slot_token_id, audio_assistant_delay_slot_token_id;
  int audio_pad_code;
  int max_batch, max_seq;       /* KV-cache capacity */
  int max_tokens;               /* largest B*T a single forward call may see */
  int weights_tiled;            /* 1: every Linear weight (w_qkv/w_o/w_gate_up/w_down/heads) is tile-major, see MT_GEMM_W_TILED */
} mt_lm_config;

/* All weights bf16, row-major [out,in] as in nn.Linear / nn.Embedding. */
typedef struct {
  const void* embed_tokens;        /* [vocab, H] */
  const void* emb_ext;             /* [n_vq, audio_head_size, H] */
  const void* const* w_qkv;        /* per layer: rows = q | k | v, [q+2kv, H] */
  const void* const* w_o;          /* per layer [H, q] */
  const void* const* w_gate_up;    /* per layer: rows = gate | up, [2I, H] */
  const void* const* w_down;       /* per layer [H, I] */
  const void* const* ln_in;        /* per layer [H] */
  const void* const* ln_post;      /* per layer [H] */
  const void* const* q_norm;       /* per layer [head_dim] */
  const void* const* k_norm;       /* per layer [head_dim] */
  const void* final_norm;          /* [H] */
  const void* text_head;           /* [vocab, H] */
  const void* audio_heads;         /* [n_vq*audio_head_size, H] */
  const void* rope_cos;            /* [max_seq, head_dim/2] bf16 (host builds it with HF's formula) */
  const void* rope_sin;
} mt_lm_weights;

typedef struct mt_lm mt_lm;

int mt_lm_query_sizes(const mt_lm_config* cfg, long long* kv_bytes, long long* workspace_bytes);
int mt_lm_create(const mt_lm_config* cfg, const mt_lm_weights* w, void* kv_cache, void* workspace, mt_lm** out);
int mt_lm_destroy(mt_lm* h);

/* MossTTSDelayModel.forward (modeling_moss_tts.py:225-300) without the loss branch.
 * ids [B,T,1+n_vq] int64; attn_mask [B, mask_ld] uint8 (1 = attend) covering positions [0,past_len+T);
 * writes K/V for positions [past_len,past_len+T).  Logit outputs are bf16 and may be NULL:
 *   text_logits  [B, Tout, vocab]                  audio_logits [B, Tout, n_vq, audio_head_size]
 * with Tout = T (last_only=0) or 1 (last_only=1: only the last position, what generate() consumes).
 * audio_logits[..., audio_head_size-1] is set to -inf as the reference does (:298-299). */
int mt_lm_forward(mt_lm* h, const int64_t* ids, const uint8_t* attn_mask, long long mask_ld,
                  int B, int T, int past_len, int last_only,
                  void* text_logits, void* audio_logits, mt_stream s);
/* MossTTSDelayModel.get_input_embeddings (modeling_moss_tts.py:196-213): out [n_tokens, H] bf16 =
 * E_text[ids[t,0]] + sum_i emb_ext[i][ids[t,1+i]], added left to right with a bf16 rounding after every add.
 * ids [n_tokens, 1+n_vq] int64 (device).  Does not touch the KV cache or the workspace. */
int mt_lm_embed(mt_lm* h, const int64_t* ids, long long n_tokens, void* out, mt_stream s);

/* generate() (modeling_moss_tts.py:392-525). */
typedef struct {
  int max_new_tokens;
  float text_temperature, text_top_p;   int text_top_k;
  float audio_temperature, audio_top_p; int audio_top_k;
  float audio_repetition_penalty;
  unsigned int seed;
  int use_cuda_graph;          /* capture the decode step once and replay it */
  int check_stop_every;        /* host polls the all-stopped flag every this many steps (>=1) */
} mt_gen_params;

/* input_ids [B,S0,1+n_vq] int64, attention_mask [B,S0] uint8 (device).
 * gen_ids [B, S0+max_new_tokens, 1+n_vq] int64 (device, caller-allocated) receives prompt + new rows.
 * start_indices [B] int64 (device): last im_start index + 3 (:518).
 * Returns through *n_steps the number of decode iterations executed (host-synchronous at the end). */
int mt_lm_generate(mt_lm* h, const int64_t* input_ids, const uint8_t* attention_mask, int B, int S0,
                   const mt_gen_params* p, int64_t* gen_ids, int64_t* start_indices, int* n_steps,
                   float* step_ms /* optional [max_new_tokens] host array of per-step device ms, or NULL */,
                   mt_stream s);

/* Sliced form of generate() for streaming synthesis (LM decode overlapped with chunked codec
 * decode, the structure of moss_tts_realtime's streaming session, streaming_mossttsrealtime.py:301-357 + :682-799,
 * applied to MossTTSDelay).  mt_lm_stream_begin = everything mt_lm_generate does up to and including time_step 0
 * (prefill + first sample); mt_lm_stream_steps enqueues up to n further decode steps (CUDA-graph replays when
 * use_cuda_graph) and returns WITHOUT synchronising; rows [0, S0 + steps) of gen_ids are final once the work enqueued
 * so far has run.  status_dev (optional, device int32[2]) receives {time steps executed, all rows stopped} as of the
 * end of this slice; *launched (optional, host) = steps enqueued so far including time_step 0.  Steps enqueued after
 * every row has stopped change nothing.  The token sequence is the one mt_lm_generate produces for the same params. */
int mt_lm_stream_begin(mt_lm* h, const int64_t* input_ids, const uint8_t* attention_mask, int B, int S0,
                       const mt_gen_params* p, int64_t* gen_ids, int64_t* start_indices, mt_stream s);
int mt_lm_stream_steps(mt_lm* h, int n, int* status_dev, int* launched, mt_stream s);
int mt_lm_stream_end(mt_lm* h);
/* De-delay (processing_moss_tts.py:528-537) of selected frames straight out of gen_ids [B, max_len, C]:
 * rows [B, T] int32 (device) = generated-row index of the gen-slot row opening each frame (-1 = none -> code 0);
 * out [C-1, B, T] int64 = gen_ids[b, S0 + rows[b,f] + q, 1 + q], the layout mt_codec_stream_push takes. */
int mt_stream_gather_frames(const int64_t* gen_ids, int B, int max_len, int C, int S0, const int* rows, int T,
                            int64_t* out, mt_stream s);

/* inference_utils.sample_token (:111-145) on bf16 logits [N,V] (row stride ld elements).
 * prev_bitmap: optional device bitmap (ceil(V/32) uint32) of ids the repetition penalty applies to
 * (the union the reference builds with torch.unique, :78-86).  uniforms [N] float in [0,1) drive the
 * inverse-CDF draw when do_sample=1.  out [N] int64. */
int mt_sample_tokens(const void* logits, long long ld, int N, int V, float temperature,
                     const uint32_t* prev_bitmap, float repetition_penalty, float top_p, int top_k,
                     int do_sample, const float* uniforms, int64_t* out, mt_stream s);

/* The same call for sampling WITHOUT a top-k filter (top_k <= 0, i.e. the reference's `top_k is None or <= 0`, :135) or with
 * top_k > 1024 over long rows: a counting sort over the 65,536 possible bf16 values replaces the candidate list.
 * scratch: mt_sample_full_scratch_bytes(N) bytes of device memory (256 KB per row), contents irrelevant.
 * Any (top_k, V) mt_sample_tokens accepts is accepted here too (the scratch is then unused). */
long long mt_sample_full_scratch_bytes(int N);
int mt_sample_tokens_full(const void* logits, long long ld, int N, int V, float temperature,
                          const uint32_t* prev_bitmap, float repetition_penalty, float top_p, int top_k,
                          int do_sample, const float* uniforms, int64_t* out, void* scratch, long long scratch_bytes,
                          mt_stream s);

/* inference_utils.find_last_equal_C (:148-165): t [B,S] int64 -> out [B] int64 (-1 if absent). */
int mt_find_last_equal(const int64_t* t, int B, int S, long long ld, long long value, int64_t* out, mt_stream s);

/* processing_moss_tts.py:516-525 / :528-537.  codes [T,n] -> out [T+n-1,n]; delayed [Td,n] -> out [Td-n+1,n]. */
int mt_apply_delay_pattern(const int64_t* codes, int T, int n, long long pad_code, int64_t* out, mt_stream s);
int mt_apply_de_delay_pattern(const int64_t* delayed, int Td, int n, int64_t* out, mt_stream s);

/* ------------------------------------------------------------------------------------------
 * MossTTSLocal (reference moss_tts_local/modeling_moss_tts.py): global Qwen3 backbone + depth transformer.
 * Replaces MossTTSDelayModel(CustomMixin)._sample :316-477 (the body of generate()) and the modules it calls:
 * MosiTTSModel.forward :532-569, MossTTSLocalTransformer :178-292, MossTTSMLP :47-95, MossTTSRMSNorm :34-44.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  mt_lm_config lm;              /* backbone shape, vocab / n_vq / audio_head_size, token ids, capacity, weights_tiled */
  int local_hidden_size, local_ffn_hidden_size, local_num_layers, additional_mlp_ffn_hidden_size;
} mt_local_config;

/* bf16, [out,in] (tile-major when lm.weights_tiled).  The local layers copy the backbone's head layout
 * (num_heads / num_kv_heads / head_dim, :585-590). */
typedef struct {
  mt_lm_weights lm;                /* embed_tokens = embedding_list[0], emb_ext = embedding_list[1..n_vq] stacked
                                      [n_vq, audio_head_size, H]; text_head / audio_heads unused (NULL) */
  const void* const* l_w_qkv;      /* per local layer [q+2kv, LH] */
  const void* const* l_w_o;        /* [LH, q] */
  const void* const* l_w_gate_up;  /* [2*LF, LH] */
  const void* const* l_w_down;     /* [LH, LF] */
  const void* const* l_ln_in;      /* [LH] */
  const void* const* l_ln_post;    /* [LH] */
  const void* const* l_q_norm;     /* [head_dim] */
  const void* const* l_k_norm;
  const void* l_final_norm;        /* [LH] */
  const void* in_gate_up;          /* speech_embedding_to_local_mlp: gate|up [2*AF, H] */
  const void* in_down;             /* [LH, AF] */
  const void* const* out_gate_up;  /* per channel local_to_speech_embedding_mlps[c]: [2*AF, LH] */
  const void* const* out_down;     /* per channel [H, AF] */
  const void* const* head_norm;    /* per channel layer_norm_before_lm_heads[c]: [H] */
  const void* const* heads;        /* per channel lm_heads[c]: [vocab | audio_head_size, H] */
} mt_local_weights;

typedef struct mt_local mt_local;
int mt_local_query_sizes(const mt_local_config* cfg, long long* kv_bytes, long long* workspace_bytes);
int mt_local_create(const mt_local_config* cfg, const mt_local_weights* w, void* kv_cache, void* workspace, mt_local** out);
int mt_local_destroy(mt_local* h);

/* One teacher-forced frame (the body of the :379-450 loop with the sampled tokens replaced by frame_ids):
 * backbone over input_ids [B,S0,1+n_vq], then the depth loop feeding frame_ids [B,1+n_vq] (device int64).
 * logits_out: bf16 [min(1+n_vq, 1+n_vq_for_inference)][B][ldv]; channel c fills its first vocab / audio_head_size
 * columns, audio channels carry -inf at the pad code (:411-412).  Parity tests compare these with the oracle. */
int mt_local_forward(mt_local* h, const int64_t* input_ids, const uint8_t* attention_mask /* [B,S0] or NULL */, int B, int S0,
                     int n_vq_for_inference, const int64_t* frame_ids, void* logits_out, long long ldv, mt_stream s);

/* generation_config of the reference (:359-370): per-channel host arrays of 1+n_vq entries.  Channels with
 * do_sample[c] == 0 take the argmax of the raw logits; the others apply RepetitionPenalty (c > 0 only, over the
 * row's own history of that channel) -> Temperature -> TopK -> TopP and draw by inverse CDF from (seed, step, row, c). */
typedef struct {
  int max_new_tokens;
  int n_vq_for_inference;          /* channels 1..n_vq_for_inference are generated, the rest stay 0 (:397,:427) */
  int eos_token_id;                /* -1: no EOS handling */
  const int* do_sample;
  const float* temperature;        /* 1.0 = none */
  const float* top_p;              /* 1.0 = none */
  const int* top_k;                /* 1..1024 */
  const float* repetition_penalty; /* 1.0 = none */
  unsigned int seed;
  int use_cuda_graph;
  int check_stop_every;
  const uint8_t* attention_mask;   /* [B,S0] uint8 (device) or NULL = all ones.  Left padding is masked out of the keys and
                                      every generated frame is visible, as HF's generation loop extends the mask (:391-395) */
} mt_local_gen_params;

/* input_ids [B,S0,1+n_vq] int64 (device).
 * gen_ids [B, S0+max_new_tokens, 1+n_vq] int64 (device) receives prompt + generated frames; *n_steps = frames made. */
int mt_local_generate(mt_local* h, const int64_t* input_ids, int B, int S0, const mt_local_gen_params* p,
                      int64_t* gen_ids, int* n_steps, float* step_ms, mt_stream s);

/* ------------------------------------------------------------------------------------------
 * MOSS-TTS-Realtime step-wise decode (reference moss_tts_realtime/mossttsrealtime/):
 * a Qwen3 backbone over [text | rvq audio] embedding sums and a Qwen3-shaped local transformer WITH rotary embedding that
 * emits the rvq codes of a frame one after the other from the backbone's last hidden state.
 * Replaces MossTTSRealtimeInference.prefill / step / finish (streaming_mossttsrealtime.py:121-299),
 * generate_local_transformer :301-357, apply_repetition_penalty :359-375, sample_token :377-421 and the modules they
 * call: MossTTSRealtime.forward (modeling_mossttsrealtime.py:101-142), MossTTSRealtimeLocalTransformerForCausalLM
 * (modeling_mossttsrealtime_local.py:258-448).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  mt_lm_config lm;              /* backbone; n_vq = rvq, audio_head_size = audio_vocab_size (1027), capacity, weights_tiled */
  int local_num_layers, local_intermediate_size, local_num_heads, local_num_kv_heads, local_head_dim;
  float local_rms_norm_eps;
  int max_frames;               /* history capacity: frames one session may generate */
  int audio_eos_token;          /* channel-0 code that stops a row (:48) */
  long long text_pad_id;        /* text token fed when the caller passes none (:236-237) */
} mt_rt_config;

/* bf16, [out,in] (tile-major when lm.weights_tiled).  The local transformer shares the backbone's hidden size. */
typedef struct {
  mt_lm_weights lm;                /* embed_tokens = embed_tokens[0], emb_ext = embed_tokens[1..rvq] stacked; no heads */
  const void* const* l_w_qkv;      /* per local layer [lq+2lkv, H] */
  const void* const* l_w_o;        /* [H, lq] */
  const void* const* l_w_gate_up;  /* [2*LI, H] */
  const void* const* l_w_down;     /* [H, LI] */
  const void* const* l_ln_in;      /* [H] */
  const void* const* l_ln_post;    /* [H] */
  const void* const* l_q_norm;     /* [local_head_dim] */
  const void* const* l_k_norm;
  const void* l_final_norm;        /* [H] */
  const void* l_embed;             /* local_transformer.model.embed_tokens stacked [rvq-1, audio_head_size, H] */
  const void* const* l_heads;      /* local_lm_heads[c]: [audio_head_size, H], rvq entries */
  const void* l_rope_cos;          /* [rvq, local_head_dim/2] bf16 */
  const void* l_rope_sin;
} mt_rt_weights;

/* sample_token's arguments (:377) + the penalty of :335-341 */
typedef struct {
  int do_sample;
  float temperature, top_p;
  int top_k;                       /* 1..1024 when sampling */
  float repetition_penalty;        /* 1.0 or 0 = none */
  int repetition_window;           /* <= 0: the whole history */
  unsigned int seed;
} mt_rt_sample_params;

typedef struct mt_rt mt_rt;
int mt_rt_query_sizes(const mt_rt_config* cfg, long long* kv_bytes, long long* workspace_bytes);
int mt_rt_create(const mt_rt_config* cfg, const mt_rt_weights* w, void* kv_cache, void* workspace, mt_rt** out);
int mt_rt_destroy(mt_rt* h);

/* prefill() (:121-218) from the point where the padded batch exists: input_ids [B,S0,1+rvq] int64 and attention_mask
 * [B,S0] uint8 (device).  keep_cache = 0 starts a new conversation; 1 appends the turn behind the positions already in
 * the KV cache (:177-179; host-synchronous: the filled length is read back).  Runs the backbone, then the depth loop with
 * no history (gen_step 0).  audio_tokens [B,rvq] int64 (device) receives the first frame. */
int mt_rt_prefill(mt_rt* h, const int64_t* input_ids, const uint8_t* attention_mask, int B, int S0, int keep_cache,
                  const mt_rt_sample_params* p, int64_t* audio_tokens, mt_stream s);
/* step() (:220-272): text_tokens [B] int64 (device) or NULL (text_pad for every row).  Asynchronous; once every row has
 * stopped a step changes nothing and returns the last frame again (:232-233). */
int mt_rt_step(mt_rt* h, const int64_t* text_tokens, const mt_rt_sample_params* p, int use_cuda_graph,
               int64_t* audio_tokens, mt_stream s);
/* finish() (:274-299): steps with no text until every row has stopped or max_steps frames were made.
 * frames [max_steps, B, rvq] int64 (device); *n_steps = frames produced (host-synchronous at the end). */
int mt_rt_finish(mt_rt* h, int max_steps, const mt_rt_sample_params* p, int use_cuda_graph, int64_t* frames, int* n_steps,
                 mt_stream s);
/* Session state read-back: stopping [B] uint8, attention mask [B, mask_ld >= len] uint8 (device, either may be NULL);
 * *len / *step_idx / *finished (host, host-synchronous). */
int mt_rt_state(mt_rt* h, uint8_t* stopping, uint8_t* mask, long long mask_ld, int* len, int* step_idx, int* finished,
                mt_stream s);
/* generate_local_transformer (:301-357) alone.  hidden [B,H] bf16 = backbone last_hidden_state of the newest position;
 * history [B, gen_step, rvq] int64 or NULL = generated_tokens[:, :gen_step]; forced [B,rvq] int64 or NULL teacher-forces
 * the chosen tokens; logits_out (optional) bf16 [rvq][B][audio_head_size] = every channel's RAW logits;
 * audio_tokens [B,rvq] (optional when forced). */
int mt_rt_local_generate(mt_rt* h, const void* hidden, int B, const int64_t* history, int gen_step,
                         const mt_rt_sample_params* p, const int64_t* forced, void* logits_out, int64_t* audio_tokens,
                         mt_stream s);
/* backbone last_hidden_state [B,H] bf16 of the newest position (what prefill/step handed to the depth loop) */
int mt_rt_last_hidden(mt_rt* h, int B, void* out, mt_stream s);

/* ------------------------------------------------------------------------------------------
 * MOSS-Audio-Tokenizer: decode side (mt_rvq_dequant, mt_codec_decode) and encode side (mt_rlfq_quantize,
 * mt_codec_encode).  A handle is either a decoder (config.encoder = 0) or an encoder (1).
 * ------------------------------------------------------------------------------------------ */
#define MT_CODEC_MAX_STAGES 8
typedef struct {
  int num_stages;
  int input_dim[MT_CODEC_MAX_STAGES], d_model[MT_CODEC_MAX_STAGES], num_heads[MT_CODEC_MAX_STAGES];
  int num_layers[MT_CODEC_MAX_STAGES], dim_ff[MT_CODEC_MAX_STAGES], output_dim[MT_CODEC_MAX_STAGES];
  int patch[MT_CODEC_MAX_STAGES], context[MT_CODEC_MAX_STAGES];
  int code_dim, rvq_dim, num_quantizers, codebook_size, codebook_dim;
  float max_period;
  int encoder;   /* 0: decoder stages, patch[s] = upsample AFTER stage s.  1: encoder stages (encoder_kwargs order),
                  * patch[s] = downsample BEFORE stage s and input_dim[s] the width after it. */
  int layer_gemm_f16;  /* 0: the four Linear weights of every transformer layer are fp32 and run as TF32 tensor-core GEMMs.
                        * 1: they are IEEE half (converted once by the caller) and LayerNorm / attention / GELU hand their
                        *    outputs to the next GEMM as half: kind::f16 MMAs with fp32 accumulation - the same 10-bit operand
                        *    mantissa as TF32 (round-to-nearest instead of truncation) at half the operand bytes.  The residual
                        *    stream, q/k/v, softmax, LayerNorm statistics and every stage input/output projection stay fp32. */
} mt_codec_config;

typedef struct {            /* one transformer layer; fp32, except the four matrices when layer_gemm_f16 = 1 (half) */
  const float *norm1_w, *norm1_b, *norm2_w, *norm2_b;
  const void *in_proj;      /* [3d, d] */
  const void *out_proj;     /* [d, d] */
  const void *linear1;      /* [ff, d] */
  const void *linear2;      /* [d, ff] */
  const float *scale1, *scale2;   /* LayerScale [d] */
} mt_codec_layer;

/* Encode-side quantiser parameters (ResidualLFQ :1217-1236, LFQ :1046-1053), fp32 on device, weight-norm
 * (W = g * v / ||v||, WNConv1d :967-969) already applied by the caller. */
typedef struct {
  const float* in_w;          /* [rvq_dim, code_dim]   quantizer.input_proj */
  const float* in_b;          /* [rvq_dim] */
  const float* q_in_w;        /* [nq, codebook_dim, rvq_dim]   quantizers[q].in_proj */
  const float* q_in_b;        /* [nq, codebook_dim] */
  const float* codebook;      /* [nq, codebook_size, codebook_dim] raw rows (what z_q is built from, :1089) */
  const float* codebook_n;    /* [nq, codebook_size, codebook_dim] F.normalize'd rows (:1082) */
  const float* codebook_nsq;  /* [nq, codebook_size] sum of squares of the normalised rows (:1087) */
  const float* q_out_w;       /* [nq, rvq_dim, codebook_dim]   quantizers[q].out_proj */
  const float* q_out_b;       /* [nq, rvq_dim] */
} mt_codec_quant_weights;

typedef struct {
  /* RVQ: folded per-quantizer tables, table[q][code][c] = W_out * (W_q * C_q[code] + b_q)   (see
   * mt_rvq_fold_tables); out_bias [code_dim]. */
  const float* rvq_table;   /* [num_quantizers, codebook_size, code_dim] */
  const float* rvq_out_bias;
  const float* input_proj[MT_CODEC_MAX_STAGES];   /* [d_model, input_dim] or NULL (Identity) */
  const float* output_proj[MT_CODEC_MAX_STAGES];  /* [output_dim, d_model] or NULL */
  const mt_codec_layer* layers[MT_CODEC_MAX_STAGES];  /* host array per stage */
  const mt_codec_quant_weights* quant;                /* encoder handles only (host struct, copied); else NULL */
} mt_codec_weights;

typedef struct mt_codec mt_codec;
int mt_codec_create(const mt_codec_config* cfg, const mt_codec_weights* w, mt_codec** out);
int mt_codec_destroy(mt_codec* h);
int mt_codec_workspace_bytes(const mt_codec* h, int B, int T, long long* bytes);

/* Builds the folded RVQ table from the reference's parameters (weight-norm g/v pairs, codebooks,
 * biases; modeling_moss_audio_tokenizer.py:967-969,1046-1053,1217-1224), all fp32 on device:
 *   codebook [nq,cs,cd], q_g [nq,rvq], q_v [nq,rvq,cd], q_b [nq,rvq], o_g [code_dim], o_v [code_dim,rvq]
 *   -> table [nq,cs,code_dim]. */
int mt_rvq_fold_tables(const float* codebook, const float* q_g, const float* q_v, const float* q_b,
                       const float* o_g, const float* o_v, int nq, int cs, int cd, int rvq, int code_dim,
                       float* table, mt_stream s);

/* ResidualLFQ.decode_codes (:1276-1283): codes [nq,B,T] int64 -> fp32.
 * channel_major=1: out [B, code_dim, T] (the reference layout); 0: out [B, T, code_dim]. */
int mt_rvq_dequant(const mt_codec* h, const int64_t* codes, int nq, int B, int T, int channel_major,
                   float* out, mt_stream s);

/* _decode_frame (:1513-1534) for codes [nq,B,T] -> wav [B, T*downsample] fp32.
 * chunk_frames > 0 reproduces decode(chunk_duration=...)'s streaming ring-cache visibility
 * (:1728-1750, RingKVCache :413-480) in a single pass; 0 = unchunked.
 * gemm_impl: MT_GEMM_AUTO | MT_GEMM_SIMT. */
int mt_codec_decode(mt_codec* h, const int64_t* codes, int nq, int B, int T, int chunk_frames,
                    float* wav, void* workspace, long long workspace_bytes, mt_stream s);

/* ResidualLFQ.forward (:1238-1274) on encoder hidden states: hidden [B, T, code_dim] fp32 token-major,
 * code_lengths [B] int64 device (valid frames per sample; NULL = all T) -> codes [nq, B, T] int64.
 * The 768 -> 512 input projection and every distance run in fp32 FMA arithmetic (no TF32), ties -> lowest index.
 * workspace: >= B*T*rvq_dim*4 bytes. */
int mt_rlfq_quantize(mt_codec* h, const float* hidden, int B, int T, const int64_t* code_lengths, int nq,
                     int64_t* codes, void* workspace, long long workspace_bytes, mt_stream s);

/* _encode_frame (:1478-1511): wav [B, T*downsample] fp32 (zero-padded to whole frames by the caller) ->
 * codes [nq, B, T] int64 and, if hidden != NULL, encoder_hidden_states [B, T, code_dim] fp32 token-major.
 * chunk_frames > 0 reproduces encode(chunk_duration=...)'s streaming ring-cache visibility in one pass. */
int mt_codec_encode(mt_codec* h, const float* wav, int B, int T, const int64_t* code_lengths, int nq, int chunk_frames,
                    int64_t* codes, float* hidden, void* workspace, long long workspace_bytes, mt_stream s);

/* Incremental streaming: `with model.streaming(batch_size): model._decode_frame(chunk)` / `model._encode_frame(chunk)` per
 * push (:1393-1400; StreamingModule :121-177, RingKVCache :413-480, MHAState :488-499 - how moss_tts_realtime's
 * AudioStreamDecoder drives the codec).  Every layer keeps a ring cache of `context` tokens (K after RoPE, V; half) in the
 * caller's `cache` buffer (mt_codec_stream_bytes; contents need no initialisation); a push of T frames (at most `context`
 * tokens at every stage) appends and attends.  Decoder stream: wav [B, T*downsample] receives the new samples.  Encoder
 * stream (a handle created with cfg.encoder = 1): mt_codec_stream_push_encode takes wav [B, T*downsample] and writes the
 * new codes [nq, B, T] and, if hidden != NULL, encoder_hidden_states [B, T, code_dim].  layer_gemm_f16 = 1 only. */
typedef struct mt_codec_stream mt_codec_stream;
int mt_codec_stream_bytes(const mt_codec* h, int B, long long* bytes);
int mt_codec_stream_begin(mt_codec* h, int B, void* cache, long long cache_bytes, mt_codec_stream** out);
int mt_codec_stream_push(mt_codec_stream* st, const int64_t* codes /* [nq, B, T] */, int nq, int T, float* wav,
                         void* workspace, long long workspace_bytes, mt_stream s);
int mt_codec_stream_push_encode(mt_codec_stream* st, const float* wav /* [B, T*downsample] */, int T,
                                const int64_t* code_lengths /* [B] or NULL */, int nq, int64_t* codes, float* hidden,
                                void* workspace, long long workspace_bytes, mt_stream s);
int mt_codec_stream_end(mt_codec_stream* st);

/* AudioStreamDecoder._apply_crossfade (moss_tts_realtime/mossttsrealtime/streaming_mossttsrealtime.py:776-799):
 * out[i] = prev_tail[i] * fade_out[i] + wav[i] * (1 - fade_out[i]), fade_out = linspace(1, 0, overlap) in fp32
 * (torch's two-sided formula).  All pointers device fp32 [overlap]; out may alias wav. */
int mt_crossfade(const float* prev_tail, const float* wav, int overlap, float* out, mt_stream s);

/* ------------------------------------------------------------------------------------------
 * Prompt side of MossTTSDelayProcessor (moss_tts_delay/processing_moss_tts.py)
 * ------------------------------------------------------------------------------------------ */
/* loudness_normalize (:736-748) of B zero-padded rows wav[b*ld .. b*ld+lengths[b]) in place (rows of length 0 untouched). */
int mt_loudness_normalize(float* wav, long long ld, const int64_t* lengths, int B, float target_dbfs, float gain_lo,
                          float gain_hi, mt_stream s);
/* _get_unified_codes (:588-641): out[r] = (text_codes[r], audio channels).  blocks [n_blocks][4] int32 on device, in row
 * order: {first row, rows, T, first code row}; T >= 0: the delay pattern (:516-525) of codes[first code row .. +T) [.., n_vq],
 * T = -1: pad rows.  Rows outside every block are pad_code. */
int mt_pack_unified_codes(const int64_t* text_codes, const int* blocks, int n_blocks, const int64_t* codes, int n_vq,
                          long long pad_code, int rows, int64_t* out, mt_stream s);
/* _pad (:410-431): rows[b] -> device pointer to [lens[b], C] int64; out [B, Lmax, C] left-padded with
 * (pad_token, pad_code, ...), mask [B, Lmax] uint8 (1 = real row). */
int mt_pad_left(const int64_t* const* rows, const int* lens, int B, int Lmax, int C, long long pad_token,
                long long pad_code, int64_t* out, uint8_t* mask, mt_stream s);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif

#ifdef __cplusplus
}
#endif
#endif /* MOSSTTS_H_ */
