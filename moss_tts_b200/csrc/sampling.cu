// Fused sampling + delay-pattern state machine: one launch samples every (row, channel) of a step,
// a second tiny launch commits the state.  No host synchronisation inside a step.
//   a4  text-logit masking / forced tokens     modeling_moss_tts.py:451-464
//   a5  sample_token                           inference_utils.py:111-145 (penalty -> argmax | top-k -> top-p -> draw)
//   a6  audio active mask + scatter            modeling_moss_tts.py:475-503
//   a7  state update / history append          modeling_moss_tts.py:472-473,505-516
//   a8  find_last_equal_C                      inference_utils.py:148-165
//   a9  (de-)delay pattern                     processing_moss_tts.py:516-537
// Logits are bf16; every transform returns the value a bf16 PyTorch tensor would hold.
#include <cstddef>

#include "sampling.cuh"
#include "sampling_full.cuh"

namespace mt {

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// ---- uniform hash shared with oracle/sampling.py::uniform_u ----------------------------------
__host__ __device__ inline uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7FEB352Du; x ^= x >> 15; x *= 0x846CA68Bu; x ^= x >> 16;
  return x;
}
__host__ __device__ inline float mt_uniform(uint32_t seed, uint32_t step, uint32_t row, uint32_t channel) {
  uint32_t h = mix32(seed ^ 0x9E3779B9u);
  h = mix32(h ^ (step * 0x85EBCA6Bu));
  h = mix32(h ^ (row * 0xC2B2AE35u));
  h = mix32(h ^ (channel * 0x27D4EB2Fu));
  return static_cast<float>(h >> 8) * (1.0f / 16777216.0f);
}

// ---- greedy: first index of the maximum ------------------------------------------------------
__device__ int row_argmax(const RowXform& x, uint32_t* s_u32) {
  uint32_t best_k = 0; int best_i = 0x7fffffff;
  for (int i = threadIdx.x; i < x.V; i += blockDim.x) {
    const uint32_t k = key16(x.get(i));
    if (k > best_k || (k == best_k && i < best_i)) { best_k = k; best_i = i; }
  }
  // pack (key, ~index) so that max() prefers the larger key, then the smaller index
  unsigned long long p = (static_cast<unsigned long long>(best_k) << 32) | (0xFFFFFFFFu - static_cast<uint32_t>(best_i));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long q = __shfl_xor_sync(0xffffffffu, p, o);
    p = q > p ? q : p;
  }
  unsigned long long* s64 = reinterpret_cast<unsigned long long*>(s_u32);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) s64[warp] = p;
  __syncthreads();
  if (warp == 0) {
    unsigned long long v = lane < (blockDim.x >> 5) ? s64[lane] : 0ull;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long q = __shfl_xor_sync(0xffffffffu, v, o);
      v = q > v ? q : v;
    }
    if (lane == 0) s64[32] = v;
  }
  __syncthreads();
  const unsigned long long r = s64[32];
  __syncthreads();
  return static_cast<int>(0xFFFFFFFFu - static_cast<uint32_t>(r & 0xFFFFFFFFu));
}

// ---- sampling: radix-select the k-th largest 16-bit key, collect, sort, top-p, draw ----------
struct SampSmem {
  int hist[256];
  int scan[kSampThreads + 1];
  unsigned long long cand[kMaxTopK];
  int misc[8];
  unsigned long long red[34];
};

// from the top bin down, find the bin where the running count reaches `need`
__device__ void find_bin(const int* hist, int need, int* out_bin, int* out_above) {
  int acc = 0;
  for (int b = 255; b >= 0; --b) {
    if (acc + hist[b] >= need) { *out_bin = b; *out_above = acc; return; }
    acc += hist[b];
  }
  *out_bin = 0; *out_above = acc - hist[0];
}

// ebuf / result: scratch outside `sm` (SelSmem<BlockG>::ebuf lies behind the end of SampSmem in the union below)
__device__ int row_sample(const RowXform& x, int top_k, float top_p, float u, SampSmem& sm, float* ebuf, int* result) {
  const int tid = threadIdx.x;
  const int k = top_k < x.V ? top_k : x.V;
  // pass 1: high byte
  for (int i = tid; i < 256; i += blockDim.x) sm.hist[i] = 0;
  __syncthreads();
  for (int i = tid; i < x.V; i += blockDim.x) atomicAdd(&sm.hist[key16(x.get(i)) >> 8], 1);
  __syncthreads();
  if (tid == 0) find_bin(sm.hist, k, &sm.misc[0], &sm.misc[1]);
  __syncthreads();
  const int hi = sm.misc[0], above_hi = sm.misc[1];
  // pass 2: low byte inside the selected high bin
  for (int i = tid; i < 256; i += blockDim.x) sm.hist[i] = 0;
  __syncthreads();
  for (int i = tid; i < x.V; i += blockDim.x) {
    const uint32_t kk = key16(x.get(i));
    if (static_cast<int>(kk >> 8) == hi) atomicAdd(&sm.hist[kk & 255], 1);
  }
  __syncthreads();
  if (tid == 0) find_bin(sm.hist, k - above_hi, &sm.misc[2], &sm.misc[3]);
  __syncthreads();
  const uint32_t thr = (static_cast<uint32_t>(hi) << 8) | static_cast<uint32_t>(sm.misc[2]);
  const int need_eq = k - above_hi - sm.misc[3];
  // pass 3: ordered collection.  Thread t owns the contiguous index segment [t*seg,(t+1)*seg) so a
  // single block scan ranks ties by index (lowest index wins, as in the oracle).
  const int seg = (x.V + blockDim.x - 1) / blockDim.x;
  const int i0 = tid * seg, i1 = min(x.V, i0 + seg);
  int n_gt = 0, n_eq = 0;
  for (int i = i0; i < i1; ++i) {
    const uint32_t kk = key16(x.get(i));
    n_gt += kk > thr;
    n_eq += kk == thr;
  }
  // exclusive scans of n_gt and n_eq over threads (packed: eq in the high half)
  sm.scan[tid] = n_gt | (n_eq << 16);
  __syncthreads();
  if (tid == 0) {
    int g = 0, e = 0;
    for (int t = 0; t < static_cast<int>(blockDim.x); ++t) {
      const int v = sm.scan[t];
      sm.scan[t] = g | (e << 16);
      g += v & 0xFFFF;
      e = min(e + (v >> 16), 32767);     // only "rank < need_eq (<=1024)" matters; clamp to stay in 15 bits
    }
    sm.misc[4] = g;
  }
  __syncthreads();
  const int total_gt = sm.misc[4];
  int off_gt = sm.scan[tid] & 0xFFFF, off_eq = sm.scan[tid] >> 16;
  for (int i = tid; i < kMaxTopK; i += blockDim.x) sm.cand[i] = ~0ull;
  __syncthreads();
  for (int i = i0; i < i1; ++i) {
    const uint32_t kk = key16(x.get(i));
    int slot = -1;
    if (kk > thr) slot = off_gt++;
    else if (kk == thr) { if (off_eq < need_eq) slot = total_gt + off_eq; off_eq++; }
    if (slot >= 0 && slot < kMaxTopK && kk != kKeyNegInf && kk != 0u)
      sm.cand[slot] = (static_cast<unsigned long long>(0xFFFFu - kk) << 32) | static_cast<uint32_t>(i);
  }
  __syncthreads();
  // bitonic sort ascending on (0xFFFF-key, index)  ==  value descending, index ascending
  int n2 = 1;
  while (n2 < k) n2 <<= 1;
  for (int size = 2; size <= n2; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = tid; i < n2; i += blockDim.x) {
        const int j = i ^ stride;
        if (j > i) {
          const unsigned long long a = sm.cand[i], b = sm.cand[j];
          const bool up = (i & size) == 0;
          if ((a > b) == up) { sm.cand[i] = b; sm.cand[j] = a; }
        }
      }
      __syncthreads();
    }
  }
  // softmax -> top-p cut -> softmax -> inverse-CDF draw on the sorted list: the same routine as the one-pass path
  // (exponentials and divisions by all threads, the order-sensitive fp32 sums by one)
  if (tid == 0) sm.misc[6] = 0;
  __syncthreads();
  for (int j = tid; j < k; j += blockDim.x)
    if (sm.cand[j] != ~0ull && (j + 1 == k || sm.cand[j + 1] == ~0ull)) sm.misc[6] = j + 1;   // valid entries sort first
  __syncthreads();
  const int n = sm.misc[6];
  __syncthreads();
  return draw_sorted<BlockG>(sm.cand, n, top_p, u, ebuf, result);
}

// ---------------------------------------------------------------------------------------------
// token of one row: one-pass selection (sampling_select.cuh), exact multi-pass fallback
// ---------------------------------------------------------------------------------------------
struct BlockSmem {
  union {
    SelSmem<BlockG> sel;
    SampSmem old;
  };
};
static_assert(offsetof(SelSmem<BlockG>, ebuf) >= sizeof(SampSmem),
              "row_sample's draw scratch (sel.ebuf, sel.result) must lie behind the multi-pass selection state");
union SampleShared {
  BlockSmem blk;
  SelSmem<WarpG> warp[kSampThreads / 32];
  FullSmem full;
};

// mode 2 (text channel inside audio): only the gen / delay slots survive the mask -- two candidates, no row scan.
// Greedy: the larger one, the lower index on a tie, 0 when both are -inf (what torch.argmax gives on an all -inf row).
__device__ int two_candidates(const RowXform& x, int k, SelSmem<BlockG>& sm) {
  if (threadIdx.x == 0) {
    const int i0 = min(x.keep0, x.keep1), i1 = max(x.keep0, x.keep1);
    unsigned long long e[2];
    int n = 0;
    for (int t = 0; t < 2; ++t) {
      const int i = t ? i1 : i0;
      if (t && i1 == i0) break;
      const uint32_t fk = (i >= 0 && i < x.V) ? key16(x.get(i)) : 0u;
      if (fk > kKeyNegInf) e[n++] = (static_cast<unsigned long long>(0xFFFFu - fk) << 32) | static_cast<uint32_t>(i);
    }
    if (n == 2 && e[1] < e[0]) { const unsigned long long t = e[0]; e[0] = e[1]; e[1] = t; }
    n = min(n, k);
    for (int j = 0; j < n; ++j) sm.list[j] = e[j];
    sm.n = n;
  }
  __syncthreads();
  const int n = sm.n;
  __syncthreads();
  return n;
}

__device__ int block_pick(const RowXform& x, bool do_sample, int top_k, float top_p, float u, BlockSmem& sm) {
  const int k = do_sample ? min(top_k, x.V) : 1;
  const int n = x.mode == 2 ? two_candidates(x, k, sm.sel)
                            : (x.bitmap == nullptr ? select_topk<BlockG, true>(x, k, sm.sel) : select_topk<BlockG, false>(x, k, sm.sel));
  if (n < 0) {
    __syncthreads();
    return do_sample ? row_sample(x, top_k, top_p, u, sm.old, sm.sel.ebuf, &sm.sel.result) : row_argmax(x, reinterpret_cast<uint32_t*>(sm.old.red));
  }
  if (!do_sample) return n ? static_cast<int>(sm.sel.list[0] & 0xFFFFFFFFu) : 0;
  return draw_sorted<BlockG>(sm.sel.list, n, top_p, u, sm.sel.ebuf, &sm.sel.result);
}

// one warp per row (short rows, k <= kWarpMaxK)
constexpr int kWarpMaxK = 58;          // k + 6 banned ids <= the 64 half-stream maxima of a warp
constexpr int kWarpMaxV = 4096;
__device__ int warp_pick(const RowXform& x, bool do_sample, int top_k, float top_p, float u, SelSmem<WarpG>& sm) {
  const int k = do_sample ? min(top_k, x.V) : 1;
  int n = x.bitmap == nullptr ? select_topk<WarpG, true>(x, k, sm) : select_topk<WarpG, false>(x, k, sm);
  if (n < 0) {
    __syncwarp();
    n = warp_exact_select(x, k, sm);
  }
  if (!do_sample) return n ? static_cast<int>(sm.list[0] & 0xFFFFFFFFu) : 0;
  return draw_sorted<WarpG>(sm.list, n, top_p, u, sm.ebuf, &sm.result);
}

// ---------------------------------------------------------------------------------------------
// WA = true : blocks [0, B) take the text rows (one block per row); every later block takes eight
//             (row, audio channel) pairs, one per warp
// WA = false: one block per (channel, row), grid (C, B)   (audio heads too long or top-k too large for a warp)
// ---------------------------------------------------------------------------------------------
template <bool WA>
__global__ void __launch_bounds__(kSampThreads, 3)
gen_sample_kernel(GenState st, const bf16* __restrict__ text_logits, int vocab,
                  const bf16* __restrict__ audio_logits, int audio_head, SampleCfg sc, TokenIds tk) {
  __shared__ SampleShared sm;
  pdl_trigger();
  pdl_wait();
  if (*st.done) return;
  const int n_vq = st.C - 1;
  const int warp = threadIdx.x >> 5;
  int ch, b;
  bool warp_mode = false;
  if (WA) {
    if (static_cast<int>(blockIdx.x) < st.B) { ch = 0; b = blockIdx.x; }
    else {
      const int pair = (blockIdx.x - st.B) * (kSampThreads / 32) + warp;
      if (pair >= st.B * n_vq) return;
      b = pair / n_vq; ch = 1 + pair % n_vq; warp_mode = true;
    }
  } else {
    ch = blockIdx.x; b = blockIdx.y;
  }
  const bool leader = warp_mode ? (threadIdx.x & 31) == 0 : threadIdx.x == 0;
  const int step = *st.step;
  const int delayed = st.delayed[b];
  int64_t* dst = st.next_ids + static_cast<long long>(b) * st.C + ch;
  RowXform x;
  x.mode = 0; x.nban = 0; x.keep0 = x.keep1 = -1; x.bitmap = nullptr; x.penalty = 1.f; x.penalty_first = 0;
  if (ch == 0) {
    // ---- text channel: forced tokens first (modeling_moss_tts.py:453-458)
    int forced = -1;
    if (st.stopping[b]) forced = tk.pad;
    else if (delayed >= 0 && delayed < n_vq) forced = tk.delay_slot;
    else if (delayed == n_vq) forced = tk.audio_end;
    if (forced >= 0) { if (leader) *dst = forced; return; }
    x.row = text_logits + static_cast<long long>(b) * vocab; x.V = vocab; x.temp = sc.text_temperature;
    if (st.is_audio[b]) {
      x.mode = 2; x.keep0 = tk.gen_slot; x.keep1 = tk.delay_slot;
    } else {
      x.mode = 1;
      x.ban[0] = tk.pad; x.ban[1] = tk.gen_slot; x.ban[2] = tk.delay_slot; x.ban[3] = tk.audio_end; x.nban = 4;
    }
    if (step == 0) x.ban[x.nban++] = 151662;              // hard-coded id in the reference (:462)
    if (step <= n_vq) x.ban[x.nban++] = tk.im_end;
    // no top-k filter, or one beyond the shared-memory candidate list: the full-vocabulary counting-sort sampler
    const bool full = sc.text_do_sample && (sc.text_top_k <= 0 || sc.text_top_k > kMaxTopK);
    const float u = mt_uniform(sc.seed, step, b, 0);
    const int tok = (full && x.mode != 2)
                        ? full_pick(x, sc.text_top_k, sc.text_top_p, u, st.full_hist + static_cast<size_t>(b) * kFullBuckets, sm.full)
                        : block_pick(x, sc.text_do_sample, full ? 2 : sc.text_top_k, sc.text_top_p, u, sm.blk);
    if (leader) *dst = tok;
  } else {
    // ---- audio channel i = ch-1 (modeling_moss_tts.py:475-503); lengths are the pre-update ones
    const int i = ch - 1;
    const bool active = st.audio_len[b] > i && (delayed < 0 || i > delayed - 1);
    if (!active) { if (leader) *dst = tk.audio_pad_code; return; }
    x.row = audio_logits + (static_cast<long long>(b) * n_vq + i) * audio_head; x.V = audio_head;
    x.temp = sc.audio_temperature;
    x.ban[0] = tk.audio_pad_code; x.nban = 1;
    if (sc.penalty != 1.0f) { x.bitmap = i == 0 ? st.seen0 : st.seen_rest; x.penalty = sc.penalty; }
    const float u = mt_uniform(sc.seed, step, b, ch);
    const int tok = warp_mode ? warp_pick(x, sc.audio_do_sample, sc.audio_top_k, sc.audio_top_p, u, sm.warp[warp])
                              : block_pick(x, sc.audio_do_sample, sc.audio_top_k, sc.audio_top_p, u, sm.blk);
    if (leader) *dst = tok;
  }
}

// state update + history append (:472-473,505-516); a single block
__global__ void gen_commit_kernel(GenState st, TokenIds tk) {
  pdl_trigger();
  pdl_wait();
  if (*st.done) return;
  const int n_vq = st.C - 1;
  const int len = *st.len;
  int all_stopped = 1, need_full = 0;
  for (int b = threadIdx.x; b < st.B; b += blockDim.x) {
    const int64_t* nx = st.next_ids + static_cast<long long>(b) * st.C;
    const int tok = static_cast<int>(nx[0]);
    int delayed = st.delayed[b];
    uint8_t stopping = st.stopping[b], is_audio = st.is_audio[b];
    if (!stopping && delayed == n_vq) is_audio = 0;          // is_audio[is_audio_eos] = False (:457)
    if (tok == tk.audio_start) is_audio = 1;
    if (tok == tk.im_end) stopping = 1;
    int al = st.audio_len[b];
    if (tok == tk.audio_start || tok == tk.gen_slot || tok == tk.delay_slot) al += 1;
    if (tok == tk.audio_end) al = 0;
    if (delayed < 0 && tok == tk.delay_slot) delayed = 0;
    if (delayed >= 0) delayed += 1;
    if (delayed > n_vq) delayed = -1;
    st.audio_len[b] = al; st.delayed[b] = delayed; st.stopping[b] = stopping; st.is_audio[b] = is_audio;
    if (len < st.max_len) st.mask[static_cast<long long>(b) * st.max_len + len] = !stopping;
    all_stopped &= stopping;
    // will the next step's text sampler read this row's whole logit row?  (forced rows read nothing,
    // rows inside an audio span read only the gen/delay-slot logits; see gen_sample_kernel)
    const bool forced = stopping || (delayed >= 0 && delayed <= n_vq);
    need_full |= !forced && !is_audio;
  }
  // history append + repetition-penalty id sets (pad code included, as torch.unique sees it), one thread
  // per (row, channel)
  for (int i = threadIdx.x; i < st.B * st.C; i += blockDim.x) {
    const int b = i / st.C, c = i % st.C;
    const long long a = st.next_ids[i];
    if (len < st.max_len) st.gen_ids[(static_cast<long long>(b) * st.max_len + len) * st.C + c] = a;
    if (c >= 1 && a >= 0 && a < st.bitmap_words * 32) atomicOr(c == 1 ? &st.seen0[a >> 5] : &st.seen_rest[a >> 5], 1u << (a & 31));
  }
  all_stopped = __syncthreads_and(all_stopped);
  need_full = __syncthreads_or(need_full);
  if (threadIdx.x == 0) {
    *st.len = len + 1;
    *st.pos = len;
    *st.step = *st.step + 1;
    if (all_stopped) *st.done = 1;
    *st.text_full = need_full;
  }
}

// prologue of generate(): continuation detection, history/mask copy, penalty sets (:417-436,518)
__global__ void gen_init_kernel(GenState st, const int64_t* __restrict__ input_ids,
                                const uint8_t* __restrict__ attn_mask, int S0, TokenIds tk,
                                int64_t* __restrict__ start_indices) {
  const int b = blockIdx.x;
  const int64_t* src = input_ids + static_cast<long long>(b) * S0 * st.C;
  int64_t* g = st.gen_ids + static_cast<long long>(b) * st.max_len * st.C;
  __shared__ int s_last_start, s_last_im;
  if (threadIdx.x == 0) { s_last_start = -1; s_last_im = -1; }
  __syncthreads();
  for (int i = threadIdx.x; i < S0 * st.C; i += blockDim.x) {
    const int64_t v = src[i];
    g[i] = v;
    const int t = i / st.C, c = i % st.C;
    if (c == 0) {
      if (v == tk.audio_start) atomicMax(&s_last_start, t);
      if (v == tk.im_start) atomicMax(&s_last_im, t);
    } else if (v >= 0 && v < st.bitmap_words * 32) {
      atomicOr(c == 1 ? &st.seen0[v >> 5] : &st.seen_rest[v >> 5], 1u << (v & 31));
    }
  }
  for (int t = threadIdx.x; t < st.max_len; t += blockDim.x)
    st.mask[static_cast<long long>(b) * st.max_len + t] = t < S0 ? (attn_mask ? attn_mask[static_cast<long long>(b) * S0 + t] : 1) : 0;
  __syncthreads();
  if (threadIdx.x == 0) {
    const int64_t last = src[static_cast<long long>(S0 - 1) * st.C];
    const bool cont = (last == tk.audio_start || last == tk.gen_slot) && s_last_start >= 0;
    st.audio_len[b] = cont ? S0 - s_last_start : 0;
    st.is_audio[b] = cont;
    st.stopping[b] = 0;
    st.delayed[b] = -1;
    start_indices[b] = s_last_im + 3;
    if (b == 0) { *st.len = S0; *st.step = 0; *st.done = 0; *st.pos = S0 - 1; *st.text_full = 1; }
  }
}


size_t gen_state_bytes(int B, int C, int max_len, int audio_head) {
  const int words = (audio_head + 31) / 32;
  size_t n = 256;                                       // scalars
  n += align_up(B, 256) * 2;                            // stopping, is_audio
  n += align_up(B * sizeof(int), 256) * 2;              // audio_len, delayed
  n += align_up(static_cast<size_t>(B) * C * 8, 256);   // next_ids
  n += align_up(static_cast<size_t>(B) * max_len, 256); // mask
  n += align_up(words * 4, 256) * 2;
  return n;
}

GenState gen_state_carve(void* buf, int B, int C, int max_len, int audio_head, int64_t* gen_ids) {
  GenState st{};
  uint8_t* p = reinterpret_cast<uint8_t*>(buf);
  const int words = (audio_head + 31) / 32;
  st.len = reinterpret_cast<int*>(p); st.step = st.len + 1; st.done = st.len + 2; st.pos = st.len + 3;
  st.text_full = st.len + 4; p += 256;
  st.stopping = p; p += align_up(B, 256);
  st.is_audio = p; p += align_up(B, 256);
  st.audio_len = reinterpret_cast<int*>(p); p += align_up(B * sizeof(int), 256);
  st.delayed = reinterpret_cast<int*>(p); p += align_up(B * sizeof(int), 256);
  st.next_ids = reinterpret_cast<int64_t*>(p); p += align_up(static_cast<size_t>(B) * C * 8, 256);
  st.mask = p; p += align_up(static_cast<size_t>(B) * max_len, 256);
  st.seen0 = reinterpret_cast<uint32_t*>(p); p += align_up(words * 4, 256);
  st.seen_rest = reinterpret_cast<uint32_t*>(p);
  st.gen_ids = gen_ids; st.B = B; st.C = C; st.max_len = max_len; st.bitmap_words = words;
  return st;
}

int gen_init(const GenState& st, const int64_t* input_ids, const uint8_t* attn_mask, int S0, const TokenIds& tk,
             int64_t* start_indices, cudaStream_t s) {
  MT_CUDA(cudaMemsetAsync(st.seen0, 0, st.bitmap_words * 4, s));
  MT_CUDA(cudaMemsetAsync(st.seen_rest, 0, st.bitmap_words * 4, s));
  gen_init_kernel<<<st.B, 256, 0, s>>>(st, input_ids, attn_mask, S0, tk, start_indices);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

int gen_sample_step(const GenState& st, const bf16* text_logits, int vocab, const bf16* audio_logits, int audio_head,
                    const SampleCfg& sc, const TokenIds& tk, cudaStream_t s) {
  const int n_vq = st.C - 1;
  if (audio_head <= kWarpMaxV && (!sc.audio_do_sample || sc.audio_top_k <= kWarpMaxK)) {
    const int grid = st.B + cdiv(st.B * n_vq, kSampThreads / 32);
    MT_CUDA(launch(gen_sample_kernel<true>, dim3(grid), dim3(kSampThreads), 0, s, st, text_logits, vocab, audio_logits, audio_head, sc, tk));
  } else {
    MT_CUDA(launch(gen_sample_kernel<false>, dim3(st.C, st.B), dim3(kSampThreads), 0, s, st, text_logits, vocab, audio_logits, audio_head, sc, tk));
  }
  MT_CUDA(launch(gen_commit_kernel, dim3(1), dim3(1024), 0, s, st, tk));
  count_launch(2);
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// MOSS-TTS-Realtime (reference moss_tts_realtime/mossttsrealtime/streaming_mossttsrealtime.py): session state and
// the per-channel sampler of generate_local_transformer :301-357.
//   rt_window_sets : ids of history[b, max(0, step - window) : step, c] as one bitmap per (row, channel) -- the set
//                    apply_repetition_penalty :359-375 gathers / scatters (duplicates are penalised once)
//   rt_sample      : penalty on the raw logits (greedy too), then sample_token :377-397 (argmax, or / temperature ->
//                    top-k -> top-p -> inverse-CDF draw of this library)
//   rt_begin_step  : step_ids = [text | last frame], mask column = !stopping (:244-247)
//   rt_commit      : history append, stopping |= (channel 0 == eos), counters (:263-267)
// ---------------------------------------------------------------------------------------------
__global__ void rt_window_sets_kernel(RtState st, const int64_t* __restrict__ history, long long ld_frames,
                                      const int* __restrict__ step_dev, int step_host, int window) {
  pdl_trigger();
  pdl_wait();
  if (*st.done) return;
  const int b = blockIdx.x;
  const int step = step_dev ? *step_dev : step_host;
  uint32_t* bm = st.seen + static_cast<long long>(b) * st.R * st.words;
  for (int i = threadIdx.x; i < st.R * st.words; i += blockDim.x) bm[i] = 0u;
  __syncthreads();
  if (!history || step <= 0) return;
  const int t0 = window > 0 ? max(0, step - window) : 0;
  const int n = (step - t0) * st.R;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int t = t0 + i / st.R, c = i % st.R;
    const long long v = history[(static_cast<long long>(b) * ld_frames + t) * st.R + c];
    if (v >= 0 && v < st.words * 32) atomicOr(&bm[c * st.words + (v >> 5)], 1u << (v & 31));
  }
}

template <bool W>
__global__ void __launch_bounds__(kSampThreads, 3)
rt_sample_kernel(RtState st, const bf16* __restrict__ logits, int V, int c, int do_sample, float temperature, float penalty,
                 float top_p, int top_k, unsigned seed, const int* __restrict__ step_dev, int step_host,
                 int64_t* __restrict__ out) {
  __shared__ SampleShared sm;
  pdl_trigger();
  pdl_wait();
  if (*st.done) return;
  const int warp = threadIdx.x >> 5;
  const int b = W ? blockIdx.x * (kSampThreads / 32) + warp : blockIdx.x;
  if (b >= st.B) return;
  RowXform x;
  x.row = logits + static_cast<long long>(b) * V; x.V = V; x.mode = 0; x.nban = 0; x.keep0 = x.keep1 = -1;
  x.temp = do_sample ? temperature : 1.f; x.bitmap = nullptr; x.penalty = 1.f; x.penalty_first = 1;
  if (penalty != 1.0f) {
    x.bitmap = st.seen + (static_cast<long long>(b) * st.R + c) * st.words;
    x.penalty = penalty;
  }
  const float u = mt_uniform(seed, step_dev ? *step_dev : step_host, b, c);
  const int tok = W ? warp_pick(x, do_sample, top_k, top_p, u, sm.warp[warp]) : block_pick(x, do_sample, top_k, top_p, u, sm.blk);
  if (W ? (threadIdx.x & 31) == 0 : threadIdx.x == 0) out[static_cast<long long>(b) * st.R + c] = tok;
}

__global__ void rt_prefill_init_kernel(RtState st, const uint8_t* __restrict__ attn_mask, int S0, int past) {
  const int b = blockIdx.x;
  for (int t = threadIdx.x; t < S0; t += blockDim.x)
    st.mask[static_cast<long long>(b) * st.max_seq + past + t] = attn_mask[static_cast<long long>(b) * S0 + t] ? 1 : 0;
  if (threadIdx.x == 0) {
    st.stopping[b] = 0;
    if (b == 0) { *st.len = past + S0; *st.step = 0; *st.done = 0; *st.pos = past + S0 - 1; }
  }
}

__global__ void rt_set_text_kernel(RtState st, const int64_t* __restrict__ text_tokens, long long text_pad) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < st.B) st.text_in[b] = text_tokens ? text_tokens[b] : text_pad;
}

__global__ void rt_begin_step_kernel(RtState st) {
  pdl_trigger();
  pdl_wait();
  if (*st.done) return;
  const int len = *st.len, C = st.R + 1;
  for (int i = threadIdx.x; i < st.B * C; i += blockDim.x) {
    const int b = i / C, c = i % C;
    st.step_ids[i] = c == 0 ? st.text_in[b] : st.last[static_cast<long long>(b) * st.R + c - 1];
  }
  if (len < st.max_seq)
    for (int b = threadIdx.x; b < st.B; b += blockDim.x) st.mask[static_cast<long long>(b) * st.max_seq + len] = st.stopping[b] ? 0 : 1;
  if (threadIdx.x == 0) *st.pos = min(len, st.max_seq - 1);
}

__global__ void rt_commit_kernel(RtState st, int eos, int advance_len, int64_t* __restrict__ tokens_out) {
  pdl_trigger();
  pdl_wait();
  const bool done = *st.done != 0;
  __shared__ int any_running;
  if (threadIdx.x == 0) any_running = 0;
  __syncthreads();
  const int step = *st.step;
  for (int i = threadIdx.x; i < st.B * st.R; i += blockDim.x) {
    const int b = i / st.R, c = i % st.R;
    const long long t = st.last[i];
    if (tokens_out) tokens_out[i] = t;                      // a finished session keeps returning its last frame (:232-233)
    if (!done && step < st.max_frames) st.history[(static_cast<long long>(b) * st.max_frames + step) * st.R + c] = t;
  }
  if (done) return;
  for (int b = threadIdx.x; b < st.B; b += blockDim.x) {
    uint8_t sp = st.stopping[b];
    if (st.last[static_cast<long long>(b) * st.R] == eos) sp = 1;
    st.stopping[b] = sp;
    if (!sp) atomicOr(&any_running, 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    *st.step = step + 1;
    if (advance_len) *st.len = *st.len + 1;
    if (!any_running) *st.done = 1;
  }
}

size_t rt_state_bytes(int B, int R, int max_frames, int max_seq, int audio_head) {
  const int words = (audio_head + 31) / 32;
  return 256 + align_up(B, 256) + align_up(static_cast<size_t>(B) * 8, 256) + align_up(static_cast<size_t>(B) * (R + 1) * 8, 256) +
         align_up(static_cast<size_t>(B) * R * 8, 256) + align_up(static_cast<size_t>(B) * max_frames * R * 8, 256) +
         align_up(static_cast<size_t>(B) * max_seq, 256) + align_up(static_cast<size_t>(B) * R * words * 4, 256);
}

RtState rt_state_carve(void* buf, int B, int R, int max_frames, int max_seq, int audio_head) {
  RtState st{};
  uint8_t* p = reinterpret_cast<uint8_t*>(buf);
  const int words = (audio_head + 31) / 32;
  st.len = reinterpret_cast<int*>(p); st.step = st.len + 1; st.done = st.len + 2; st.pos = st.len + 3; p += 256;
  st.stopping = p; p += align_up(B, 256);
  st.text_in = reinterpret_cast<int64_t*>(p); p += align_up(static_cast<size_t>(B) * 8, 256);
  st.step_ids = reinterpret_cast<int64_t*>(p); p += align_up(static_cast<size_t>(B) * (R + 1) * 8, 256);
  st.last = reinterpret_cast<int64_t*>(p); p += align_up(static_cast<size_t>(B) * R * 8, 256);
  st.history = reinterpret_cast<int64_t*>(p); p += align_up(static_cast<size_t>(B) * max_frames * R * 8, 256);
  st.mask = p; p += align_up(static_cast<size_t>(B) * max_seq, 256);
  st.seen = reinterpret_cast<uint32_t*>(p);
  st.B = B; st.R = R; st.max_frames = max_frames; st.max_seq = max_seq; st.words = words;
  return st;
}

int rt_prefill_init(const RtState& st, const uint8_t* attn_mask, int S0, int past, cudaStream_t s) {
  rt_prefill_init_kernel<<<st.B, 256, 0, s>>>(st, attn_mask, S0, past);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

int rt_set_text(const RtState& st, const int64_t* text_tokens, long long text_pad, cudaStream_t s) {
  rt_set_text_kernel<<<cdiv(st.B, 128), 128, 0, s>>>(st, text_tokens, text_pad);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

int rt_begin_step(const RtState& st, cudaStream_t s) {
  MT_CUDA(launch(rt_begin_step_kernel, dim3(1), dim3(1024), 0, s, st));
  count_launch();
  return MT_OK;
}

int rt_window_sets(const RtState& st, const int64_t* history, long long ld_frames, const int* step_dev, int step_host,
                   int window, cudaStream_t s) {
  MT_CUDA(launch(rt_window_sets_kernel, dim3(st.B), dim3(256), 0, s, st, history, ld_frames, step_dev, step_host, window));
  count_launch();
  return MT_OK;
}

int rt_sample(const RtState& st, const bf16* logits, int V, int c, int do_sample, float temperature, float penalty,
              float top_p, int top_k, unsigned seed, const int* step_dev, int step_host, int64_t* out, cudaStream_t s) {
  if (do_sample && (top_k <= 0 || top_k > kMaxTopK))
    return fail(MT_ERR_UNSUPPORTED, "sampling requires 1 <= top_k <= %d (got %d)", kMaxTopK, top_k);
  if (V <= kWarpMaxV && (!do_sample || top_k <= kWarpMaxK))
    MT_CUDA(launch(rt_sample_kernel<true>, dim3(cdiv(st.B, kSampThreads / 32)), dim3(kSampThreads), 0, s, st, logits, V, c,
                   do_sample, temperature, penalty, top_p, top_k, seed, step_dev, step_host, out));
  else
    MT_CUDA(launch(rt_sample_kernel<false>, dim3(st.B), dim3(kSampThreads), 0, s, st, logits, V, c, do_sample, temperature,
                   penalty, top_p, top_k, seed, step_dev, step_host, out));
  count_launch();
  return MT_OK;
}

int rt_commit(const RtState& st, int eos, int advance_len, int64_t* tokens_out, cudaStream_t s) {
  MT_CUDA(launch(rt_commit_kernel, dim3(1), dim3(1024), 0, s, st, eos, advance_len, tokens_out));
  count_launch();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// standalone sample_token (inference_utils.py:111-145) on [N,V] bf16 logits
// ---------------------------------------------------------------------------------------------
template <bool W>
__global__ void __launch_bounds__(kSampThreads, W ? 4 : 3)
sample_tokens_kernel(const bf16* __restrict__ logits, long long ld, int N, int V, float temperature,
                     const uint32_t* __restrict__ bitmap, float penalty, float top_p, int top_k, int do_sample,
                     const float* __restrict__ uniforms, int64_t* __restrict__ out) {
  __shared__ SampleShared sm;
  const int warp = threadIdx.x >> 5;
  const int r = W ? blockIdx.x * (kSampThreads / 32) + warp : blockIdx.x;
  if (r >= N) return;
  RowXform x;
  x.row = logits + r * ld; x.V = V; x.temp = temperature; x.mode = 0; x.nban = 0; x.keep0 = x.keep1 = -1;
  x.bitmap = penalty != 1.0f ? bitmap : nullptr; x.penalty = penalty; x.penalty_first = 0;
  const float u = uniforms ? uniforms[r] : 0.f;
  const int tok = W ? warp_pick(x, do_sample, top_k, top_p, u, sm.warp[warp]) : block_pick(x, do_sample, top_k, top_p, u, sm.blk);
  if (W ? (threadIdx.x & 31) == 0 : threadIdx.x == 0) out[r] = tok;
}

__global__ void __launch_bounds__(kSampThreads, 3)
sample_full_kernel(const bf16* __restrict__ logits, long long ld, int N, int V, float temperature,
                   const uint32_t* __restrict__ bitmap, float penalty, float top_p, int top_k,
                   const float* __restrict__ uniforms, int64_t* __restrict__ out, uint32_t* __restrict__ hist) {
  __shared__ FullSmem sm;
  const int r = blockIdx.x;
  RowXform x;
  x.row = logits + r * ld; x.V = V; x.temp = temperature; x.mode = 0; x.nban = 0; x.keep0 = x.keep1 = -1;
  x.bitmap = penalty != 1.0f ? bitmap : nullptr; x.penalty = penalty; x.penalty_first = 0;
  const int tok = full_pick(x, top_k, top_p, uniforms ? uniforms[r] : 0.f, hist + static_cast<size_t>(r) * kFullBuckets, sm);
  if (threadIdx.x == 0) out[r] = tok;
}

size_t sample_full_scratch_bytes() { return kFullScratchPerRow; }

int sample_tokens(const bf16* logits, long long ld, int N, int V, float temperature, const uint32_t* prev_bitmap,
                  float penalty, float top_p, int top_k, int do_sample, const float* uniforms, int64_t* out,
                  cudaStream_t s, void* scratch, long long scratch_bytes) {
  MT_CHECK(N > 0 && V > 0, "sample_tokens: empty input");
  MT_CHECK(temperature > 0.f, "temperature must be > 0");
  if (do_sample && (top_k <= 0 || top_k > kMaxTopK)) {
    if (top_k <= 0 && V <= kMaxTopK) {
      top_k = V;                                           // keeping every entry of a short row IS no filter
    } else {
      // no top-k filter (inference_utils.py:135) or a top-k beyond the shared-memory list: counting sort over the row
      if (!scratch || scratch_bytes < static_cast<long long>(N * kFullScratchPerRow))
        return fail(MT_ERR_ARG, "sampling with top_k = %d over %d logits needs %lld bytes of scratch (mt_sample_tokens_full)",
                    top_k, V, static_cast<long long>(N * kFullScratchPerRow));
      sample_full_kernel<<<N, kSampThreads, 0, s>>>(logits, ld, N, V, temperature, prev_bitmap, penalty, top_p, top_k, uniforms,
                                                     out, reinterpret_cast<uint32_t*>(scratch));
      count_launch();
      MT_LAUNCH_CHECK();
      return MT_OK;
    }
  }
  if (V <= kWarpMaxV && (!do_sample || top_k <= kWarpMaxK))
    sample_tokens_kernel<true><<<cdiv(N, kSampThreads / 32), kSampThreads, 0, s>>>(logits, ld, N, V, temperature, prev_bitmap, penalty,
                                                                                   top_p, top_k, do_sample, uniforms, out);
  else
    sample_tokens_kernel<false><<<N, kSampThreads, 0, s>>>(logits, ld, N, V, temperature, prev_bitmap, penalty, top_p, top_k,
                                                           do_sample, uniforms, out);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// MossTTSLocal (reference moss_tts_local/modeling_moss_tts.py:316-477): one channel of one frame per launch.
//   local_sample  : logits [B,V] of channel c -> next_ids[b][c]; greedy channels use the raw logits (:361-362),
//                   sampled ones RepetitionPenalty (per row, over that row's history of the channel, c > 0)
//                   -> Temperature -> TopK -> TopP (:359-370) and the inverse-CDF draw of this library
//   local_commit  : EOS / pad substitution for finished rows (:430-433), history append, stop flags
// ---------------------------------------------------------------------------------------------
template <bool W>
__global__ void __launch_bounds__(kSampThreads, 3)
local_sample_kernel(LocalState st, const bf16* __restrict__ logits, int V, int c, int do_sample, float temperature,
                    float penalty, float top_p, int top_k, unsigned seed) {
  __shared__ SampleShared sm;
  pdl_trigger();
  pdl_wait();
  if (*st.done) return;
  const int warp = threadIdx.x >> 5;
  const int b = W ? blockIdx.x * (kSampThreads / 32) + warp : blockIdx.x;
  if (b >= st.B) return;
  RowXform x;
  x.row = logits + static_cast<long long>(b) * V; x.V = V; x.mode = 0; x.nban = 0; x.keep0 = x.keep1 = -1;
  x.temp = do_sample ? temperature : 1.f; x.bitmap = nullptr; x.penalty = 1.f; x.penalty_first = 1;
  if (c > 0) { x.ban[0] = V - 1; x.nban = 1; }                 // audio pad code (:411-412)
  if (do_sample && c > 0 && penalty != 1.0f) {
    x.bitmap = st.seen + (static_cast<long long>(b) * (st.C - 1) + (c - 1)) * st.bitmap_words;
    x.penalty = penalty;
  }
  const float u = mt_uniform(seed, *st.step, b, c);
  const int tok = W ? warp_pick(x, do_sample, top_k, top_p, u, sm.warp[warp]) : block_pick(x, do_sample, top_k, top_p, u, sm.blk);
  if (W ? (threadIdx.x & 31) == 0 : threadIdx.x == 0) st.next_ids[static_cast<long long>(b) * st.C + c] = tok;
}

__global__ void local_commit_kernel(LocalState st, int eos, int pad_code, int n_channels_sampled) {
  pdl_trigger();
  pdl_wait();
  if (*st.done) return;
  const int len = *st.len;
  __shared__ int any_unfinished;
  if (threadIdx.x == 0) any_unfinished = 0;
  __syncthreads();
  for (int i = threadIdx.x; i < st.B * st.C; i += blockDim.x) {
    const int b = i / st.C, c = i % st.C;
    long long t = c < n_channels_sampled ? st.next_ids[i] : 0;          // channels beyond n_vq_for_inference: 0 (:427-428)
    if (eos >= 0 && !st.unfinished[b]) t = c == 0 ? eos : pad_code;
    st.next_ids[i] = t;
    if (len < st.max_len) st.gen_ids[(static_cast<long long>(b) * st.max_len + len) * st.C + c] = t;
    if (c >= 1 && t >= 0 && t < st.bitmap_words * 32)
      atomicOr(&st.seen[(static_cast<long long>(b) * (st.C - 1) + (c - 1)) * st.bitmap_words + (t >> 5)], 1u << (t & 31));
  }
  __syncthreads();
  for (int b = threadIdx.x; b < st.B; b += blockDim.x) {
    uint8_t u = st.unfinished[b];
    if (eos >= 0 && st.next_ids[static_cast<long long>(b) * st.C] == eos) u = 0;
    st.unfinished[b] = u;
    if (u) atomicOr(&any_unfinished, 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    *st.len = len + 1;
    *st.pos = len;
    *st.step = *st.step + 1;
    if (!any_unfinished) *st.done = 1;
  }
}

__global__ void local_init_kernel(LocalState st, const int64_t* __restrict__ input_ids, int S0,
                                  const uint8_t* __restrict__ attn_mask) {
  const int b = blockIdx.x;
  const int64_t* src = input_ids + static_cast<long long>(b) * S0 * st.C;
  int64_t* g = st.gen_ids + static_cast<long long>(b) * st.max_len * st.C;
  for (int i = threadIdx.x; i < S0 * st.C; i += blockDim.x) {
    const int64_t v = src[i];
    if (st.gen_ids) g[i] = v;
    const int c = i % st.C;
    if (c >= 1 && v >= 0 && v < st.bitmap_words * 32)
      atomicOr(&st.seen[(static_cast<long long>(b) * (st.C - 1) + (c - 1)) * st.bitmap_words + (v >> 5)], 1u << (v & 31));
  }
  // the prompt's mask (left padding), ones for every generated frame (HF extends the mask it is given with ones)
  for (int t = threadIdx.x; t < st.max_len; t += blockDim.x)
    st.mask[static_cast<long long>(b) * st.max_len + t] = (attn_mask && t < S0) ? (attn_mask[static_cast<long long>(b) * S0 + t] ? 1 : 0) : 1;
  if (b == 0)
    for (int t = threadIdx.x; t < st.C; t += blockDim.x) st.ones[t] = 1;
  if (threadIdx.x == 0) {
    st.unfinished[b] = 1;
    if (b == 0) { *st.len = S0; *st.step = 0; *st.done = 0; *st.pos = S0 - 1; }
  }
}

size_t local_state_bytes(int B, int C, int max_len, int audio_head) {
  const int words = (audio_head + 31) / 32;
  return 256 + align_up(B, 256) + align_up(static_cast<size_t>(B) * C * 8, 256) + align_up(static_cast<size_t>(B) * max_len, 256) +
         align_up(C, 256) + align_up(static_cast<size_t>(B) * (C - 1) * words * 4, 256);
}

LocalState local_state_carve(void* buf, int B, int C, int max_len, int audio_head, int64_t* gen_ids) {
  LocalState st{};
  uint8_t* p = reinterpret_cast<uint8_t*>(buf);
  const int words = (audio_head + 31) / 32;
  st.len = reinterpret_cast<int*>(p); st.step = st.len + 1; st.done = st.len + 2; st.pos = st.len + 3; p += 256;
  st.unfinished = p; p += align_up(B, 256);
  st.next_ids = reinterpret_cast<int64_t*>(p); p += align_up(static_cast<size_t>(B) * C * 8, 256);
  st.mask = p; p += align_up(static_cast<size_t>(B) * max_len, 256);
  st.ones = p; p += align_up(C, 256);
  st.seen = reinterpret_cast<uint32_t*>(p);
  st.gen_ids = gen_ids; st.B = B; st.C = C; st.max_len = max_len; st.bitmap_words = words;
  return st;
}

int local_init(const LocalState& st, const int64_t* input_ids, int S0, cudaStream_t s, const uint8_t* attn_mask) {
  MT_CUDA(cudaMemsetAsync(st.seen, 0, static_cast<size_t>(st.B) * (st.C - 1) * st.bitmap_words * 4, s));
  MT_CUDA(launch(local_init_kernel, dim3(st.B), dim3(256), 0, s, st, input_ids, S0, attn_mask));
  count_launch();
  return MT_OK;
}

__global__ void __launch_bounds__(kSampThreads, 3)
local_sample_full_kernel(LocalState st, const bf16* __restrict__ logits, int V, int c, float temperature, float penalty,
                         float top_p, int top_k, unsigned seed, uint32_t* __restrict__ hist) {
  __shared__ FullSmem sm;
  pdl_trigger();
  pdl_wait();
  if (*st.done) return;
  const int b = blockIdx.x;
  RowXform x;
  x.row = logits + static_cast<long long>(b) * V; x.V = V; x.mode = 0; x.nban = 0; x.keep0 = x.keep1 = -1;
  x.temp = temperature; x.bitmap = nullptr; x.penalty = 1.f; x.penalty_first = 1;
  if (c > 0) { x.ban[0] = V - 1; x.nban = 1; }
  if (c > 0 && penalty != 1.0f) {
    x.bitmap = st.seen + (static_cast<long long>(b) * (st.C - 1) + (c - 1)) * st.bitmap_words;
    x.penalty = penalty;
  }
  const int tok = full_pick(x, top_k, top_p, mt_uniform(seed, *st.step, b, c), hist + static_cast<size_t>(b) * kFullBuckets, sm);
  if (threadIdx.x == 0) st.next_ids[static_cast<long long>(b) * st.C + c] = tok;
}

int local_sample(const LocalState& st, const bf16* logits, int V, int c, int do_sample, float temperature, float penalty,
                 float top_p, int top_k, unsigned seed, cudaStream_t s, uint32_t* full_hist) {
  if (do_sample && top_k <= 0 && V <= kMaxTopK) top_k = V;        // no top-k warper on a short row: keep everything
  if (do_sample && (top_k <= 0 || top_k > kMaxTopK)) {
    // no TopKLogitsWarper (or one beyond the candidate list) on a long row: full-vocabulary counting sort
    if (!full_hist) return fail(MT_ERR_UNSUPPORTED, "sampling requires 1 <= top_k <= %d (got %d)", kMaxTopK, top_k);
    MT_CUDA(launch(local_sample_full_kernel, dim3(st.B), dim3(kSampThreads), 0, s, st, logits, V, c, temperature, penalty, top_p,
                   top_k, seed, full_hist));
    count_launch();
    return MT_OK;
  }
  if (V <= kWarpMaxV && (!do_sample || top_k <= kWarpMaxK))
    MT_CUDA(launch(local_sample_kernel<true>, dim3(cdiv(st.B, kSampThreads / 32)), dim3(kSampThreads), 0, s, st, logits, V, c,
                   do_sample, temperature, penalty, top_p, top_k, seed));
  else
    MT_CUDA(launch(local_sample_kernel<false>, dim3(st.B), dim3(kSampThreads), 0, s, st, logits, V, c, do_sample, temperature,
                   penalty, top_p, top_k, seed));
  count_launch();
  return MT_OK;
}

int local_commit(const LocalState& st, int eos, int pad_code, int n_channels_sampled, cudaStream_t s) {
  MT_CUDA(launch(local_commit_kernel, dim3(1), dim3(1024), 0, s, st, eos, pad_code, n_channels_sampled));
  count_launch();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// find_last_equal_C: one warp per row, scanning from the end
// ---------------------------------------------------------------------------------------------
__global__ void find_last_equal_kernel(const int64_t* __restrict__ t, int B, int S, long long ld, long long value,
                                       int64_t* __restrict__ out) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= B) return;
  int best = -1;
  for (int i = lane; i < S; i += 32)
    if (t[row * ld + i] == value) best = i;       // ascending per lane: the last hit survives
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (lane == 0) out[row] = best;
}
int find_last_equal(const int64_t* t, int B, int S, long long ld, long long value, int64_t* out, cudaStream_t s) {
  MT_CHECK(B > 0, "find_last_equal: empty batch");
  find_last_equal_kernel<<<cdiv(B * 32, 128), 128, 0, s>>>(t, B, S, ld, value, out);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

// ---------------------------------------------------------------------------------------------
// delay pattern: out[t+i, i] = codes[t, i], pad elsewhere; de-delay: out[t, i] = in[t+i, i]
// ---------------------------------------------------------------------------------------------
__global__ void delay_kernel(const int64_t* __restrict__ codes, int T, int n, long long pad, int64_t* __restrict__ out) {
  const long long total = static_cast<long long>(T + n - 1) * n;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(idx / n), i = static_cast<int>(idx % n);
    const int t = r - i;
    out[idx] = (t >= 0 && t < T) ? codes[static_cast<long long>(t) * n + i] : pad;
  }
}
__global__ void dedelay_kernel(const int64_t* __restrict__ in, int T, int n, int64_t* __restrict__ out) {
  const long long total = static_cast<long long>(T) * n;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int t = static_cast<int>(idx / n), i = static_cast<int>(idx % n);
    out[idx] = in[static_cast<long long>(t + i) * n + i];
  }
}
int apply_delay_pattern(const int64_t* codes, int T, int n, long long pad, int64_t* out, cudaStream_t s) {
  MT_CHECK(T >= 0 && n > 0, "apply_delay_pattern: bad shape");
  const long long total = static_cast<long long>(T + n - 1) * n;
  if (total <= 0) return MT_OK;
  delay_kernel<<<static_cast<unsigned>(min(cdivll(total, 256), 4096LL)), 256, 0, s>>>(codes, T, n, pad, out);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}
int apply_de_delay_pattern(const int64_t* delayed, int Td, int n, int64_t* out, cudaStream_t s) {
  const int T = Td - n + 1;
  MT_CHECK(n > 0, "apply_de_delay_pattern: bad shape");
  if (T <= 0) return MT_OK;
  const long long total = static_cast<long long>(T) * n;
  dedelay_kernel<<<static_cast<unsigned>(min(cdivll(total, 256), 4096LL)), 256, 0, s>>>(delayed, T, n, out);
  count_launch();
  MT_LAUNCH_CHECK();
  return MT_OK;
}

}  // namespace mt
