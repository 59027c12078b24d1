// One-pass top-k selection over a bf16 logit row, for a thread group that is either a whole block (long rows: the
// text vocabulary) or one warp (short rows: a 1025-entry audio codebook head; eight rows per block).
//
// The row is streamed exactly once with 16-byte loads.  Two bf16 logits are turned into two sortable 16-bit keys
// with four integer instructions on the packed word (or, when a repetition penalty makes the per-element transform
// non-monotone, element by element through RowXform::xf), a packed VIMNMX.U16x2 max per 8 elements decides whether
// any of them can still be a candidate, and the rare survivors are appended to a shared-memory list:
//
//   chunk 0   32 elements per thread stay in registers; the (k + #banned)-th largest of the group's
//             2 * threads "half-stream" maxima is a threshold T that at least k + #banned elements reach
//   floor     the temperature division merges neighbouring bf16 values: the list threshold R is the lowest raw key
//             whose transformed value equals that of T, so everything that could tie with the k-th value is kept
//   stream    the remaining chunks append elements >= R; when the list passes kRaise entries, T is raised to the
//             (k + #banned)-th largest listed key and the list is compacted
//   final     listed keys -> transformed keys ((0xFFFF - key) << 32 | index), banned / -inf entries dropped,
//             bitonic sort: value descending, index ascending -- the order sample_token's top-k needs
//             (reference inference_utils.py:111-145; ties: lowest index first, as oracle/sampling.py)
//
// Rows the fast path cannot finish (more ties than the list holds, NaNs, k beyond 2 * threads) return -1 and the
// caller runs the exact multi-pass selection instead (row_sample / row_argmax, or warp_exact_select).
#pragma once
#include "sampling.cuh"

namespace mt {

constexpr int kSampThreads = 256;

// ---- per-row logit transform -----------------------------------------------------------------
struct RowXform {
  const bf16* row; int V;
  float temp;                 // value = bf16(l / temp)
  int mode;                   // 0 plain, 1 text outside audio, 2 text inside audio
  int ban[6]; int nban;       // ids forced to -inf (all modes)
  int keep0, keep1;           // mode 2: everything else is -inf
  const uint32_t* bitmap; float penalty;   // repetition penalty on ids whose bit is set
  int penalty_first;          // HF processor order (MossTTSLocal): penalty on the raw logit, then temperature
  __device__ __forceinline__ bool banned(int i) const {
#pragma unroll
    for (int j = 0; j < 6; ++j)
      if (j < nban && i == ban[j]) return true;
    return false;
  }
  // transformed value of element i whose raw logit is v
  __device__ __forceinline__ float xf(float v, int i) const {
    if (penalty_first) {
      if (bitmap && (bitmap[i >> 5] >> (i & 31) & 1u)) v = v < 0.f ? rbf(v * penalty) : rbf(v / penalty);
      v = rbf(v / temp);
      return banned(i) ? -INFINITY : v;
    }
    v = rbf(v / temp);
    if (mode == 2 && i != keep0 && i != keep1) return -INFINITY;
    if (banned(i)) return -INFINITY;
    if (bitmap && (bitmap[i >> 5] >> (i & 31) & 1u)) v = v > 0.f ? rbf(v / penalty) : rbf(v * penalty);
    return v;
  }
  __device__ __forceinline__ float get(int i) const { return xf(bf2f(row[i]), i); }
};

// ---- sortable 16-bit keys ---------------------------------------------------------------------
__device__ __forceinline__ uint32_t key16(float v) {
  if (v != v) return 0u;                                  // NaN sorts below everything
  const uint32_t b = __float_as_uint(v) >> 16;            // v is bf16-representable
  return (b & 0x8000u) ? (~b & 0xFFFFu) : (b | 0x8000u);
}
__device__ __forceinline__ float unkey16(uint32_t k) {
  const uint32_t b = (k & 0x8000u) ? (k & 0x7FFFu) : (~k & 0xFFFFu);
  return __uint_as_float(b << 16);
}
constexpr uint32_t kKeyNegInf = 0x007Fu;
constexpr uint32_t kKeyPosInf = 0xFF80u;
// keys of the two raw bf16 halves of a word (same map as key16, except that NaNs land above +inf / below -inf)
__device__ __forceinline__ uint32_t rawkeys2(uint32_t w) {
  const uint32_t s = (w >> 15) & 0x00010001u;
  return w ^ (s * 0x7FFFu | 0x80008000u);
}

// ---- thread groups ------------------------------------------------------------------------------
struct BlockG {
  static constexpr int kSize = kSampThreads;
  static constexpr int kCap = 2048;      // list entries
  static constexpr int kRaise = 128;     // compact the list when it is longer than this at a chunk boundary
  __device__ static int tid() { return threadIdx.x; }
  __device__ static void sync() { __syncthreads(); }
  __device__ static int sum(int v, int* scratch) {
    v = __reduce_add_sync(0xffffffffu, v);
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    __syncthreads();
    int t = 0;
#pragma unroll
    for (int w = 0; w < kSize / 32; ++w) t += scratch[w];
    __syncthreads();
    return t;
  }
  __device__ static unsigned long long min64(unsigned long long v, unsigned long long* scratch) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long q = __shfl_xor_sync(0xffffffffu, v, o);
      v = q < v ? q : v;
    }
    if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
    __syncthreads();
    unsigned long long t = scratch[0];
#pragma unroll
    for (int w = 1; w < kSize / 32; ++w) t = scratch[w] < t ? scratch[w] : t;
    __syncthreads();
    return t;
  }
};
struct WarpG {
  static constexpr int kSize = 32;
  static constexpr int kCap = 256;
  static constexpr int kRaise = 128;
  __device__ static int tid() { return threadIdx.x & 31; }
  __device__ static void sync() { __syncwarp(); }
  __device__ static int sum(int v, int*) { return __reduce_add_sync(0xffffffffu, v); }
  __device__ static unsigned long long min64(unsigned long long v, unsigned long long*) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long q = __shfl_xor_sync(0xffffffffu, v, o);
      v = q < v ? q : v;
    }
    return v;
  }
};

template <class G>
struct SelSmem {
  unsigned long long list[G::kCap];
  unsigned long long red[G::kSize / 32 + 1];
  float ebuf[2 * (G::kCap < kMaxTopK ? 64 : kMaxTopK) + 1];   // exp(v_j - max) and probabilities of the sorted candidates
  int scratch[G::kSize / 32 + 1];
  int n;            // entries appended so far (may run past kCap: then `flag` is set)
  int flag;         // the fast path gave up on this row
  uint32_t thr;     // current list threshold R
  uint32_t kth;     // kth_listed broadcast
  int result;
};

// lowest key whose transformed value equals that of key t (the transform is monotone, so the class is an interval)
__device__ inline uint32_t class_floor(uint32_t t, float temp, int* flag) {
  if (temp == 1.0f || t == 0u) return t;
  const uint32_t ft = key16(rbf(unkey16(t) / temp));
  uint32_t r = t;
  for (int it = 0; it < 4096 && r > 0u; ++it) {
    if (key16(rbf(unkey16(r - 1u) / temp)) != ft) return r;
    --r;
  }
  if (r > 0u) *flag = 1;     // a temperature that merges > 4096 neighbouring values: not worth a fast path
  return r;
}

// packed bf16 pair: maximum (a NaN operand is ignored) and >= (0xFFFF per true half, false for NaN)
__device__ __forceinline__ uint32_t bmax2(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ uint32_t bge2(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("set.ge.u32.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
constexpr uint32_t kNegInf2 = 0xFF80FF80u;               // (-inf, -inf)
// key of one raw bf16 (low 16 bits of `half`) with key16's NaN rule
__device__ __forceinline__ uint32_t rawkey(uint32_t half) {
  const uint32_t k = rawkeys2(half) & 0xFFFFu;
  return k > kKeyPosInf ? 0u : k;
}

// kk-th largest key (high words) among sm.list[0 .. n): a 16-step bisection run by the first warp of the group on
// registers (n <= 32 * NR) or on the shared-memory list; every thread of the group gets the result.  0 if n < kk.
template <class G, int NR>
__device__ uint32_t kth_listed(SelSmem<G>& sm, int n, int kk) {
  G::sync();
  const int lane = G::tid();
  if (lane < 32) {
    uint32_t prefix = 0u;
    if (n <= 32 * NR) {
      uint32_t kr[NR];
#pragma unroll
      for (int j = 0; j < NR; ++j) kr[j] = lane + 32 * j < n ? static_cast<uint32_t>(sm.list[lane + 32 * j] >> 32) : 0u;
      if (kk == 1) {
        uint32_t mx = 0u;
#pragma unroll
        for (int j = 0; j < NR; ++j) mx = max(mx, kr[j]);
        prefix = __reduce_max_sync(0xffffffffu, mx);
      } else {
        for (int bit = 15; bit >= 0; --bit) {
          const uint32_t thr = prefix | (1u << bit);
          int c = 0;
#pragma unroll
          for (int j = 0; j < NR; ++j) c += kr[j] >= thr;           // absent entries have key 0 < thr
          if (__reduce_add_sync(0xffffffffu, c) >= kk) prefix = thr;
        }
      }
    } else {
      for (int bit = 15; bit >= 0; --bit) {
        const uint32_t thr = prefix | (1u << bit);
        int c = 0;
        for (int i = lane; i < n; i += 32) c += static_cast<uint32_t>(sm.list[i] >> 32) >= thr;
        if (__reduce_add_sync(0xffffffffu, c) >= kk) prefix = thr;
      }
    }
    if (lane == 0) sm.kth = prefix;
  }
  G::sync();
  return sm.kth;
}

// bitonic sort of sm.list[0 .. n2) (n2 a power of two), ascending on ((0xFFFF - key) << 32 | index) == value descending,
// index ascending; returns the number of valid entries among the first k
template <class G>
__device__ int sort_count(SelSmem<G>& sm, int n2, int k) {
  constexpr int GS = G::kSize;
  const int tid = G::tid();
  for (int size = 2; size <= n2; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = tid; i < n2; i += GS) {
        const int j = i ^ stride;
        if (j > i) {
          const unsigned long long a = sm.list[i], b = sm.list[j];
          const bool up = (i & size) == 0;
          if ((a > b) == up) { sm.list[i] = b; sm.list[j] = a; }
        }
      }
      G::sync();
    }
  }
  int c = 0;
  for (int i = tid; i < min(k, n2); i += GS) c += sm.list[i] != ~0ull;
  return G::sum(c, sm.scratch);
}

// One warp, at most 64 entries with indices below 65536: the same order on 32-bit entries ((0xFFFF - key) << 16 | index),
// two per lane, sorted in registers with shuffles; the sorted entries go back to sm.list in the 64-bit format.
__device__ inline int warp_sort64(SelSmem<WarpG>& sm, uint32_t e0, uint32_t e1, int k) {
  const int lane = threadIdx.x & 31;
  uint32_t e[2] = {e0, e1};
#pragma unroll
  for (int size = 2; size <= 64; size <<= 1) {
#pragma unroll
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      if (stride == 32) {
        const uint32_t lo = min(e[0], e[1]), hi = max(e[0], e[1]);
        e[0] = lo; e[1] = hi;
      } else {
#pragma unroll
        for (int r = 0; r < 2; ++r) {
          const uint32_t p = __shfl_xor_sync(0xffffffffu, e[r], stride);
          const bool up = ((lane + 32 * r) & size) == 0;
          const bool lower = (lane & stride) == 0;
          e[r] = (lower == up) ? min(e[r], p) : max(e[r], p);
        }
      }
    }
  }
  int c = 0;
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const bool valid = e[r] != 0xFFFFFFFFu;
    sm.list[lane + 32 * r] = valid ? (static_cast<unsigned long long>(e[r] >> 16) << 32) | (e[r] & 0xFFFFu) : ~0ull;
    c += valid && lane + 32 * r < k;
  }
  __syncwarp();
  return __reduce_add_sync(0xffffffffu, c);
}

// Sorted (value descending, index ascending) top-k candidates of the row in sm.list[0 .. return value); -1 = give up.
// k == 1 skips the sort and leaves the single best entry in sm.list[0].
// FAST (no repetition penalty: the transform is monotone): the stream works on the raw bf16 words -- a packed
// HMNMX2.BF16 maximum per 8 elements and one packed compare against the threshold; keys exist only for survivors.
// Otherwise every element goes through RowXform::xf and the stream compares packed 16-bit keys.
template <class G, bool FAST>
__device__ int select_topk(const RowXform& x, int k, SelSmem<G>& sm) {
  constexpr int GS = G::kSize, CAP = G::kCap;
  const int tid = G::tid();
  const int kk = k + (FAST ? x.nban : 0);
  if (kk > 2 * GS || x.mode == 2) return -1;
  const int raise_at = max(G::kRaise, 2 * kk);
  const int V = x.V;
  const uintptr_t addr = reinterpret_cast<uintptr_t>(x.row);
  const int head = min(V, static_cast<int>(((16u - static_cast<unsigned>(addr & 15u)) & 15u) >> 1));
  const int nvec = (V - head) >> 3;
  const int tail0 = head + 8 * nvec;
  const uint4* vrow = reinterpret_cast<const uint4*>(x.row + head);

  // the four vectors of chunk c owned by this thread: raw words (FAST) or packed keys; absent vectors = lowest
  auto load_chunk = [&](int c, uint4 (&w)[4]) {
    const int v0 = c * 4 * GS + tid;
    if ((c + 1) * 4 * GS <= nvec) {
#pragma unroll
      for (int j = 0; j < 4; ++j) w[j] = __ldg(vrow + v0 + j * GS);
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        w[j] = v0 + j * GS < nvec ? __ldg(vrow + v0 + j * GS) : make_uint4(kNegInf2, kNegInf2, kNegInf2, kNegInf2);
    }
  };
  auto to_keys = [&](const uint4& w, int vec, uint32_t (&kw)[4]) {
    const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
    if (FAST) {
#pragma unroll
      for (int q = 0; q < 4; ++q) kw[q] = ww[q];
    } else {
      const int i0 = head + 8 * vec;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        uint32_t lo = 0u, hi = 0u;
        if (vec < nvec) {
          lo = key16(x.xf(__uint_as_float(ww[q] << 16), i0 + 2 * q));
          hi = key16(x.xf(__uint_as_float(ww[q] & 0xFFFF0000u), i0 + 2 * q + 1));
        }
        kw[q] = lo | (hi << 16);
      }
    }
  };
  auto append = [&](uint32_t sk, int idx) {
    const int slot = atomicAdd(&sm.n, 1);
    if (slot < CAP) sm.list[slot] = (static_cast<unsigned long long>(sk) << 32) | static_cast<uint32_t>(idx);
    else sm.flag = 1;
  };
  // R: key threshold; Rv2: the same threshold as a packed bf16 pair (FAST)
  auto filter = [&](const uint32_t (&kw)[4], int vec, uint32_t R, uint32_t Rv2) {
    bool hit;
    if (FAST) {
      hit = bge2(bmax2(bmax2(kw[0], kw[1]), bmax2(kw[2], kw[3])), Rv2) != 0u;
    } else {
      const uint32_t vm = __vmaxu2(__vmaxu2(kw[0], kw[1]), __vmaxu2(kw[2], kw[3]));
      hit = (vm >> 16) >= R || (vm & 0xFFFFu) >= R;
    }
    if (hit && vec < nvec) {
      const int i0 = head + 8 * vec;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t lo = FAST ? rawkey(kw[q]) : kw[q] & 0xFFFFu;
        const uint32_t hi = FAST ? rawkey(kw[q] >> 16) : kw[q] >> 16;
        if (lo >= R) append(lo, i0 + 2 * q);
        if (hi >= R) append(hi, i0 + 2 * q + 1);
      }
    }
  };
  auto pair_of = [&](uint32_t R) -> uint32_t {            // raw bf16 bits of key R in both halves; -inf when R is below it
    const uint32_t b = R <= kKeyNegInf ? 0xFF80u : ((R & 0x8000u) ? (R & 0x7FFFu) : (~R & 0xFFFFu));
    return b | (b << 16);
  };
  // kk-th largest of the listed keys -> new threshold, compacted list
  auto raise = [&](int n) {
    constexpr int EPT = CAP / GS;
    unsigned long long e[EPT];
#pragma unroll
    for (int j = 0; j < EPT; ++j) e[j] = tid + j * GS < n ? sm.list[tid + j * GS] : 0ull;
    const uint32_t prefix = kth_listed<G, 8>(sm, n, kk);
    if (tid == 0) {
      sm.thr = FAST ? class_floor(prefix, x.temp, &sm.flag) : prefix;
      sm.n = 0;
    }
    G::sync();
    const uint32_t R = sm.thr;
#pragma unroll
    for (int j = 0; j < EPT; ++j)
      if (tid + j * GS < n && static_cast<uint32_t>(e[j] >> 32) >= R) append(static_cast<uint32_t>(e[j] >> 32), static_cast<int>(e[j] & 0xFFFFFFFFu));
    G::sync();
  };

  if (tid == 0) { sm.n = 0; sm.flag = 0; }
  // ---- chunk 0: 4 vectors per thread, threshold from the half-stream maxima
  uint32_t cur[4][4];
  {
    uint4 w[4];
    load_chunk(0, w);
#pragma unroll
    for (int j = 0; j < 4; ++j) to_keys(w[j], j * GS + tid, cur[j]);
  }
  {
    uint32_t lo, hi;
    if (FAST) {
      uint32_t m = kNegInf2;
#pragma unroll
      for (int j = 0; j < 4; ++j) m = bmax2(m, bmax2(bmax2(cur[j][0], cur[j][1]), bmax2(cur[j][2], cur[j][3])));
      lo = rawkey(m); hi = rawkey(m >> 16);
    } else {
      uint32_t m = 0u;
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int q = 0; q < 4; ++q) m = __vmaxu2(m, cur[j][q]);
      lo = m & 0xFFFFu; hi = m >> 16;
    }
    sm.list[2 * tid] = static_cast<unsigned long long>(lo) << 32;
    sm.list[2 * tid + 1] = static_cast<unsigned long long>(hi) << 32;
    const uint32_t prefix = kth_listed<G, 2 * GS / 32>(sm, 2 * GS, kk);
    if (tid == 0) sm.thr = FAST ? class_floor(prefix, x.temp, &sm.flag) : prefix;
  }
  G::sync();
  // ---- stream: filter the chunk in registers while the next one is in flight
  const int nchunks = max(1, (nvec + 4 * GS - 1) / (4 * GS));
  for (int c = 0; c < nchunks; ++c) {
    uint4 nxt[4];
    const bool more = c + 1 < nchunks;
    if (more) load_chunk(c + 1, nxt);
    const uint32_t R = sm.thr;
    const uint32_t Rv2 = pair_of(R);
#pragma unroll
    for (int j = 0; j < 4; ++j) filter(cur[j], c * 4 * GS + j * GS + tid, R, Rv2);
    if (more) {
#pragma unroll
      for (int j = 0; j < 4; ++j) to_keys(nxt[j], (c + 1) * 4 * GS + j * GS + tid, cur[j]);
    }
    G::sync();
    const int n = sm.n;
    const int give_up = sm.flag;
    G::sync();
    if (give_up) return -1;
    if (n > raise_at) {
      raise(n);
      if (sm.n > raise_at || sm.flag) return -1;       // ties: the list cannot shrink
      G::sync();
    }
  }
  // ---- the (at most 14) elements in front of / behind the 16-byte aligned body
  {
    const int n_extra = head + (V - tail0);
    const uint32_t R = sm.thr;
    if (tid < n_extra) {
      const int i = tid < head ? tid : tail0 + (tid - head);
      const float v = bf2f(x.row[i]);
      const uint32_t sk = FAST ? rawkey(__float_as_uint(v) >> 16) : key16(x.xf(v, i));
      if (sk >= R) append(sk, i);
    }
  }
  G::sync();
  if (sm.flag || sm.n > CAP) return -1;
  if (sm.n > 64 && sm.n > 2 * kk) {                      // a last raise keeps the sort short
    const int n_all = sm.n;
    G::sync();
    raise(n_all);
    if (sm.flag) return -1;
  }
  const int n = sm.n;
  // ---- listed keys -> final entries
  auto final_key = [&](unsigned long long l) -> uint32_t {
    const uint32_t sk = static_cast<uint32_t>(l >> 32);
    if (!FAST) return sk;
    return x.banned(static_cast<int>(l & 0xFFFFFFFFu)) ? 0u : key16(rbf(unkey16(sk) / x.temp));
  };
  if constexpr (G::kSize == 32) if (k > 1 && n <= 64 && V <= 65536) {
    uint32_t e32[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      e32[r] = 0xFFFFFFFFu;
      if (tid + 32 * r < n) {
        const unsigned long long l = sm.list[tid + 32 * r];
        const uint32_t fk = final_key(l);
        if (fk > kKeyNegInf) e32[r] = ((0xFFFFu - fk) << 16) | static_cast<uint32_t>(l & 0xFFFFu);
      }
    }
    G::sync();
    return warp_sort64(reinterpret_cast<SelSmem<WarpG>&>(sm), e32[0], e32[1], k);
  }
  int n2 = 1;
  while (n2 < n) n2 <<= 1;
  unsigned long long best = ~0ull;
  for (int i = tid; i < n2; i += GS) {
    unsigned long long e = ~0ull;
    if (i < n) {
      const unsigned long long l = sm.list[i];
      const uint32_t fk = final_key(l);
      if (fk > kKeyNegInf) e = (static_cast<unsigned long long>(0xFFFFu - fk) << 32) | (l & 0xFFFFFFFFull);
    }
    if (k > 1) sm.list[i] = e;
    best = e < best ? e : best;
  }
  if (k == 1) {
    best = G::min64(best, sm.red);
    if (tid == 0) sm.list[0] = best;
    G::sync();
    return best != ~0ull;
  }
  G::sync();
  return sort_count<G>(sm, n2, k);
}

// Exact selection for one warp without any assumption on the data: 16-step bisection for the k-th largest
// transformed key (the row is re-read from L1/L2 each step), then an index-ordered collection so that ties at the
// threshold go to the lowest indices.  Slow; only for rows select_topk gave up on.  k <= WarpG::kCap.
__device__ inline int warp_exact_select(const RowXform& x, int k, SelSmem<WarpG>& sm) {
  const int lane = threadIdx.x & 31, V = x.V;
  uint32_t T = 0u;
  for (int bit = 15; bit >= 0; --bit) {
    const uint32_t thr = T | (1u << bit);
    int c = 0;
    for (int i = lane; i < V; i += 32) c += key16(x.get(i)) >= thr;
    if (__reduce_add_sync(0xffffffffu, c) >= k) T = thr;
  }
  int cg = 0;
  for (int i = lane; i < V; i += 32) cg += key16(x.get(i)) > T;
  const int need_eq = k - __reduce_add_sync(0xffffffffu, cg);
  const unsigned lt = (1u << lane) - 1u;
  int n = 0, eq_taken = 0;
  for (int base = 0; base < V; base += 32) {
    const int i = base + lane;
    const uint32_t key = i < V ? key16(x.get(i)) : 0u;
    const bool gt = i < V && key > T, eq = i < V && key == T;
    const unsigned beq = __ballot_sync(0xffffffffu, eq);
    const bool take = gt || (eq && eq_taken + __popc(beq & lt) < need_eq);
    const unsigned bt = __ballot_sync(0xffffffffu, take);
    const int pos = n + __popc(bt & lt);
    if (take && pos < WarpG::kCap)
      sm.list[pos] = key > kKeyNegInf ? (static_cast<unsigned long long>(0xFFFFu - key) << 32) | static_cast<uint32_t>(i) : ~0ull;
    n += __popc(bt);
    eq_taken += __popc(beq);
  }
  n = min(n, WarpG::kCap);
  int n2 = 1;
  while (n2 < n) n2 <<= 1;
  for (int i = n + lane; i < n2; i += 32) sm.list[i] = ~0ull;
  if (n == 0 && lane == 0) sm.list[0] = ~0ull;
  __syncwarp();
  return sort_count<WarpG>(sm, n2, k);
}

// sample_token's tail on the sorted candidates: softmax -> top-p cut -> softmax -> inverse-CDF draw
// (inference_utils.py:123-145; the sums run in candidate order on one thread, as oracle/sampling.py does)
template <class G>
__device__ int draw_sorted(const unsigned long long* cand, int n, float top_p, float u, float* ebuf, int* result) {
  if (n <= 0) return 0;
  const int tid = G::tid();
  const float mx = unkey16(0xFFFFu - static_cast<uint32_t>(cand[0] >> 32));
  for (int j = tid; j < n; j += G::kSize) ebuf[j] = expf(unkey16(0xFFFFu - static_cast<uint32_t>(cand[j] >> 32)) - mx);
  G::sync();
  // the sums run in candidate order on one thread (fp32 addition is not associative); the divisions are done by all
  float* pbuf = ebuf + n;                                 // rbf(e_j / sum)
  int keep = n;
  if (top_p < 1.0f) {
    if (tid == 0) {
      float sum = 0.f;
#pragma unroll 4
      for (int j = 0; j < n; ++j) sum += ebuf[j];
      pbuf[n] = sum;
    }
    G::sync();
    const float sum = pbuf[n];
    for (int j = tid; j < n; j += G::kSize) pbuf[j] = rbf(ebuf[j] / sum);
    G::sync();
    if (tid == 0) {
      const float thr_p = rbf(top_p);
      float run = 0.f;
      int kp = 0;
      for (int j = 0; j < n; ++j) {
        // entry j is removed when the rounded cumulative sum of entries < j exceeds top_p
        if (j > 0 && rbf(run) > thr_p) break;
        ++kp;
        run += pbuf[j];
      }
      *result = kp;
    }
    G::sync();
    keep = *result;
    G::sync();
  }
  if (tid == 0) {
    float sum2 = 0.f;
#pragma unroll 4
    for (int j = 0; j < keep; ++j) sum2 += ebuf[j];
    pbuf[n] = sum2;
  }
  G::sync();
  const float sum2 = pbuf[n];
  for (int j = tid; j < keep; j += G::kSize) pbuf[j] = rbf(ebuf[j] / sum2);
  G::sync();
  if (tid == 0) {
    float total = 0.f;
#pragma unroll 4
    for (int j = 0; j < keep; ++j) total += pbuf[j];
    const float target = u * total;
    float run2 = 0.f;
    int r = static_cast<int>(cand[keep - 1] & 0xFFFFFFFFu);
    for (int j = 0; j < keep; ++j) {
      run2 += pbuf[j];
      if (run2 > target) { r = static_cast<int>(cand[j] & 0xFFFFFFFFu); break; }
    }
    *result = r;
  }
  G::sync();
  const int r = *result;
  G::sync();
  return r;
}

}  // namespace mt
