// Full-vocabulary sampling: sample_token (reference moss_tts_delay/inference_utils.py:111-145) with NO top-k filter
// (`top_k is None or <= 0`, :135) or a top-k beyond the 1024 candidates the shared-memory paths hold.
//
// A bf16 logit has at most 65,536 distinct values, so the (value descending, index ascending) order the reference's
// torch.sort / torch.topk produce is a COUNTING sort: one pass builds a 65,536-bucket histogram of the transformed
// keys in global scratch (256 KB per row, L2-resident), and everything sample_token needs is a prefix computation over
// buckets in descending key order, because all elements of a bucket share one value and hence one probability:
//   top-k        the bucket where the running count reaches k; ties at the k-th value keep the lowest indices
//   softmax      Z = sum_b kept_b * exp(v_b - max);  p_b = bf16(exp(v_b - max) / Z)
//   top-p        element r of bucket b is removed iff bf16(C_b + r * p_b) > bf16(top_p), C_b = probability mass of the
//                buckets above (the reference removes entry j when the rounded cumulative sum of the entries before it
//                exceeds top_p, :52-55); the first entry always stays
//   draw         second softmax over what is left, inverse CDF at u * total: bucket, then rank r inside the bucket
//   index        the r-th lowest index whose key equals the bucket (one ordered pass over the row)
// Bucket sums stand in for the reference's element-by-element fp32 sums (count * p instead of p added count times);
// the two differ by fp32 rounding only, which can move a decision only when a cumulative sum lands within ~1e-6 of a
// boundary.  One block per row; thread t owns the 65536 / blockDim keys [hi_t - per + 1, hi_t] in descending order.
#pragma once
#include "sampling_select.cuh"

namespace mt {

constexpr int kFullBuckets = 65536;
constexpr size_t kFullScratchPerRow = static_cast<size_t>(kFullBuckets) * sizeof(uint32_t);

struct FullSmem {
  float fscan[kSampThreads + 1];
  int iscan[kSampThreads + 1];
  int misc[8];
  float fmisc[4];
};

// exclusive scans over the block in thread order, done by one thread (fixed order: deterministic)
__device__ inline float block_excl_scan_f(float v, FullSmem& sm, float* total) {
  sm.fscan[threadIdx.x] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    float run = 0.f;
    for (int t = 0; t < static_cast<int>(blockDim.x); ++t) { const float x = sm.fscan[t]; sm.fscan[t] = run; run += x; }
    sm.fscan[blockDim.x] = run;
  }
  __syncthreads();
  const float r = sm.fscan[threadIdx.x];
  *total = sm.fscan[blockDim.x];
  __syncthreads();
  return r;
}
__device__ inline int block_excl_scan_i(int v, FullSmem& sm, int* total) {
  sm.iscan[threadIdx.x] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0;
    for (int t = 0; t < static_cast<int>(blockDim.x); ++t) { const int x = sm.iscan[t]; sm.iscan[t] = run; run += x; }
    sm.iscan[blockDim.x] = run;
  }
  __syncthreads();
  const int r = sm.iscan[threadIdx.x];
  *total = sm.iscan[blockDim.x];
  __syncthreads();
  return r;
}

// hist: this row's kFullBuckets counters (global).  Returns the sampled index (0 when nothing is finite, like argmax of an
// all -inf row).  top_k <= 0: no top-k filter.
__device__ int full_pick(const RowXform& x, int top_k, float top_p, float u, uint32_t* __restrict__ hist, FullSmem& sm) {
  const int tid = threadIdx.x, nt = blockDim.x;
  const int per = kFullBuckets / nt;
  for (int i = tid; i < kFullBuckets; i += nt) hist[i] = 0u;
  __syncthreads();
  for (int i = tid; i < x.V; i += nt) {
    const uint32_t k = key16(x.get(i));
    if (k > kKeyNegInf) atomicAdd(&hist[k], 1u);          // -inf and NaN never survive the softmax
  }
  __syncthreads();
  const int hi = kFullBuckets - 1 - tid * per;            // this thread's keys: hi, hi-1, ..., hi-per+1
  // ---- counts -> top-k threshold
  int cnt_local = 0;
  for (int j = 0; j < per; ++j) cnt_local += static_cast<int>(hist[hi - j]);
  int total_cnt = 0;
  const int cnt_before = block_excl_scan_i(cnt_local, sm, &total_cnt);
  if (total_cnt == 0) return 0;
  const int k = (top_k > 0 && top_k < total_cnt) ? top_k : total_cnt;
  if (tid == 0) { sm.misc[0] = -1; sm.misc[1] = 0; sm.misc[2] = -1; }
  __syncthreads();
  if (cnt_before < k && cnt_before + cnt_local >= k) {
    int acc = cnt_before;
    for (int j = 0; j < per; ++j) {
      const int c = static_cast<int>(hist[hi - j]);
      if (c && acc + c >= k) { sm.misc[0] = hi - j; sm.misc[1] = k - acc; break; }   // threshold key, entries kept at it
      acc += c;
    }
  }
  if (cnt_local > 0 && cnt_before == 0) {                  // the row maximum: first non-empty bucket from the top
    for (int j = 0; j < per; ++j)
      if (hist[hi - j]) { sm.misc[2] = hi - j; break; }
  }
  __syncthreads();
  const int thr = sm.misc[0], need_eq = sm.misc[1];
  const float vmax = unkey16(static_cast<uint32_t>(sm.misc[2]));
  __syncthreads();
  auto kept1 = [&](int key) -> int {
    const int c = static_cast<int>(hist[key]);
    return key > thr ? c : (key == thr ? min(c, need_eq) : 0);
  };
  // ---- first softmax
  float m_local = 0.f;
  for (int j = 0; j < per; ++j) {
    const int key = hi - j;
    const int c = kept1(key);
    if (c) m_local += static_cast<float>(c) * expf(unkey16(static_cast<uint32_t>(key)) - vmax);
  }
  float Z1 = 0.f;
  block_excl_scan_f(m_local, sm, &Z1);
  // ---- top-p cut: bucket and number of its entries that stay
  int cut_key = thr, cut_keep = need_eq;                   // default: everything the top-k kept
  if (top_p < 1.0f) {
    const float thr_p = rbf(top_p);
    float p_local = 0.f;
    for (int j = 0; j < per; ++j) {
      const int key = hi - j;
      const int c = kept1(key);
      if (c) p_local += static_cast<float>(c) * rbf(expf(unkey16(static_cast<uint32_t>(key)) - vmax) / Z1);
    }
    float P = 0.f;
    const float c_before = block_excl_scan_f(p_local, sm, &P);
    if (tid == 0) { sm.misc[3] = 0x7fffffff; sm.misc[4] = 0; }
    __syncthreads();
    // the cut is the first entry (descending order) that sees a rounded cumulative sum > top_p.  A chunk can only hold
    // such an entry when even the sum after its last entry exceeds top_p; the lowest thread that finds one owns the cut.
    int my_key = -1, my_keep = 0;
    if (p_local > 0.f && rbf(c_before + p_local) > thr_p) {
      float run = c_before;
      int seen = cnt_before;                               // entries before this bucket (0 means the row's first entry)
      for (int j = 0; j < per; ++j) {
        const int key = hi - j;
        const int c = kept1(key);
        if (!c) continue;
        const float p = rbf(expf(unkey16(static_cast<uint32_t>(key)) - vmax) / Z1);
        // entry r of the bucket stays iff it is the very first entry or !(bf16(run + r * p) > top_p)
        auto stays = [&](int r) { return (seen == 0 && r == 0) || !(rbf(run + static_cast<float>(r) * p) > thr_p); };
        if (stays(c - 1)) { run += static_cast<float>(c) * p; seen += c; continue; }
        int lo = 0, hi_r = c - 1;                          // smallest r that does not stay
        while (lo < hi_r) {
          const int mid = (lo + hi_r) >> 1;
          if (stays(mid)) lo = mid + 1; else hi_r = mid;
        }
        my_key = key; my_keep = lo;
        break;
      }
      if (my_key >= 0) atomicMin(&sm.misc[3], tid);
    }
    __syncthreads();
    if (sm.misc[3] == tid) { sm.misc[4] = my_key; sm.misc[5] = my_keep; }
    __syncthreads();
    if (sm.misc[3] != 0x7fffffff) {
      cut_key = sm.misc[4]; cut_keep = sm.misc[5];
      if (cut_keep == 0) {                                 // nothing of that bucket stays: the cut is the bucket above
        cut_key += 1; cut_keep = 0x7fffffff;               // (keys above are all kept in full)
      }
    }
    __syncthreads();
  }
  auto kept2 = [&](int key) -> int {
    const int c = kept1(key);
    return key > cut_key ? c : (key == cut_key ? min(c, cut_keep) : 0);
  };
  // ---- second softmax + inverse-CDF draw
  float m2_local = 0.f;
  for (int j = 0; j < per; ++j) {
    const int key = hi - j;
    const int c = kept2(key);
    if (c) m2_local += static_cast<float>(c) * expf(unkey16(static_cast<uint32_t>(key)) - vmax);
  }
  float Z2 = 0.f;
  block_excl_scan_f(m2_local, sm, &Z2);
  float q_local = 0.f;
  int n2_local = 0;
  for (int j = 0; j < per; ++j) {
    const int key = hi - j;
    const int c = kept2(key);
    if (c) { q_local += static_cast<float>(c) * rbf(expf(unkey16(static_cast<uint32_t>(key)) - vmax) / Z2); n2_local += c; }
  }
  float total = 0.f;
  const float q_before = block_excl_scan_f(q_local, sm, &total);
  int n2_total = 0;
  const int n2_before = block_excl_scan_i(n2_local, sm, &n2_total);
  const float target = u * total;
  if (tid == 0) { sm.misc[5] = -1; sm.misc[6] = 0; }
  __syncthreads();
  // the drawn entry is the first whose running sum exceeds the target: exactly one chunk's [q_before, q_before + q_local)
  // holds the target (the scan values are the same floats on both sides of a chunk boundary); if rounding leaves the
  // target at or past the total, the last kept entry is drawn, as in draw_sorted
  const bool last_owner = n2_local > 0 && n2_before + n2_local == n2_total;
  const bool crosses = q_before + q_local > target && !(q_before > target);
  if (n2_local > 0 && (crosses || (last_owner && !(q_before + q_local > target)))) {
    float run = q_before;
    int last_key = -1, last_c = 0;
    bool found = false;
    for (int j = 0; j < per && !found; ++j) {
      const int key = hi - j;
      const int c = kept2(key);
      if (!c) continue;
      const float q = rbf(expf(unkey16(static_cast<uint32_t>(key)) - vmax) / Z2);
      last_key = key; last_c = c;
      if (run + static_cast<float>(c) * q > target) {
        int lo = 0, hi_r = c - 1;                          // smallest r with run + (r + 1) q > target
        while (lo < hi_r) {
          const int mid = (lo + hi_r) >> 1;
          if (run + static_cast<float>(mid + 1) * q > target) hi_r = mid; else lo = mid + 1;
        }
        sm.misc[5] = key; sm.misc[6] = lo;
        found = true;
      }
      run += static_cast<float>(c) * q;
    }
    if (!found) { sm.misc[5] = last_key; sm.misc[6] = last_c - 1; }     // the chunk's own last entry (rounding only)
  }
  __syncthreads();
  const int pick_key = sm.misc[5], pick_rank = sm.misc[6];
  __syncthreads();
  if (pick_key < 0) return 0;
  // ---- the pick_rank-th lowest index whose key is pick_key
  const int seg = (x.V + nt - 1) / nt;
  const int i0 = tid * seg, i1 = min(x.V, i0 + seg);
  int n_eq = 0;
  for (int i = i0; i < i1; ++i) n_eq += key16(x.get(i)) == static_cast<uint32_t>(pick_key);
  int tot_eq = 0;
  int off = block_excl_scan_i(n_eq, sm, &tot_eq);
  if (tid == 0) sm.misc[7] = 0;
  __syncthreads();
  if (off <= pick_rank && pick_rank < off + n_eq) {
    for (int i = i0; i < i1; ++i)
      if (key16(x.get(i)) == static_cast<uint32_t>(pick_key)) {
        if (off == pick_rank) { sm.misc[7] = i; break; }
        ++off;
      }
  }
  __syncthreads();
  const int r = sm.misc[7];
  __syncthreads();
  return r;
}

}  // namespace mt
