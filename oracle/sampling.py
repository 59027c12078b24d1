"""Oracle for token sampling (TEST INFRASTRUCTURE ONLY).

Restates reference moss_tts_delay/inference_utils.py:
  apply_repetition_penalty_delay_pattern :62-108  (only "case 1", the one generate() reaches)
  apply_top_k                            :19-26
  apply_top_p_optimized                  :44-59
  sample_token                           :111-145
  find_last_equal_C                      :148-165

Rows are processed one at a time with explicit loops.  The random draw is made
explicit: instead of torch.multinomial's Philox stream the caller supplies a
uniform u in [0,1) per row; the token is the first candidate, in (prob desc,
index asc) order, whose running sum exceeds u * total.  The CUDA path uses the
same rule with the same u, so sampled ids are comparable one-for-one.

Tie rule for top-k: among equal logits the lower index wins (torch.topk leaves
this unspecified).
"""
from __future__ import annotations

from typing import Iterable, Optional

import numpy as np
import torch


def penalty_set(prev_tokens: torch.Tensor) -> np.ndarray:
    """Sorted unique ids over *everything* passed in — the union over batch, time
    and channels (reference :78-80)."""
    return np.unique(prev_tokens.reshape(-1).cpu().numpy())


def apply_penalty_row(row: torch.Tensor, ids: Iterable[int], penalty: float) -> torch.Tensor:
    """logit>0 ? logit/p : logit*p on the listed columns, in the row's dtype (reference :82-86)."""
    row = row.clone()
    idx = torch.as_tensor(np.asarray(list(ids), dtype=np.int64))
    idx = idx[idx < row.numel()]
    if idx.numel() == 0 or penalty == 1.0:
        return row
    vals = row[idx]
    pos = vals > 0
    out = torch.where(pos, vals / penalty, vals * penalty)
    row[idx] = out
    return row


def apply_penalty_heads(logits: torch.Tensor, prev_tokens: torch.Tensor, penalty: float) -> torch.Tensor:
    """"Case 2" of the reference (:88-106): logits [B, H, V], prev_tokens [B, T, H] or [B, H]; head h is penalised on the
    unique ids of prev_tokens[..., h] only (union over the batch and time), on every batch row of that head."""
    out = logits.clone()
    B, H, _ = logits.shape
    for h in range(H):
        ids = penalty_set(prev_tokens[..., h])
        for b in range(B):
            out[b, h] = apply_penalty_row(out[b, h], ids, penalty)
    return out


def topk_survivors(row: torch.Tensor, k: int):
    """Indices of the k largest logits, ordered (value desc, index asc)."""
    V = row.numel()
    k = min(int(k), V)
    vals = row.float().numpy()
    order = np.lexsort((np.arange(V), -vals))  # primary: -value, secondary: index
    return order[:k]


def filtered_distribution(row: torch.Tensor, top_k: Optional[int], top_p: Optional[float]):
    """Candidate ids (value desc, index asc) and their final probabilities in the
    row's dtype after top-k -> top-p -> softmax (reference :131-144).
    Returns (ids[int64], probs[float32 view of dtype-rounded values])."""
    dtype = row.dtype
    V = row.numel()
    if top_k is not None and top_k > 0:
        cand = topk_survivors(row, top_k)
    else:
        cand = topk_survivors(row, V)
    cand = np.asarray([c for c in cand if row[c] != float("-inf")], dtype=np.int64)
    logits = row[torch.as_tensor(cand)]
    if top_p is not None and top_p < 1.0:
        # softmax in the row dtype (F.softmax on bf16 computes in fp32 and rounds once)
        probs = torch.softmax(logits.float(), dim=-1).to(dtype)
        # candidates are already sorted descending by logit == descending by prob
        # fp32 accumulate, rounded output.  Beyond the 1,024 candidates a top-k filter leaves (no top-k: the whole
        # vocabulary) a sequential fp32 sum stagnates (adding 1e-8 to 0.9 is a no-op), which torch's own scans avoid
        # on the GPU; there the restatement accumulates in float64, i.e. the sums are the true ones.
        if probs.numel() > 1024:
            csum = torch.cumsum(probs.double(), dim=-1).float().to(dtype)
        else:
            csum = torch.cumsum(probs.float(), dim=-1).to(dtype)
        thr = torch.tensor(top_p, dtype=dtype)                     # python scalar is cast to the tensor dtype
        remove = csum > thr
        remove = torch.cat([torch.zeros(1, dtype=torch.bool), remove[:-1]])
        keep = ~remove
        cand = cand[keep.numpy()]
        logits = logits[keep]
    probs = torch.softmax(logits.float(), dim=-1).to(dtype)
    return cand, probs.float().numpy()


def draw(cand: np.ndarray, probs: np.ndarray, u: float) -> int:
    """Inverse-CDF draw in candidate order with fp32 running sums (float64 beyond 1,024 candidates, where a
    sequential fp32 sum stagnates; the CUDA path's bucket sums approximate the true sums to fp32 rounding)."""
    if len(probs) > 1024:
        cum = np.cumsum(np.asarray(probs, dtype=np.float64))
        target = float(np.float32(u)) * cum[-1]
        j = int(np.searchsorted(cum, target, side="right"))
        return int(cand[min(j, len(cand) - 1)])
    total = np.float32(0)
    for p in probs:
        total = np.float32(total + np.float32(p))
    target = np.float32(np.float32(u) * total)
    run = np.float32(0)
    for c, p in zip(cand, probs):
        run = np.float32(run + np.float32(p))
        if run > target:
            return int(c)
    return int(cand[-1])


def sample_row(row: torch.Tensor, *, prev_ids=None, repetition_penalty: float = 1.0,
               top_p=None, top_k=None, do_sample: bool = True, u: float = 0.0) -> int:
    """One row of reference sample_token :111-145."""
    if prev_ids is not None and repetition_penalty != 1.0:
        row = apply_penalty_row(row, prev_ids, repetition_penalty)
    if not do_sample:
        return int(torch.argmax(row.float()).item())   # first max index on ties
    cand, probs = filtered_distribution(row, top_k, top_p)
    return draw(cand, probs, u)


def find_last_equal(t: torch.Tensor, c: int) -> torch.Tensor:
    """Last column index equal to c per row, -1 if absent (reference :148-165)."""
    arr = t.cpu().numpy()
    out = np.full(arr.shape[0], -1, dtype=np.int64)
    for b in range(arr.shape[0]):
        hits = np.nonzero(arr[b] == c)[0]
        if hits.size:
            out[b] = hits[-1]
    return torch.from_numpy(out)


# ---- counter-based uniforms shared with the CUDA path --------------------------------

_M32 = 0xFFFFFFFF


def _mix32(x: int) -> int:
    x &= _M32
    x ^= x >> 16
    x = (x * 0x7FEB352D) & _M32
    x ^= x >> 15
    x = (x * 0x846CA68B) & _M32
    x ^= x >> 16
    return x


def uniform_u(seed: int, step: int, row: int, channel: int) -> float:
    """Stateless uniform in [0,1): the same integer hash the CUDA sampler uses
    (moss_tts_b200/csrc/sampling.cu `mt_uniform`)."""
    h = _mix32(seed ^ 0x9E3779B9)
    h = _mix32(h ^ ((step * 0x85EBCA6B) & _M32))
    h = _mix32(h ^ ((row * 0xC2B2AE35) & _M32))
    h = _mix32(h ^ ((channel * 0x27D4EB2F) & _M32))
    return float(np.float32(h >> 8) * np.float32(1.0 / 16777216.0))
