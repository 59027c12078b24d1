"""Sampler / state-machine helpers vs the oracle, through the C ABI."""
import numpy as np
import pytest
import torch

from moss_tts_b200.inference_utils import sample_token, find_last_equal_C
from moss_tts_b200.processing_moss_tts import MossTTSDelayProcessor
from oracle import sampling as S
from oracle.codec import apply_delay_pattern, apply_de_delay_pattern
from gpu_util import DEV

pytestmark = pytest.mark.gpu


def _rows(n, V, seed, scale=2.0):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(n, V, generator=g) * scale).to(torch.bfloat16)
    x[:, -1] = float("-inf")
    return x


@pytest.mark.parametrize("V", [1025, 151936])
def test_greedy_first_max_index(V):
    x = _rows(8, V, 1)
    x[0, 5] = x[0, 900] = 50.0          # exact tie -> lowest index
    x[1, :] = -3.0                      # everything tied
    got = sample_token(x.to(DEV), do_sample=False).cpu()
    for r in range(x.shape[0]):
        assert int(got[r]) == S.sample_row(x[r], do_sample=False)


def test_greedy_with_penalty_matches_oracle():
    x = _rows(6, 1025, 2)
    prev = torch.randint(0, 1025, (3, 40, 4), generator=torch.Generator().manual_seed(3))
    got = sample_token(x.to(DEV), prev_tokens=prev.to(DEV), repetition_penalty=1.3, do_sample=False).cpu()
    ids = S.penalty_set(prev)
    for r in range(x.shape[0]):
        assert int(got[r]) == S.sample_row(x[r], prev_ids=ids, repetition_penalty=1.3, do_sample=False)


@pytest.mark.parametrize("V,k,p", [(1025, 25, 0.8), (1025, 50, 0.9), (1025, 5, 0.5), (1025, 25, 1.0),
                                   (151936, 50, 1.0), (151936, 50, 0.9), (4000, 1024, 0.95)])
def test_sampling_matches_oracle_draws(V, k, p):
    n = 48 if V < 10000 else 6
    x = _rows(n, V, 5 + k)
    g = torch.Generator().manual_seed(9)
    match = 0
    total = 0
    for rep in range(4 if V < 10000 else 2):
        u = torch.rand(n, generator=g)
        got = sample_token(x.to(DEV), top_p=p, top_k=k, do_sample=True, uniforms=u.to(DEV)).cpu()
        for r in range(n):
            want = S.sample_row(x[r], top_p=p, top_k=k, do_sample=True, u=float(u[r]))
            match += int(got[r]) == want
            total += 1
            cand, _ = S.filtered_distribution(x[r], k, p)
            assert int(got[r]) in set(cand.tolist())        # never outside the filtered support
    assert match / total >= 0.97, (match, total)


@pytest.mark.parametrize("V,k,p", [(151936, None, 0.9), (151936, None, 1.0), (151936, 0, 0.5), (151936, 5000, 0.95),
                                   (151936, 1025, 1.0), (4000, 2000, 0.8), (2050, None, 0.7)])
def test_full_vocabulary_sampling_matches_oracle(V, k, p):
    """a5: sample_token with top_k None / <= 0 (no filter, reference :135) or top_k > 1024: the counting-sort sampler
    (csrc/sampling_full.cuh) against the oracle's full sort with the same uniforms.  u close to 1 draws the LAST kept
    candidate, which pins the top-k / top-p cut itself; a draw may differ from the oracle's only by landing on a
    neighbouring candidate of the same sorted order (bucket sums vs element-by-element fp32 sums)."""
    n = 6
    x = _rows(n, V, 77, scale=2.0)
    g = torch.Generator().manual_seed(13)
    match = total = 0
    for rep in range(3):
        u = torch.rand(n, generator=g)
        if rep == 2:
            u = torch.tensor([0.0, 0.99999, 0.9999, 0.5, 1e-7, 0.999])
        got = sample_token(x.to(DEV), top_p=p, top_k=k, do_sample=True, uniforms=u.to(DEV)).cpu()
        for r in range(n):
            cand, probs = S.filtered_distribution(x[r], k, p)
            want = S.draw(cand, probs, float(u[r]))
            total += 1
            if int(got[r]) == want:
                match += 1
                continue
            # the sorted order continues past the oracle's cut: the probabilities are bf16-rounded quotients e / Z, so a
            # last-bit difference in Z moves the cut (and with it the last kept candidate) by a position or two
            order = S.topk_survivors(x[r], V)
            pos = {int(c): j for j, c in enumerate(order.tolist())}
            pg, pw = pos[int(got[r])], pos[want]
            cdf = np.cumsum(np.asarray(probs, dtype=np.float64)) / float(np.sum(probs, dtype=np.float64))
            close = abs(pg - pw) <= 2 or (pg < len(cand) and abs(cdf[pg] - cdf[pw]) <= 2e-6)
            assert pg <= len(cand) + 1 and close, (r, float(u[r]), pg, pw, len(cand))
    print(f"[full-vocabulary sampling V={V} k={k} p={p}] {match}/{total} draws identical to the oracle")
    assert match / total >= 0.8, (match, total)


def test_full_vocabulary_sampling_with_penalty_and_ties():
    """Repetition penalty (non-monotone transform) and massive ties (a quantised row) through the counting sort."""
    V = 30000
    g = torch.Generator().manual_seed(3)
    x = (torch.randint(-8, 9, (4, V), generator=g).float() * 0.25).to(torch.bfloat16)       # 33 distinct values
    prev = torch.randint(0, V, (4, 50), generator=g)
    u = torch.tensor([0.3, 0.77, 0.001, 0.9999])
    ids = S.penalty_set(prev)
    for k, p in ((None, 0.6), (3000, 1.0), (None, 1.0)):
        got = sample_token(x.to(DEV), prev_tokens=prev.to(DEV), repetition_penalty=1.5, top_k=k, top_p=p, do_sample=True,
                           uniforms=u.to(DEV)).cpu()
        for r in range(4):
            row = S.apply_penalty_row(x[r], ids, 1.5)
            cand, probs = S.filtered_distribution(row, k, p)
            want = S.draw(cand, probs, float(u[r]))
            pos = {int(c): j for j, c in enumerate(cand.tolist())}
            assert int(got[r]) in pos
            assert abs(pos[int(got[r])] - pos[want]) <= 2, (k, p, r, pos[int(got[r])], pos[want], len(cand))


def test_sampling_distribution():
    """Empirical frequencies of one row follow the filtered distribution."""
    x = _rows(1, 1025, 21, scale=1.5)
    cand, probs = S.filtered_distribution(x[0], 25, 0.8)
    probs = probs / probs.sum()
    N = 20000
    xs = x.to(DEV).expand(N, -1)
    got = sample_token(xs, top_p=0.8, top_k=25, do_sample=True,
                       uniforms=torch.rand(N, generator=torch.Generator().manual_seed(4)).to(DEV)).cpu().numpy()
    assert set(np.unique(got)).issubset(set(cand.tolist()))
    freq = np.array([(got == c).mean() for c in cand])
    assert np.abs(freq - probs).max() < 0.015


def test_find_last_equal():
    g = torch.Generator().manual_seed(1)
    t = torch.randint(0, 6, (9, 77), generator=g)
    for c in (3, 99):
        got = find_last_equal_C(t.to(DEV), c).cpu()
        np.testing.assert_array_equal(got.numpy(), S.find_last_equal(t, c).numpy())


@pytest.mark.parametrize("T,n", [(1, 1), (21, 8), (125, 32), (3, 16)])
def test_delay_pattern_roundtrip(T, n):
    g = torch.Generator().manual_seed(T)
    codes = torch.randint(0, 1024, (T, n), generator=g)
    d = MossTTSDelayProcessor.apply_delay_pattern(codes.to(DEV), 1024)
    np.testing.assert_array_equal(d.cpu().numpy(), apply_delay_pattern(codes.numpy(), 1024))
    back = MossTTSDelayProcessor.apply_de_delay_pattern(d)
    np.testing.assert_array_equal(back.cpu().numpy(), codes.numpy())
    np.testing.assert_array_equal(back.cpu().numpy(), apply_de_delay_pattern(d.cpu().numpy()))


# ---- one-pass selection (sampling_select.cuh): edge cases of the candidate list and its exact fallbacks ----------
def _native_sample(x, *, temperature=1.0, top_k=0, top_p=1.0, do_sample=True, u=None, bitmap=None, penalty=1.0):
    import ctypes as C
    from moss_tts_b200 import _lib as L
    x = x.to(DEV).contiguous()
    N, V = x.shape
    out = torch.empty(N, dtype=torch.int64, device=DEV)
    uu = None if u is None else u.to(device=DEV, dtype=torch.float32).contiguous()
    bm = None if bitmap is None else bitmap.to(DEV).contiguous()
    L.check(L.lib().mt_sample_tokens(L.ptr(x), C.c_longlong(x.stride(0)), N, V, C.c_float(temperature), L.ptr(bm),
                                     C.c_float(penalty), C.c_float(top_p), int(top_k), int(do_sample), L.ptr(uu),
                                     L.ptr(out), L.stream_ptr(DEV)), "mt_sample_tokens")
    return out.cpu()


def _quantised_rows(n, V, seed, levels=23):
    """Few distinct values -> many exact ties, so the top-k boundary always cuts through a tie class."""
    g = torch.Generator().manual_seed(seed)
    x = (torch.randint(0, levels, (n, V), generator=g).float() * 0.125 - 1.0).to(torch.bfloat16)
    x[:, -1] = float("-inf")
    return x


@pytest.mark.parametrize("V,k", [(1025, 25), (1025, 58), (1025, 59), (151936, 50), (151936, 300), (151936, 600),
                                 (70, 25), (10, 25), (4097, 25)])
def test_topk_support_with_ties_is_lowest_index_first(V, k):
    """u -> 0 draws the best candidate and u -> 1 the last one kept: with top_p = 1 those are the first and the k-th
    entry of the (value desc, index asc) order, which pins the tie handling of every selection path (warp / block,
    one-pass / exact fallback: 23 levels over a 1025-entry row overflow neither list, over 151936 entries both)."""
    n = 16 if V < 10000 else 4
    x = _quantised_rows(n, V, V + k)
    x[0, :] = 0.5                                                     # a row of nothing but ties
    for u_val in (0.0, 0.999999):
        u = torch.full((n,), u_val)
        got = _native_sample(x, top_k=k, top_p=1.0, u=u)
        for r in range(n):
            assert int(got[r]) == S.sample_row(x[r], top_k=k, top_p=1.0, do_sample=True, u=u_val), (r, u_val)


@pytest.mark.parametrize("V,k,temp", [(1025, 25, 1.7), (1025, 50, 0.6), (151936, 50, 1.5), (151936, 50, 3.3)])
def test_temperature_merges_neighbouring_values(V, k, temp):
    """bf16(l / T) maps neighbouring logits onto one value: the candidate set must be chosen on the transformed
    values (ties -> lowest index), not on the raw ones."""
    n = 12 if V < 10000 else 3
    g = torch.Generator().manual_seed(V + k)
    x = (torch.randn(n, V, generator=g) * 0.05 + 4.0).to(torch.bfloat16)      # values packed into a few bf16 steps
    x[:, -1] = float("-inf")
    xt = (x.float() / temp).to(torch.bfloat16)
    for u_val in (0.0, 0.37, 0.999999):
        u = torch.full((n,), u_val)
        got = _native_sample(x, temperature=temp, top_k=k, top_p=1.0, u=u)
        for r in range(n):
            cand, _ = S.filtered_distribution(xt[r], k, 1.0)
            assert int(got[r]) in set(cand.tolist()), (r, u_val)
            if u_val != 0.37:
                assert int(got[r]) == S.sample_row(xt[r], top_k=k, top_p=1.0, do_sample=True, u=u_val), (r, u_val)
    got = _native_sample(x, temperature=temp, do_sample=False)
    for r in range(n):
        assert int(got[r]) == S.sample_row(xt[r], do_sample=False)


@pytest.mark.parametrize("V,k", [(1025, 25), (151936, 50)])
def test_sampling_with_penalty_matches_oracle(V, k):
    n = 24 if V < 10000 else 4
    x = _rows(n, V, 77)
    prev = torch.randint(0, min(V, 1025), (2, 60, 3), generator=torch.Generator().manual_seed(5))
    ids = S.penalty_set(prev)
    u = torch.rand(n, generator=torch.Generator().manual_seed(6))
    got = sample_token(x.to(DEV), prev_tokens=prev.to(DEV), repetition_penalty=1.2, top_k=k, top_p=0.9, do_sample=True,
                       uniforms=u.to(DEV)).cpu()
    match = 0
    for r in range(n):
        match += int(got[r]) == S.sample_row(x[r], prev_ids=ids, repetition_penalty=1.2, top_k=k, top_p=0.9,
                                             do_sample=True, u=float(u[r]))
    assert match >= n - 1, (match, n)


def test_rows_with_nan_and_inf_do_not_hang():
    x = _rows(8, 1025, 3)
    x[0, 7] = float("nan")
    x[1, :] = float("-inf")
    x[2, 100] = float("inf")
    got = _native_sample(x, top_k=25, top_p=0.8, u=torch.full((8,), 0.5))
    assert int(got[1]) == 0                              # nothing survives -> 0 (torch.argmax of an all -inf row)
    g2 = _native_sample(x, do_sample=False)
    assert int(g2[1]) == 0 and int(g2[2]) == 100
    for r in range(3, 8):
        assert int(g2[r]) == S.sample_row(x[r], do_sample=False)


def test_penalty_per_head_on_3d_logits_matches_reference():
    """sample_token on [B, H, V] logits: per-head id sets (reference inference_utils.py:88-106), checked against the
    reference-executed golden; greedy and sampled (same uniforms -> the oracle's draw on the per-head penalised rows)."""
    import os
    from conftest import GOLDEN
    g = np.load(os.path.join(GOLDEN, "sampling3d.npz"), allow_pickle=False)
    x = torch.from_numpy(g["logits"]).to(torch.bfloat16)
    prev = torch.from_numpy(g["prev"])
    got = sample_token(x.to(DEV), prev_tokens=prev.to(DEV), repetition_penalty=1.3, do_sample=False).cpu()
    np.testing.assert_array_equal(got.numpy(), g["argmax"])
    assert tuple(got.shape) == (3, 4) and (got.numpy().reshape(-1) != g["argmax_union"]).any()
    u = torch.rand(3, 4, generator=torch.Generator().manual_seed(1))
    got = sample_token(x.to(DEV), prev_tokens=prev.to(DEV), repetition_penalty=1.3, top_k=25, top_p=0.8, do_sample=True,
                       uniforms=u.to(DEV)).cpu()
    pen = S.apply_penalty_heads(x, prev, 1.3)
    want = [[S.sample_row(pen[b, h], top_k=25, top_p=0.8, do_sample=True, u=float(u[b, h])) for h in range(4)] for b in range(3)]
    assert (got.numpy() == np.asarray(want)).mean() >= 0.9
    # top_k=None on a row of at most 1024 entries = no filter
    y = x[:, :, :1024].contiguous()
    got = sample_token(y.to(DEV), top_k=None, top_p=0.9, do_sample=True, uniforms=u.to(DEV)).cpu()
    want = [[S.sample_row(y[b, h], top_k=None, top_p=0.9, do_sample=True, u=float(u[b, h])) for h in range(4)] for b in range(3)]
    assert (got.numpy() == np.asarray(want)).mean() >= 0.9
