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
