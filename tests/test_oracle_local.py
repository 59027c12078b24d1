"""MossTTSLocal oracle (oracle/local.py) vs goldens minted by the reference's own modules
(oracle/make_golden_local.py -> tests/golden/local_*.npz).  CPU only."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from moss_tts_b200.config import LocalConfig
from oracle.local import OracleLocal
from oracle.weights import make_local_weights


def bf16_ulp(x: float) -> float:
    x = abs(float(x))
    return 2.0 ** (np.floor(np.log2(x)) - 7) if x > 0 else 2.0 ** -133


def load_local_golden(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    meta = json.loads(str(g["meta"]))
    cfg = LocalConfig(**meta["cfg"])
    return g, meta, cfg


def first_divergence(mine: np.ndarray, ref: np.ndarray, S0: int):
    """(frame, channel) of the first differing cell in generation order, or None."""
    bad = np.argwhere(mine != ref)
    if len(bad) == 0:
        return None
    r, c = sorted(map(tuple, bad))[0]
    return int(r) - S0, int(c)


@pytest.mark.parametrize("name", ["local_tiny_f32", "local_mid_f32"])
def test_local_oracle_fp32_is_exact(name):
    g, meta, cfg = load_local_golden(name)
    w = make_local_weights(cfg, seed=meta["seed"])
    ids = torch.from_numpy(g["input_ids"])
    out, trace = OracleLocal(cfg, w, torch.float32).generate(ids, meta["steps"], eos_token_id=cfg.audio_end_token_id,
                                                             return_trace=True)
    S0 = ids.shape[1]
    for b, (sl, rows) in enumerate(out):
        start = S0 - sl - 1
        assert np.array_equal(rows.numpy(), g["gen_ids"][b, start:]), (name, b)
    # the largest logit of every (frame, channel, row) agrees with the reference's
    tv = g["top2_val"]
    for f, step in enumerate(trace):
        for c, lg in enumerate(step):
            np.testing.assert_allclose(lg.float().max(dim=-1).values.numpy(), tv[f, c, :, 0], rtol=0, atol=2e-4)


@pytest.mark.parametrize("name", ["local_tiny_bf16", "local_mid_bf16"])
def test_local_oracle_bf16_exact_or_near_tie(name):
    """bf16: the reference runs the local transformer under sdpa, the oracle restates eager attention; ids must
    agree up to the first bf16 near-tie (reference margin <= 2 ulp), after which the sequences legitimately fork."""
    g, meta, cfg = load_local_golden(name)
    w = make_local_weights(cfg, seed=meta["seed"])
    ids = torch.from_numpy(g["input_ids"])
    out = OracleLocal(cfg, w, torch.bfloat16).generate(ids, meta["steps"], eos_token_id=cfg.audio_end_token_id)
    S0 = ids.shape[1]
    tv = g["top2_val"]
    for b, (sl, rows) in enumerate(out):
        start = S0 - sl - 1
        mine, ref = rows.numpy(), g["gen_ids"][b, start:]
        assert mine.shape == ref.shape
        d = first_divergence(mine[sl + 1:], ref[sl + 1:], 0)
        if d is None:
            continue
        f, c = d
        margin = float(tv[f, c, b, 0] - tv[f, c, b, 1])
        ulps = margin / bf16_ulp(tv[f, c, b, 0])
        print(f"[{name}] row {b}: first divergence at frame {f} channel {c}, reference margin {ulps:.1f} bf16 ulp")
        assert ulps <= 2.0, (name, b, f, c, ulps)
        assert f >= 1 or c >= 1, "the very first token must agree"


def test_local_eos_and_partial_channels():
    """EOS substitution (:430-433) and n_vq_for_inference (:397, :427) on the oracle itself."""
    cfg = LocalConfig.tiny(4)
    w = make_local_weights(cfg, seed=11)
    ids = torch.from_numpy(np.load(os.path.join(GOLDEN, "local_tiny_f32.npz"))["input_ids"])
    o = OracleLocal(cfg, w, torch.float32)
    base = o.generate(ids, 3, eos_token_id=None)
    tok0 = int(base[0][1][-3, 0])                   # first generated text token of row 0
    out = o.generate(ids, 6, eos_token_id=tok0)      # row 0 finishes at once
    r0 = out[0][1][-out[0][1].shape[0] + (out[0][0] + 1):]
    assert int(r0[0, 0]) == tok0
    if r0.shape[0] > 1:
        assert (r0[1:, 0] == tok0).all() and (r0[1:, 1:] == cfg.audio_pad_code).all()
    part = o.generate(ids, 2, n_vq_for_inference=2, eos_token_id=None)
    assert (part[0][1][-2:, 3:] == 0).all()


def test_local_oracle_left_padded_prompts():
    """Left-padded batch (rows of 12, 9 and 7 real tokens): the reference modules were fed the mask extended with ones and
    HF's cumulative-sum position_ids; the oracle masks the padding and keeps column positions.  fp32: ids exact, top
    logits <= 2e-4 (the per-row position shift is invisible to rotary attention); bf16: exact or near-tie."""
    g, meta, cfg = load_local_golden("local_tiny_pad_f32")
    w = make_local_weights(cfg, seed=meta["seed"])
    ids, mask = torch.from_numpy(g["input_ids"]), torch.from_numpy(g["attention_mask"])
    assert not mask.all() and mask[:, -1].all()
    out, trace = OracleLocal(cfg, w, torch.float32).generate(ids, meta["steps"], eos_token_id=cfg.audio_end_token_id,
                                                             return_trace=True, attention_mask=mask)
    S0 = ids.shape[1]
    for b, (sl, rows) in enumerate(out):
        assert np.array_equal(rows.numpy(), g["gen_ids"][b, S0 - sl - 1:]), b
    for f, step in enumerate(trace):
        for c, lg in enumerate(step):
            np.testing.assert_allclose(lg.float().max(dim=-1).values.numpy(), g["top2_val"][f, c, :, 0], rtol=0, atol=2e-4)
    # and the padding matters: ignoring the mask changes the logits
    out2, trace2 = OracleLocal(cfg, w, torch.float32).generate(ids, 1, eos_token_id=cfg.audio_end_token_id, return_trace=True)
    assert (trace2[0][0][1].float() - trace[0][0][1].float()).abs().max() > 1e-3
    gb, metab, _ = load_local_golden("local_tiny_pad_bf16")
    outb = OracleLocal(cfg, w, torch.bfloat16).generate(ids, metab["steps"], eos_token_id=cfg.audio_end_token_id,
                                                        attention_mask=mask)
    for b, (sl, rows) in enumerate(outb):
        d = first_divergence(rows.numpy(), gb["gen_ids"][b, S0 - sl - 1:], sl + 1)
        if d is not None:
            f, c = d
            v = gb["top2_val"][f, c, b]
            assert (v[0] - v[1]) <= 2 * bf16_ulp(v[0]), (b, f, c, v)
