"""Pins the CPU oracle against fixtures minted by executing the reference
(oracle/make_golden.py).  CPU only."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from moss_tts_b200.config import LMConfig, CodecConfig, CodecStage
from oracle import sampling as S
from oracle.codec import (OracleCodec, apply_delay_pattern, apply_de_delay_pattern,
                          split_audio_segments, trim_first_segment)
from oracle.lm import OracleLM
from oracle.weights import make_lm_weights, make_codec_weights


def _load(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)


def _lmcfg(g):
    return LMConfig(**json.loads(str(g["cfg"])))


def _codec_cfg(js):
    d = json.loads(js)
    stages = [CodecStage(**s) for s in d.pop("stages")]
    return CodecConfig(stages=stages, **d)


@pytest.mark.parametrize("name,logit_tol", [
    ("lm_tiny_f32", 2e-5), ("lm_tiny_bf16", 2e-2), ("lm_tiny_pen_f32", 2e-5),
    ("lm_mid_f32", 5e-5), ("lm_mid_bf16", 2e-2)])
def test_lm_generate_and_logits(name, logit_tol):
    g = _load(name)
    cfg = _lmcfg(g)
    dtype = torch.bfloat16 if "bfloat16" in str(g["dtype"]) else torch.float32
    w = make_lm_weights(cfg, seed=int(g["seed"]), never_stop=bool(g["never_stop"]))
    lm = OracleLM(cfg, w, dtype)
    ids = torch.from_numpy(g["prompt_ids"])
    mask = torch.from_numpy(g["prompt_mask"])
    out = lm.generate(ids, mask, max_new_tokens=int(g["steps"]), text_temperature=0, audio_temperature=0,
                      audio_repetition_penalty=float(g["penalty"]))
    for b, (sl, gen) in enumerate(out):
        assert sl == int(g[f"start_length{b}"])
        np.testing.assert_array_equal(gen.numpy(), g[f"gen{b}"])       # greedy ids bit-exact
    # teacher-forced logits
    logits = lm.forward(torch.from_numpy(g["tf_ids"]), torch.from_numpy(g["tf_mask"]))
    pos = g["tf_positions"].tolist()
    for i in range(1, cfg.n_vq + 1):
        mine = logits[i][:, pos, :].float().numpy()
        ref = g[f"tf_audio{i}"]
        assert np.array_equal(np.isinf(mine), np.isinf(ref))
        fin = np.isfinite(ref)
        assert np.abs(mine[fin] - ref[fin]).max() <= logit_tol
    text = logits[0][:, pos, :].float()
    got = torch.gather(text, -1, torch.from_numpy(g["tf_text_top_i"])).numpy()
    assert np.abs(got - g["tf_text_top_v"]).max() <= logit_tol


def test_sampling_filters_match_reference():
    g = _load("sampling")
    for tag, dtype in (("f32", torch.float32), ("bf16", torch.bfloat16), ("f32big", torch.float32)):
        x = torch.from_numpy(g[f"{tag}_logits"]).to(dtype)
        for k, p in ((25, 0.8), (50, 0.9), (5, 0.5), (25, 1.0)):
            ref = g[f"{tag}_k{k}_p{p}_probs"]
            for r in range(x.shape[0]):
                cand, probs = S.filtered_distribution(x[r], k, p)
                dense = np.zeros(x.shape[1], dtype=np.float32)
                dense[cand] = probs
                if dtype == torch.float32:
                    np.testing.assert_allclose(dense, ref[r], rtol=1e-5, atol=1e-7)
                else:
                    # bf16 rows have ties at the cut; the support may differ by tied entries only
                    diff = np.nonzero((dense > 0) != (ref[r] > 0))[0]
                    if diff.size:
                        vals = x[r].float().numpy()
                        assert np.unique(vals[diff]).size == 1
                    else:
                        np.testing.assert_allclose(dense, ref[r], rtol=1e-2, atol=1e-3)
        prev = torch.from_numpy(g[f"{tag}_prev"])
        ids = S.penalty_set(prev)
        for r in range(x.shape[0]):
            mine = S.apply_penalty_row(x[r], ids, 1.3).float().numpy()
            np.testing.assert_array_equal(mine, g[f"{tag}_penalized"][r])
            assert S.sample_row(x[r], do_sample=False) == int(g[f"{tag}_argmax"][r])
    t = torch.from_numpy(g["fle_in"])
    np.testing.assert_array_equal(S.find_last_equal(t, 3).numpy(), g["fle_out3"])
    np.testing.assert_array_equal(S.find_last_equal(t, 9).numpy(), g["fle_out9"])


def test_draw_is_inverse_cdf():
    cand = np.array([5, 2, 9]); probs = np.array([0.5, 0.3, 0.2], dtype=np.float32)
    assert S.draw(cand, probs, 0.0) == 5
    assert S.draw(cand, probs, 0.49) == 5
    assert S.draw(cand, probs, 0.51) == 2
    assert S.draw(cand, probs, 0.99) == 9
    us = [S.uniform_u(1, s, 0, 0) for s in range(2000)]
    assert 0.0 <= min(us) and max(us) < 1.0 and abs(np.mean(us) - 0.5) < 0.03


def test_codec_matches_reference():
    g = _load("codec_mini")
    cfg = _codec_cfg(str(g["cfg"]))
    oc = OracleCodec(cfg, make_codec_weights(cfg, seed=int(g["seed"])))
    codes = torch.from_numpy(g["codes"])
    np.testing.assert_allclose(oc.decode_codes(codes).numpy(), g["emb"], atol=2e-5, rtol=1e-5)
    np.testing.assert_allclose(oc.decode_codes(codes[:3]).numpy(), g["emb_nq3"], atol=2e-5, rtol=1e-5)
    wav, lens = oc.decode(codes)
    scale = np.abs(g["wav"]).max()
    assert np.abs(wav.numpy() - g["wav"]).max() <= 1e-4 * scale
    np.testing.assert_array_equal(lens.numpy(), g["wav_lengths"])
    pm = torch.ones(2, 40, dtype=torch.bool); pm[1, 29:] = False
    wav5, lens5 = oc.decode(codes, padding_mask=pm, num_quantizers=5)
    assert np.abs(wav5.numpy() - g["wav_nq5"]).max() <= 1e-4 * scale
    np.testing.assert_array_equal(lens5.numpy(), g["wav_nq5_lengths"])
    # chunked decode == closed-form ring mask, and it genuinely differs from unchunked
    wc, _ = oc.decode(codes[:, :1, :33], chunk_duration=0.8)
    assert np.abs(wc.numpy() - g["wav_chunked"]).max() <= 1e-4 * scale
    assert np.abs(g["wav_chunked"] - g["wav_unchunked33"]).max() > 1e-2 * scale
    with pytest.raises(ValueError):
        oc.decode(codes, chunk_duration=0.8)          # B != 1
    with pytest.raises(ValueError):
        oc.decode(codes[:, :1], chunk_duration=2.0)    # > context duration


def test_delay_pattern_and_parse():
    g = _load("codec_mini")
    raw = g["dp_raw"]
    np.testing.assert_array_equal(apply_delay_pattern(raw, 1024), g["dp_delayed"])
    np.testing.assert_array_equal(apply_de_delay_pattern(g["dp_delayed"]), g["dp_dedelayed"])
    np.testing.assert_array_equal(g["dp_dedelayed"], raw)
    cfg10 = _codec_cfg(str(g["cfg10"]))
    oc = OracleCodec(cfg10, make_codec_weights(cfg10, seed=5))
    dd = apply_de_delay_pattern(g["parse_in"])
    segs = split_audio_segments(dd, 1024)
    wavs = oc.decode_audio_codes([torch.from_numpy(s) for s in segs])
    wavs = trim_first_segment(wavs, [s.shape[0] for s in segs], 1)
    assert len(wavs) == int(g["parse_n"])
    for i, wv in enumerate(wavs):
        ref = g[f"parse_wav{i}"]
        assert wv.shape == ref.shape
        assert np.abs(wv.numpy() - ref).max() <= 1e-4 * np.abs(ref).max()


def test_streaming_pushes_closed_form_matches_reference_streaming():
    """`with model.streaming(B): _decode_frame(chunk)` over uneven pushes == one pass with the per-push ring floor."""
    g = _load("codec_stream_mini")
    from moss_tts_b200.config import CodecConfig
    cfg = CodecConfig.from_dict(json.loads(str(g["cfg"])))
    oc = OracleCodec(cfg, make_codec_weights(cfg, seed=int(g["seed"])))
    codes = torch.from_numpy(g["codes"])
    wav = oc.decode_frames(codes, chunk_frames=[int(n) for n in g["pushes"]]).numpy()
    assert np.abs(wav - g["wav_stream"]).max() <= 2e-5 * max(1.0, np.abs(g["wav_stream"]).max())
    assert np.abs(g["wav_stream"] - g["wav_full"]).max() > 0.1          # and streaming genuinely differs from one shot


# ---- BASELINE shapes (oracle/make_golden_big.py): 8B widths x 2 layers, released 68-layer codec --------------------------
@pytest.fixture(scope="module")
def wide_base_weights():
    from oracle.make_golden_big import wide_cfg, wide_weights
    return wide_weights(wide_cfg(32))


@pytest.mark.parametrize("which,n_vq", [("lm32", 32), ("lm16", 16)])
def test_oracle_matches_reference_at_8b_widths(wide_base_weights, which, n_vq):
    """H 4096 / I 12288 / 32q-8kv x 128 / V 151936: greedy ids bit-exact (n_vq 32: ragged left pad; n_vq 16: codes in the
    prompt, penalty 1.1) and the teacher-forced top-1 logit of every (row, position, head) bit-equal to the reference's."""
    from oracle.make_golden_big import wide_prompt, wide_weights
    g = np.load(os.path.join(GOLDEN, f"lm_wide_nvq{n_vq}_bf16.npz"), allow_pickle=False)
    cfg = LMConfig(**json.loads(str(g["cfg"])))
    oracle = OracleLM(cfg, wide_weights(cfg, wide_base_weights), torch.bfloat16)
    ids, mask = wide_prompt(cfg, which)
    out = oracle.generate(ids, mask, max_new_tokens=int(g["steps"]), text_temperature=0, audio_temperature=0,
                          audio_repetition_penalty=float(g["penalty"]))
    for b, (sl, gen) in enumerate(out):
        assert sl == int(g[f"start_length{b}"])
        np.testing.assert_array_equal(gen.numpy(), g[f"gen{b}"].astype(np.int64))
    logits = oracle.forward(torch.from_numpy(g["tf_ids"].astype(np.int64)), torch.from_numpy(g["tf_mask"]))
    valid = g["tf_mask"].astype(bool)
    for h, lg in enumerate(logits):
        lg = lg.float().numpy()
        for k in range(2):
            idx = g["tf_top2_i"][:, :, h, k].astype(np.int64)
            got = np.take_along_axis(lg, idx[..., None], -1)[..., 0]
            assert np.array_equal(got[valid], g["tf_top2_v"][:, :, h, k][valid]), (which, h, k)


def test_oracle_matches_reference_released_codec_68_layers():
    from oracle.make_golden_big import CODEC_SEED
    g = np.load(os.path.join(GOLDEN, "codec_full.npz"), allow_pickle=False)
    cfg = CodecConfig.from_dict(json.loads(str(g["cfg"])))
    assert sum(st.num_layers for st in cfg.stages) == 68
    oc = OracleCodec(cfg, make_codec_weights(cfg, seed=CODEC_SEED, layer_scale=float(g["layer_scale"])))
    codes = torch.from_numpy(g["codes"].astype(np.int64))
    for key, kw in (("wav", {}), ("wav_chunked", {"chunk_duration": 8})):
        wav, _ = oc.decode(codes, **kw)
        ref = g[key]
        assert wav.shape == ref.shape
        assert float(np.abs(wav.numpy() - ref).max()) <= 2e-4 * float(np.abs(ref).max())
    assert float(np.abs(g["wav"] - g["wav_chunked"]).max()) > 1e-2          # the chunked path genuinely differs


def test_penalty_per_head_on_3d_logits_matches_reference():
    """[B, H, V] logits: head h is penalised with the ids of prev_tokens[..., h] only (reference "case 2")."""
    g = np.load(os.path.join(GOLDEN, "sampling3d.npz"), allow_pickle=False)
    x = torch.from_numpy(g["logits"]).to(torch.bfloat16)
    prev = torch.from_numpy(g["prev"])
    pen = S.apply_penalty_heads(x, prev, 1.3)
    np.testing.assert_array_equal(pen.float().numpy(), g["penalized"])
    np.testing.assert_array_equal(pen.float().argmax(-1).numpy(), g["argmax"])
    assert (g["argmax"].reshape(-1) != g["argmax_union"]).any()          # the fixture separates the two cases
