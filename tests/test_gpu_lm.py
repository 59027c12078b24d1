"""MossTTSDelay CUDA path vs the oracle (and the reference-minted goldens), through the C ABI."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from moss_tts_b200.config import LMConfig
from moss_tts_b200.modeling_moss_tts import MossTTSDelayModel
from oracle.lm import OracleLM
from oracle import sampling as S
from oracle.weights import make_lm_weights, make_prompt
from gpu_util import DEV, bf16_ulp

pytestmark = pytest.mark.gpu

LOGIT_TOL = 2e-2      # north_star: teacher-forced logits max-abs <= 2e-2 (bf16)


def _golden(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    return g, LMConfig(**json.loads(str(g["cfg"])))


def _check_tokens(out, ref_rows, oracle_trace, start_lengths, what):
    """Greedy ids must be bit-exact; if they are not, the first divergence must be a bf16 near-tie
    (oracle margin <= 2 ulp), which we report but tolerate."""
    exact = True
    for b, ((sl, mine), ref) in enumerate(zip(out, ref_rows)):
        mine = mine.cpu().numpy()
        assert mine.shape == ref.shape, (what, b, mine.shape, ref.shape)
        bad = np.argwhere(mine != ref)
        if len(bad) == 0:
            continue
        exact = False
        r, c = bad[0]
        step = int(r) - int(start_lengths[b])
        lg = oracle_trace[step][c][b].float()
        top = lg.max().item()
        margin = top - lg[int(mine[r, c])].item()
        ulps = margin / bf16_ulp(top)
        print(f"[{what}] row {b}: first divergence at gen row {r} ch {c}: mine {mine[r, c]} ref {ref[r, c]} "
              f"oracle margin {margin:.4g} = {ulps:.1f} bf16 ulp; {len(bad)} cells differ downstream")
        assert ulps <= 2.0, f"{what}: divergence is not a near-tie ({ulps:.1f} ulp)"
    return exact


@pytest.mark.parametrize("name", ["lm_tiny_bf16", "lm_mid_bf16"])
def test_generate_greedy_matches_reference_golden(name):
    g, cfg = _golden(name)
    w = make_lm_weights(cfg, seed=int(g["seed"]), never_stop=bool(g["never_stop"]))
    model = MossTTSDelayModel(cfg, w, device=DEV)
    ids = torch.from_numpy(g["prompt_ids"])
    mask = torch.from_numpy(g["prompt_mask"])
    B = ids.shape[0]
    kw = dict(max_new_tokens=int(g["steps"]), text_temperature=0, audio_temperature=0)
    out = model.generate(ids, mask, **kw)
    oracle = OracleLM(cfg, w, torch.bfloat16)
    o_out, trace = oracle.generate(ids, mask, return_trace=True, **kw)
    for b in range(B):
        np.testing.assert_array_equal(o_out[b][1].numpy(), g[f"gen{b}"])      # oracle == reference
        assert out[b][0] == int(g[f"start_length{b}"])
    exact = _check_tokens(out, [g[f"gen{b}"] for b in range(B)], trace, [o[0] for o in out], name)
    print(f"[{name}] greedy ids bit-exact vs reference golden: {exact}")
    # CUDA graph on/off must not change anything
    out2 = model.generate(ids, mask, use_cuda_graph=False, **kw)
    for (a, x), (b_, y) in zip(out, out2):
        assert a == b_ and torch.equal(x, y)


@pytest.mark.parametrize("name", ["lm_tiny_bf16", "lm_mid_bf16"])
def test_teacher_forced_logits(name):
    g, cfg = _golden(name)
    w = make_lm_weights(cfg, seed=int(g["seed"]), never_stop=bool(g["never_stop"]))
    model = MossTTSDelayModel(cfg, w, device=DEV)
    ids = torch.from_numpy(g["tf_ids"])
    mask = torch.from_numpy(g["tf_mask"])
    logits = model.forward(ids, mask).logits
    oracle = OracleLM(cfg, w, torch.bfloat16).forward(ids, mask)
    valid = mask.bool()
    worst, ndiff, ntot, big = 0.0, 0, 0, 0.0
    for i, (mine, ref) in enumerate(zip(logits, oracle)):
        mine = mine.float().cpu()[valid]
        ref = ref.float()[valid]
        assert torch.equal(torch.isinf(mine), torch.isinf(ref)), f"head {i}: -inf pattern differs"
        fin = torch.isfinite(ref)
        d = (mine[fin] - ref[fin]).abs()
        worst = max(worst, d.max().item())
        ndiff += int((d > 0).sum()); ntot += d.numel()
        big = max(big, ref[fin].abs().max().item())
    print(f"[{name}] teacher-forced logits max-abs vs oracle: {worst:.4g}; {ndiff}/{ntot} logits differ "
          f"({100.0 * ndiff / ntot:.3f}%), max |logit| {big:.3g}")
    assert worst <= LOGIT_TOL
    # and against the reference-minted digest
    pos = g["tf_positions"].tolist()
    for i in range(1, cfg.n_vq + 1):
        mine = logits[i][:, pos, :].float().cpu().numpy()
        ref = g[f"tf_audio{i}"]
        fin = np.isfinite(ref)
        assert np.abs(mine[fin] - ref[fin]).max() <= LOGIT_TOL


@pytest.mark.parametrize("name", ["lm_tiny_bf16", "lm_mid_bf16"])
def test_teacher_forced_argmax_along_reference_trajectory(name):
    """Every step of the reference's golden trajectory, every channel (no "first divergence, then stop"): the CUDA path,
    teacher-forced on the reference's ids, must pick the token the reference picked at every (row, step, channel); a
    different argmax is accepted only where the reference arithmetic's own top-2 margin is <= 2 bf16 ulp (the oracle,
    bit-exact to the reference on these goldens, supplies the full logit rows).  Counts are printed."""
    g, cfg = _golden(name)
    w = make_lm_weights(cfg, seed=int(g["seed"]), never_stop=bool(g["never_stop"]))
    model = MossTTSDelayModel(cfg, w, device=DEV)
    ids = torch.from_numpy(g["tf_ids"])
    mask = torch.from_numpy(g["tf_mask"])
    B, T, C = ids.shape
    S0 = int(g["prompt_len"])
    mine = [x.float().cpu() for x in model.forward(ids, mask).logits]
    ref = [x.float() for x in OracleLM(cfg, w, torch.bfloat16).forward(ids, mask)]
    valid = mask.bool()
    cells = flips = flips2 = ties = 0
    for h in range(C):
        am, ar = mine[h].argmax(-1), ref[h].argmax(-1)
        top2 = torch.topk(ref[h], 2, dim=-1).values
        cells += int(valid.sum())
        for b, t in np.argwhere(((am != ar) & valid).numpy()):
            margin = float(top2[b, t, 0] - top2[b, t, 1])
            if margin == 0.0:
                ties += 1
                continue
            # every logit is within one bf16 ulp of the oracle (test_teacher_forced_logits), so a winner can lose one ulp
            # while the runner-up gains one: margins up to 2 ulp can flip, anything larger is a real mismatch
            ulps = margin / bf16_ulp(float(top2[b, t, 0]))
            assert ulps <= 2.0, f"{name}: head {h} row {b} pos {t}: argmax differs at a {ulps:.1f}-ulp margin"
            flips += 1
            flips2 += ulps > 1.0
    # the tokens the reference actually generated (audio channels, rows where the channel was live)
    gen_cells = gen_flips = 0
    for b in range(B):
        rows = torch.from_numpy(g[f"gen{b}"])[int(g[f"start_length{b}"]):]        # generated rows only
        for r in range(rows.shape[0]):
            t = S0 - 1 + r                                                         # position whose logits chose this row
            for c in range(1, C):
                tok = int(rows[r, c])
                if tok == cfg.audio_pad_code:
                    continue
                gen_cells += 1
                row = mine[c][b, t]
                if int(row.argmax()) != tok:
                    ulps = float(row.max() - row[tok]) / bf16_ulp(float(row.max()))
                    assert ulps <= 2.0, f"{name}: generated token of row {b} step {r} ch {c} is {ulps:.1f} ulp below our argmax"
                    gen_flips += 1
    print(f"[{name}] teacher-forced argmax over {cells} cells (all {T - S0 + 1} steps): {flips} near-tie flips ({flips2} at a 2-ulp margin), {ties} exact ties, "
          f"0 real mismatches; {gen_cells} generated audio tokens re-scored: {gen_flips} near-tie flips")
    assert gen_cells > 0


def test_generate_left_pad_penalty_and_stop():
    """Ragged prompts (left padding), repetition penalty coupling, natural stop (not never_stop)."""
    cfg = LMConfig.tiny(n_vq=4)
    w = make_lm_weights(cfg, seed=11, never_stop=False)
    model = MossTTSDelayModel(cfg, w, device=DEV)
    oracle = OracleLM(cfg, w, torch.bfloat16)
    ids, mask = make_prompt(cfg, 3, 14, seed=11, left_pad=[0, 5, 2])
    kw = dict(max_new_tokens=24, text_temperature=0, audio_temperature=0, audio_repetition_penalty=1.2)
    out = model.generate(ids, mask, **kw)
    o_out, trace = oracle.generate(ids, mask, return_trace=True, **kw)
    _check_tokens(out, [o[1].numpy() for o in o_out], trace, [o[0] for o in o_out], "tiny-pen")


def test_generate_sampling_mode_follows_oracle():
    """do_sample path: same uniforms (counter hash) -> same ids unless a bf16 near-tie interferes."""
    cfg = LMConfig.tiny(n_vq=4)
    w = make_lm_weights(cfg, seed=5, never_stop=True)
    model = MossTTSDelayModel(cfg, w, device=DEV)
    oracle = OracleLM(cfg, w, torch.bfloat16)
    ids, mask = make_prompt(cfg, 2, 12, seed=5)
    kw = dict(max_new_tokens=12, text_temperature=1.5, text_top_k=50, text_top_p=1.0, audio_temperature=1.7,
              audio_top_k=25, audio_top_p=0.8, audio_repetition_penalty=1.1, seed=123)
    out = model.generate(ids, mask, **kw)
    o_out = oracle.generate(ids, mask, **kw)
    same = sum(int(torch.equal(a[1].cpu(), b[1])) for a, b in zip(out, o_out))
    rows = sum(int((a[1].cpu()[: b[1].shape[0]] == b[1][: a[1].shape[0]]).all(dim=1).sum()) for a, b in zip(out, o_out))
    print(f"[sampling] identical utterances {same}/{len(out)}, identical rows {rows}")
    for (sl, g_), (osl, og) in zip(out, o_out):
        assert sl == osl
        g_ = g_.cpu()
        assert g_.shape[1] == og.shape[1]
        # structure must hold regardless of draws: audio codes in range, pad where inactive on row 0
        assert int(g_[:, 1:].max()) <= 1024 and int(g_[:, 1:].min()) >= 0
        assert torch.equal(g_[0], og[0])
    assert rows >= 4


def test_generate_sampling_without_audio_top_k_follows_oracle():
    """audio_top_k=None (the reference skips apply_top_k): every real code stays a candidate, top-p still applies."""
    cfg = LMConfig.tiny(n_vq=4)
    w = make_lm_weights(cfg, seed=5, never_stop=True)
    model = MossTTSDelayModel(cfg, w, device=DEV)
    oracle = OracleLM(cfg, w, torch.bfloat16)
    ids, mask = make_prompt(cfg, 2, 12, seed=5)
    kw = dict(max_new_tokens=10, text_temperature=0, audio_temperature=1.2, audio_top_k=None, audio_top_p=0.9, seed=7)
    out = model.generate(ids, mask, **kw)
    o_out = oracle.generate(ids, mask, **kw)
    rows = sum(int((a[1].cpu()[: b[1].shape[0]] == b[1][: a[1].shape[0]]).all(dim=1).sum()) for a, b in zip(out, o_out))
    total = sum(b[1].shape[0] for b in o_out)
    print(f"[sampling, no audio top-k] identical rows {rows}/{total}")
    assert rows >= total // 2
    for (sl, g_), (osl, og) in zip(out, o_out):
        assert sl == osl and torch.equal(g_.cpu()[0], og[0])
        assert int(g_[:, 1:].max()) <= 1024 and int(g_[:, 1:].min()) >= 0


def _last_position_text_logits(model, ids, mask):
    """Text logits of the last prompt position through the call generate() itself makes (mt_lm_forward, last_only=1)."""
    import ctypes as C
    from moss_tts_b200 import _lib as L
    c = model.config
    B, S, _ = ids.shape
    model._ensure_engine(B, S, B * S)
    ids_d = ids.to(device=DEV, dtype=torch.int64).contiguous()
    mask_d = mask.to(device=DEV).to(torch.uint8).contiguous()
    text = torch.empty(B, 1, c.vocab_size, dtype=torch.bfloat16, device=DEV)
    audio = torch.empty(B, 1, c.n_vq, c.audio_head_size, dtype=torch.bfloat16, device=DEV)
    L.check(L.lib().mt_lm_forward(model._engine, L.ptr(ids_d), L.ptr(mask_d), C.c_longlong(S), B, S, 0, 1, L.ptr(text),
                                  L.ptr(audio), L.stream_ptr(torch.device(DEV))), "mt_lm_forward")
    return text[:, 0].cpu()


@pytest.mark.parametrize("text_top_k,text_top_p", [(None, 0.9), (None, 1.0), (5000, 0.95)])
def test_generate_text_sampling_without_top_k_follows_oracle(text_top_k, text_top_p):
    """text_top_k=None / > 1024 (reference inference_utils.py:135 skips apply_top_k): the text channel samples from the
    whole 151,936-entry row (full-vocabulary counting sort).  Prompts that do NOT end inside an audio span, so the text
    row is really sampled (mode 1 bans only).  With > 100,000 candidates of tiny mass each, a one-ulp difference in a
    handful of logits moves the inverse-CDF draw to another token, so the check is the sampler's own: the oracle's
    masking + sampling applied to the logits the CUDA path sampled from, with the same uniform, must give generate()'s
    first token (or its neighbour in the sorted order: bucket sums vs element-by-element sums)."""
    cfg = LMConfig.tiny(n_vq=4)
    w = make_lm_weights(cfg, seed=5)
    model = MossTTSDelayModel(cfg, w, device=DEV)
    ids, mask = make_prompt(cfg, 3, 12, seed=6)
    ids[:, -1, 0] = torch.tensor([1234, 77, 4242])           # plain text last: no audio continuation
    temp = 1.3
    kw = dict(max_new_tokens=8, text_temperature=temp, text_top_k=text_top_k, text_top_p=text_top_p,
              audio_temperature=0, seed=11)
    out = model.generate(ids, mask, **kw)
    logits = _last_position_text_logits(model, ids, mask)
    for b, (sl, g_) in enumerate(out):
        g_ = g_.cpu()
        row = (logits[b] / temp).clone()
        row[[cfg.pad_token_id, cfg.audio_assistant_gen_slot_token_id, cfg.audio_assistant_delay_slot_token_id,
             cfg.audio_end_token_id, 151662, cfg.im_end_token_id]] = float("-inf")      # step 0 bans (:459-464)
        cand, probs = S.filtered_distribution(row, text_top_k, text_top_p)
        want = S.draw(cand, probs, S.uniform_u(11, 0, b, 0))
        got = int(g_[sl, 0])
        pos = {int(c_): j for j, c_ in enumerate(cand.tolist())}
        assert got in pos, (b, "outside the filtered support")
        assert abs(pos[got] - pos[want]) <= 2, (b, pos[got], pos[want], len(cand))
        assert int(g_[:, 0].max()) < cfg.vocab_size and int(g_[:, 0].min()) >= 0


def test_get_input_embeddings_bit_exact():
    """a1: E_text[id0] + sum_i emb_ext[i][id_{i+1}], left to right, bf16 rounding after each add (reference :196-213)."""
    cfg = LMConfig.mid(n_vq=32)
    w = make_lm_weights(cfg, seed=3)
    model = MossTTSDelayModel(cfg, w, device=DEV)
    g = torch.Generator().manual_seed(2)
    ids = torch.cat([torch.randint(0, cfg.vocab_size, (3, 17, 1), generator=g),
                     torch.randint(0, cfg.audio_head_size, (3, 17, cfg.n_vq), generator=g)], dim=-1)
    got = model.get_input_embeddings(ids).cpu()
    ref = OracleLM(cfg, w, torch.bfloat16).embed(ids)
    assert got.dtype == torch.bfloat16 and tuple(got.shape) == (3, 17, cfg.hidden_size)
    assert torch.equal(got, ref)
    with pytest.raises(ValueError):
        model.get_input_embeddings(ids[..., :5])


def test_forward_rejects_bad_shape():
    cfg = LMConfig.tiny(n_vq=4)
    model = MossTTSDelayModel(cfg, make_lm_weights(cfg, seed=1), device=DEV)
    with pytest.raises(ValueError):
        model.forward(torch.zeros(1, 4, 3, dtype=torch.int64))
    with pytest.raises(ValueError):
        model.generate(torch.zeros(1, 4, 3, dtype=torch.int64))


def test_generate_long_context_n_vq16():
    """configs[4] shape family (TTSD: n_vq = 16, long multi-speaker prompt with codes in the audio channels,
    ragged left padding, repetition penalty 1.1): KV lengths ~900-960 through the fused decode attention."""
    cfg = LMConfig.tiny(n_vq=16)
    w = make_lm_weights(cfg, seed=21, never_stop=True)
    model = MossTTSDelayModel(cfg, w, device=DEV)
    oracle = OracleLM(cfg, w, torch.bfloat16)
    ids, mask = make_prompt(cfg, 3, 900, seed=21, left_pad=[0, 37, 411])
    g = torch.Generator().manual_seed(3)
    codes = torch.randint(0, cfg.audio_vocab_size, (3, 600, cfg.n_vq), generator=g)
    ids[:, 200:800, 1:] = torch.where(mask[:, 200:800, None], codes, ids[:, 200:800, 1:])    # reference audio in the prompt
    kw = dict(max_new_tokens=40, text_temperature=0, audio_temperature=0, audio_repetition_penalty=1.1)
    out = model.generate(ids, mask, **kw)
    o_out, trace = oracle.generate(ids, mask, return_trace=True, **kw)
    exact = _check_tokens(out, [o[1].numpy() for o in o_out], trace, [o[0] for o in o_out], "tiny-nvq16-long")
    print(f"[tiny-nvq16-long] greedy ids bit-exact vs oracle: {exact}")


@pytest.mark.parametrize("shape,cs", [("tiny", 1), ("tiny", 2), ("tiny", 8), ("mid", 4), ("mid", 8)])
def test_decode_attention_key_split_clusters(shape, cs):
    """The fused decode attention split over a thread-block cluster of `cs` CTAs per (row, kv head) (the small-batch /
    long-context launch shape) must give the oracle's greedy tokens for every cluster size, including contexts
    shorter than 16 * cs keys (empty CTAs) and ragged left padding."""
    from moss_tts_b200._lib import lib
    cfg = LMConfig.tiny(n_vq=4) if shape == "tiny" else LMConfig.mid(n_vq=8)
    w = make_lm_weights(cfg, seed=33, never_stop=True)
    model = MossTTSDelayModel(cfg, w, device=DEV)
    oracle = OracleLM(cfg, w, torch.bfloat16)
    assert lib().mt_debug_set_attn_cluster(3) != 0          # not a supported size
    assert lib().mt_debug_set_attn_cluster(cs) == 0
    try:
        for S0, left_pad, steps in ((5, [0, 2], 30), (300, [0, 123], 24)):
            ids, mask = make_prompt(cfg, 2, S0, seed=5 + S0, left_pad=left_pad)
            kw = dict(max_new_tokens=steps, text_temperature=0, audio_temperature=0)
            out = model.generate(ids, mask, **kw)
            o_out, trace = oracle.generate(ids, mask, return_trace=True, **kw)
            exact = _check_tokens(out, [o[1].numpy() for o in o_out], trace, [o[0] for o in o_out], f"{shape}-cs{cs}-S{S0}")
            print(f"[{shape} cs={cs} S0={S0}] greedy ids bit-exact vs oracle: {exact}")
    finally:
        lib().mt_debug_set_attn_cluster(0)
