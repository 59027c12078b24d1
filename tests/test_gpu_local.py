"""MossTTSLocal CUDA path (mt_local_* through the C ABI) vs the oracle and the reference-minted goldens."""
import numpy as np
import pytest
import torch

from moss_tts_b200.config import LocalConfig
from moss_tts_b200.modeling_moss_tts_local import MossTTSLocalModel
from oracle.local import OracleLocal
from oracle.weights import make_local_weights, make_prompt
from gpu_util import DEV, bf16_ulp
from test_oracle_local import load_local_golden

pytestmark = pytest.mark.gpu

LOGIT_TOL = 2e-2      # north_star: teacher-forced logits max-abs <= 2e-2 (bf16)


def _truth(cfg, w):
    """fp32 evaluation of the same bf16-rounded weights: what both bf16 implementations approximate."""
    return OracleLocal(cfg, {k: v.to(torch.bfloat16).float() for k, v in w.items()}, torch.float32)


def _noise_floor(cfg, w, ctx, frame):
    """max |bf16 oracle - fp32 truth| over one teacher-forced frame: the reference arithmetic's own bf16 noise
    for these weights.  The depth chain stacks ~45 bf16-rounded Linears per channel, and with the synthetic
    mid-size weights that noise alone exceeds 2e-2, so tolerances are max(LOGIT_TOL, this floor)."""
    r = OracleLocal(cfg, w, torch.bfloat16).frame_logits(ctx, frame)
    t = _truth(cfg, w).frame_logits(ctx, frame)
    worst = 0.0
    for a, b_ in zip(r, t):
        fin = torch.isfinite(b_)
        worst = max(worst, (a.float() - b_)[fin].abs().max().item())
    return max(LOGIT_TOL, worst)


def _check(out, o_out, trace, what, floor=LOGIT_TOL):
    """Greedy ids bit-exact vs the oracle, or first divergence = near-tie of the oracle's own logits: two logit
    rows within `floor` of each other can only swap argmax inside that margin."""
    exact = True
    for b, ((sl, mine), (osl, ref)) in enumerate(zip(out, o_out)):
        assert sl == osl, (what, b, sl, osl)
        mine, ref = mine.cpu().numpy(), ref.numpy()
        assert mine.shape == ref.shape, (what, b, mine.shape, ref.shape)
        bad = sorted(map(tuple, np.argwhere(mine != ref)))
        if not bad:
            continue
        exact = False
        r, c = bad[0]
        f = int(r) - (sl + 1)
        lg = trace[f][c][b].float()
        top = lg.max().item()
        margin = top - lg[int(mine[r, c])].item()
        print(f"[{what}] row {b}: first divergence frame {f} ch {c}: mine {mine[r, c]} oracle {ref[r, c]} "
              f"margin {margin:.4g} = {margin / bf16_ulp(top):.1f} bf16 ulp (floor {floor:.4g}); {len(bad)} cells differ downstream")
        assert margin <= floor, f"{what}: divergence is not a near-tie (margin {margin:.4g} > {floor:.4g})"
    return exact


@pytest.mark.parametrize("name", ["local_tiny_bf16", "local_mid_bf16"])
def test_local_greedy_matches_oracle_and_golden(name):
    g, meta, cfg = load_local_golden(name)
    w = make_local_weights(cfg, seed=meta["seed"])
    model = MossTTSLocalModel(cfg, w, device=DEV)
    ids = torch.from_numpy(g["input_ids"])
    S0 = ids.shape[1]
    kw = dict(max_new_tokens=meta["steps"], eos_token_id=cfg.audio_end_token_id)
    out = model.generate(ids, **kw)
    o_out, trace = OracleLocal(cfg, w, torch.bfloat16).generate(ids, meta["steps"], eos_token_id=cfg.audio_end_token_id,
                                                                return_trace=True)
    floor = _noise_floor(cfg, w, ids, torch.from_numpy(g["gen_ids"])[:, S0])
    exact = _check(out, o_out, trace, name, floor)
    # against the reference's ids: same rule, margin taken from the reference's own top-2 logits
    tv = g["top2_val"]
    ref_exact = True
    for b, (sl, mine) in enumerate(out):
        mine = mine.cpu().numpy()[sl + 1:]
        ref = g["gen_ids"][b, S0:]
        bad = sorted(map(tuple, np.argwhere(mine != ref)))
        if bad:
            ref_exact = False
            f, c = bad[0]
            margin = float(tv[f, c, b, 0] - tv[f, c, b, 1])
            print(f"[{name}] vs reference row {b}: first divergence frame {f} ch {c}, reference margin {margin:.4g} "
                  f"= {margin / bf16_ulp(tv[f, c, b, 0]):.1f} ulp (floor {floor:.4g})")
            assert margin <= floor
    print(f"[{name}] greedy ids bit-exact vs oracle: {exact}; vs reference golden: {ref_exact}")
    out2 = model.generate(ids, use_cuda_graph=False, **kw)
    for (a, x), (b_, y) in zip(out, out2):
        assert a == b_ and torch.equal(x, y)


@pytest.mark.parametrize("name", ["local_tiny_bf16", "local_mid_bf16"])
def test_local_teacher_forced_logits(name):
    """Every channel's logits of a frame, fed the reference's own tokens (frame 0 after the prompt; a later frame
    after prompt + reference frames), against (1) the bf16 oracle, (2) the reference's recorded top-1 logit and
    (3) the fp32 truth, where the CUDA path must not be further away than the bf16 oracle itself is."""
    g, meta, cfg = load_local_golden(name)
    w = make_local_weights(cfg, seed=meta["seed"])
    model = MossTTSLocalModel(cfg, w, device=DEV)
    oracle = OracleLocal(cfg, w, torch.bfloat16)
    truth = _truth(cfg, w)
    gen = torch.from_numpy(g["gen_ids"])
    S0 = g["input_ids"].shape[1]
    tv, ti = g["top2_val"], g["top2_idx"]
    for f in (0, 3):
        ctx, frame = gen[:, :S0 + f], gen[:, S0 + f]
        mine = model.frame_logits(ctx, frame)
        ref = oracle.frame_logits(ctx, frame)
        tru = truth.frame_logits(ctx, frame)
        worst = floor = 0.0
        e_mine = e_ref = 0.0
        for c, (m, r, t) in enumerate(zip(mine, ref, tru)):
            m = m.float().cpu()
            r = r.float()
            assert torch.equal(torch.isinf(m), torch.isinf(r)), f"channel {c}: -inf pattern differs"
            fin = torch.isfinite(r)
            worst = max(worst, (m[fin] - r[fin]).abs().max().item())
            floor = max(floor, (r[fin] - t[fin]).abs().max().item())
            e_mine += (m[fin] - t[fin]).abs().mean().item()
            e_ref += (r[fin] - t[fin]).abs().mean().item()
        tol = max(LOGIT_TOL, floor)
        for c, m in enumerate(mine):      # the reference's largest logit of this (frame, channel, row)
            for b in range(m.shape[0]):
                assert abs(m[b, int(ti[f, c, b, 0])].item() - float(tv[f, c, b, 0])) <= tol, (name, f, c, b)
        print(f"[{name}] frame {f}: max-abs vs bf16 oracle {worst:.4g} (bf16 noise floor {floor:.4g}); mean error vs fp32 "
              f"truth: cuda {e_mine / len(mine):.5f}, bf16 oracle {e_ref / len(mine):.5f}")
        assert worst <= tol
        assert e_mine <= 1.25 * e_ref + 1e-4 * len(mine)


def test_local_eos_and_partial_channels():
    cfg = LocalConfig.tiny(4)
    w = make_local_weights(cfg, seed=11)
    model = MossTTSLocalModel(cfg, w, device=DEV)
    oracle = OracleLocal(cfg, w, torch.bfloat16)
    ids, _ = make_prompt(cfg, 3, 9, seed=11)
    base = model.generate(ids, max_new_tokens=3, eos_token_id=None)
    tok0 = int(base[1][1][-3, 0])                          # row 1's first generated text token becomes "EOS"
    out = model.generate(ids, max_new_tokens=7, eos_token_id=tok0)
    o_out, trace = oracle.generate(ids, 7, eos_token_id=tok0, return_trace=True)
    _check(out, o_out, trace, "local-eos")
    r1 = out[1][1].cpu()[out[1][0] + 1:]
    assert int(r1[0, 0]) == tok0
    assert (r1[1:, 0] == tok0).all() and (r1[1:, 1:] == cfg.audio_pad_code).all()
    part = model.generate(ids, max_new_tokens=3, n_vq_for_inference=2, eos_token_id=None)
    o_part, tr = oracle.generate(ids, 3, n_vq_for_inference=2, eos_token_id=None, return_trace=True)
    _check(part, o_part, tr, "local-partial")
    assert (part[0][1][-3:, 3:] == 0).all()


def test_local_sampling_follows_oracle():
    """HF processor order (penalty -> temperature -> top-k -> top-p) and the shared counter-hash uniforms.
    Frame 0 is checked EXACTLY: the logits the CUDA path sampled from are re-read through frame_logits (same
    kernels, same shapes, deterministic) and pushed through the oracle's sampler with the same uniforms."""
    from oracle import sampling as S
    cfg = LocalConfig.tiny(4)
    C_ = cfg.n_vq + 1
    w = make_local_weights(cfg, seed=5)
    model = MossTTSLocalModel(cfg, w, device=DEV)
    ids, _ = make_prompt(cfg, 3, 10, seed=5)
    gen = torch.Generator().manual_seed(9)
    ids[:, -6:-1, 1:] = torch.randint(0, cfg.audio_vocab_size, (3, 5, cfg.n_vq), generator=gen)   # history for the penalty
    layers = [dict(temperature=1.5, top_k=50, top_p=1.0)] + \
             [dict(temperature=1.1, top_k=25, top_p=0.9, repetition_penalty=1.3) for _ in range(cfg.n_vq)]
    for do_samples in ([True] * C_, [False] + [True] * cfg.n_vq):
        kw = dict(do_samples=do_samples, layers=layers, eos_token_id=None, seed=77)
        out = model.generate(ids, max_new_tokens=5, **kw)
        for b, (sl, rows) in enumerate(out):
            rows = rows.cpu()
            assert rows.shape == (sl + 1 + 5, C_)
            assert int(rows[sl + 1:, 1:].max()) < cfg.audio_pad_code and int(rows.min()) >= 0     # pad code is never drawn
        frame0 = torch.stack([rows.cpu()[sl + 1] for sl, rows in out])
        logits = [l.cpu() for l in model.frame_logits(ids, frame0)]
        for c in range(C_):
            for b in range(ids.shape[0]):
                row = logits[c][b].clone()
                if do_samples[c]:
                    lc = layers[c]
                    if lc.get("repetition_penalty") is not None and c != 0:
                        row = S.apply_penalty_row(row, np.unique(ids[b, :, c].numpy()), lc["repetition_penalty"])
                    row = row / lc["temperature"]
                    tok = S.sample_row(row, top_p=lc.get("top_p"), top_k=lc.get("top_k"), do_sample=True,
                                       u=S.uniform_u(77, 0, b, c))
                else:
                    tok = int(torch.argmax(row.float()).item())
                assert tok == int(frame0[b, c]), (do_samples, b, c, tok, int(frame0[b, c]))
    # whole-run comparison with the oracle is informational: the first rounding difference forks a sampled sequence
    o_out = OracleLocal(cfg, w, torch.bfloat16).generate(ids, 5, **kw)
    same = sum(int((a[1].cpu() == b_[1]).sum()) for a, b_ in zip(out, o_out))
    print(f"[local-sampling] cells identical to the oracle's run: {same}/{sum(b_[1].numel() for b_ in o_out)}")


def test_local_rejects_bad_inputs():
    cfg = LocalConfig.tiny(4)
    model = MossTTSLocalModel(cfg, make_local_weights(cfg, seed=1), device=DEV)
    ids, _ = make_prompt(cfg, 2, 8, seed=1)
    with pytest.raises(ValueError):
        model.generate(torch.zeros(1, 4, 3, dtype=torch.int64))
    mask = torch.ones(2, 8, dtype=torch.bool)
    mask[1, 3] = False                             # a hole that is not left padding
    with pytest.raises(ValueError, match="left padding"):
        model.generate(ids, mask, max_new_tokens=2)
    # sampled text channel without top_k: the full-vocabulary sampler (no filter, like HF without a TopKLogitsWarper).
    # The oracle's sampler on the logits the CUDA path sampled from, same uniform -> the same token (or its neighbour in
    # the sorted order: with 150,000 candidates the draw resolves masses of 1e-6).
    from oracle import sampling as S
    out = model.generate(ids, max_new_tokens=2, do_samples=[True] + [False] * cfg.n_vq, layers=[{} for _ in range(5)], seed=3)
    lg = model.frame_logits(ids, torch.zeros(2, 5, dtype=torch.int64))[0].cpu()
    for b, (sl, a) in enumerate(out):
        cand, probs = S.filtered_distribution(lg[b], None, None)
        want = S.draw(cand, probs, S.uniform_u(3, 0, b, 0))
        pos = {int(c_): j for j, c_ in enumerate(cand.tolist())}
        got = int(a[sl + 1, 0])
        assert got in pos and abs(pos[got] - pos[want]) <= 2, (b, got, want)


def test_local_left_padded_prompts_match_oracle_and_golden():
    """a15: a left-padded batch (rows of 12, 9 and 7 real tokens).  The reference's modules were run with the mask HF's
    generation loop builds (given mask + ones) and its cumulative-sum position_ids (tests/golden/local_tiny_pad_*.npz);
    the CUDA path masks the padding in prefill and decode attention.  Greedy ids vs the oracle and vs the reference's:
    exact or first divergence a near-tie; teacher-forced logits of frame 0 within 2e-2; padding really changes the result."""
    g, meta, cfg = load_local_golden("local_tiny_pad_bf16")
    w = make_local_weights(cfg, seed=meta["seed"])
    model = MossTTSLocalModel(cfg, w, device=DEV)
    ids, mask = torch.from_numpy(g["input_ids"]), torch.from_numpy(g["attention_mask"])
    S0 = ids.shape[1]
    kw = dict(max_new_tokens=meta["steps"], eos_token_id=cfg.audio_end_token_id)
    out = model.generate(ids, mask, **kw)
    o_out, trace = OracleLocal(cfg, w, torch.bfloat16).generate(ids, meta["steps"], eos_token_id=cfg.audio_end_token_id,
                                                                return_trace=True, attention_mask=mask)
    _check(out, o_out, trace, "local_tiny_pad_bf16")
    for b, (sl, mine) in enumerate(out):
        mine = mine.cpu().numpy()[sl + 1:]
        ref = g["gen_ids"][b, S0:]
        bad = sorted(map(tuple, np.argwhere(mine != ref)))
        if bad:
            f, c = bad[0]
            v = g["top2_val"][f, c, b]
            assert (v[0] - v[1]) <= 2 * bf16_ulp(v[0]), ("vs reference", b, f, c, v)
    frame = torch.from_numpy(g["gen_ids"])[:, S0]
    lg = model.frame_logits(ids, frame, attention_mask=mask)
    want = OracleLocal(cfg, w, torch.bfloat16)
    h = want.global_forward(ids, [None] * cfg.num_layers, 0, cfg.n_vq, mask)
    cur = want.swiglu(h, "speech_embedding_to_local_mlp.")
    ref0 = want.channel_logits(want.local_forward(cur[:, None, :]), 0)
    assert (lg[0].float().cpu() - ref0.float()).abs().max().item() <= LOGIT_TOL
    no_mask = model.frame_logits(ids, frame)
    assert (no_mask[0][1:].float() - lg[0][1:].float()).abs().max().item() > 1e-2        # padded rows differ without the mask
    assert torch.equal(no_mask[0][0], lg[0][0])                                          # the un-padded row does not
    with pytest.raises(ValueError, match="left padding"):
        bad_mask = mask.clone(); bad_mask[1, 6] = False
        model.generate(ids, bad_mask, **kw)
