"""Mint golden vectors AT THE BASELINE SHAPES by EXECUTING THE REFERENCE (TEST INFRASTRUCTURE ONLY).

Run in the build container only:   PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden_big [lm32|lm16|codec ...]

BASELINE.json configs[1] names a "teacher-forced logit and greedy-token parity run" at the assumed released 8B shape
and a waveform gate against the released decoder.  The toy goldens (make_golden.py) never reach the widths where the
product's decode GEMMs split K 3-4 ways (K = 4096 / 12288), `attn_decode_kernel<128>` runs 32 q / 8 kv heads, or 68
fp16-operand residual layers stack up.  These fixtures do:

  lm_wide_nvq32_bf16.npz   MossTTSDelayModel, H 4096, I 12288, 32q/8kv x 128, V_t 151936, n_vq 32, L = 2 (wide-shallow: every
                           per-layer kernel shape of the 8B model, 2 of its 36 layers), B = 4 with ragged left padding,
                           S0 = 24, 40 greedy steps, + teacher-forced logits of the reference on prompt + generated ids
  lm_wide_nvq16_bf16.npz   the same backbone at n_vq = 16 (TTSD family, configs[4]) with reference codes in the prompt's audio
                           channels and audio_repetition_penalty = 1.1
  codec_full.npz           released config.json decoder, all 32+12+12+12 layers, layer_scale 0.3, B = 1, T = 130 frames:
                           waveform unchunked and with chunk_duration = 8 (the processor's path)

Stored digests (the full logits would be 95 MB): generated ids; per (row, position, head) the reference's top-2 logit
values and indices (what the teacher-forced argmax test needs); the text head's top-16; all audio-head logits at a few
positions as raw bf16 bits; fp64 sums of every head's finite logits.  Weights are regenerated from oracle/weights.py seeds.
"""
from __future__ import annotations

import json
import os
import sys
import time

import numpy as np
import torch

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

from moss_tts_b200.config import LMConfig, CodecConfig                      # noqa: E402
from oracle import ref                                                       # noqa: E402
from oracle.weights import make_lm_weights, make_codec_weights, make_prompt, _gen   # noqa: E402

WIDE_SEED = 8
CODEC_SEED = 12


def _versions():
    import transformers
    return json.dumps({"torch": torch.__version__, "transformers": transformers.__version__,
                       "reference": "xiami2019/MOSS-TTS@fe397ee"})


def wide_cfg(n_vq: int, layers: int = 2) -> LMConfig:
    """Every width of the assumed released 8B shape, `layers` of its 36 layers."""
    c = LMConfig.delay_8b(n_vq)
    c.num_layers = layers
    return c


def wide_weights(cfg: LMConfig, base=None):
    """fp32 master weights for wide_cfg(n_vq).  The per-tensor seeds depend on the tensor name only, so the n_vq = 16
    model is the n_vq = 32 model's backbone, text head and first 16 audio tables / heads: pass `base` to reuse them."""
    if base is None:
        return make_lm_weights(cfg, seed=WIDE_SEED, never_stop=True)
    keep = {}
    for k, v in base.items():
        if k.startswith("emb_ext.") and int(k.split(".")[1]) >= cfg.n_vq:
            continue
        if k.startswith("lm_heads.") and int(k.split(".")[1]) > cfg.n_vq:
            continue
        keep[k] = v
    return keep


def wide_prompt(cfg: LMConfig, which: str):
    """Prompts of the two LM goldens (regenerated identically by the GPU tests)."""
    if which == "lm32":
        return make_prompt(cfg, 4, 24, seed=WIDE_SEED, left_pad=[0, 3, 7, 1])
    ids, mask = make_prompt(cfg, 3, 48, seed=WIDE_SEED + 1, left_pad=[0, 5, 2])
    g = _gen(WIDE_SEED + 1, "prompt_codes")
    codes = torch.randint(0, cfg.audio_vocab_size, (3, 30, cfg.n_vq), generator=g)
    ids[:, 10:40, 1:] = torch.where(mask[:, 10:40, None], codes, ids[:, 10:40, 1:])     # reference audio in the prompt
    return ids, mask


def bf16_bits(t: torch.Tensor) -> np.ndarray:
    return t.to(torch.bfloat16).contiguous().view(torch.int16).numpy().view(np.uint16)


def golden_lm(which: str, base_weights=None):
    n_vq = 32 if which == "lm32" else 16
    cfg = wide_cfg(n_vq)
    t0 = time.time()
    w = wide_weights(cfg, base_weights)
    print(f"[{which}] weights {time.time() - t0:.0f}s")
    model = ref.build_lm(cfg, w, torch.bfloat16)
    ids, mask = wide_prompt(cfg, which)
    B, S0, _ = ids.shape
    steps = 40
    penalty = 1.0 if which == "lm32" else 1.1
    t0 = time.time()
    out = model.generate(ids, mask, max_new_tokens=steps, text_temperature=0, audio_temperature=0,
                         audio_repetition_penalty=penalty)
    print(f"[{which}] reference generate {time.time() - t0:.0f}s")
    rec = {"versions": _versions(), "cfg": json.dumps(cfg.to_dict()), "dtype": "torch.bfloat16", "seed": WIDE_SEED,
           "which": which, "never_stop": True, "batch": B, "prompt_len": S0, "steps": steps, "penalty": penalty,
           "prompt_ids": ids.numpy().astype(np.int32), "prompt_mask": mask.numpy()}
    for b, (sl, g) in enumerate(out):
        rec[f"start_length{b}"] = int(sl)
        rec[f"gen{b}"] = g.numpy().astype(np.int32)
    new_rows = torch.stack([g[int(sl):] for sl, g in out])
    full_ids = torch.cat([ids, new_rows], dim=1)
    full_mask = torch.cat([mask, torch.ones(B, new_rows.shape[1], dtype=torch.bool)], dim=1)
    t0 = time.time()
    with torch.inference_mode():
        logits = model(input_ids=full_ids, attention_mask=full_mask).logits
    print(f"[{which}] reference teacher-forced forward {time.time() - t0:.0f}s")
    T = full_ids.shape[1]
    top2_v = np.zeros((B, T, 1 + n_vq, 2), np.float32)
    top2_i = np.zeros((B, T, 1 + n_vq, 2), np.int32)
    sums = np.zeros((B, T, 1 + n_vq), np.float64)
    for i, lg in enumerate(logits):
        lf = lg.float()
        v, ix = torch.topk(lf, 2, dim=-1)
        top2_v[:, :, i], top2_i[:, :, i] = v.numpy(), ix.numpy()
        sums[:, :, i] = torch.where(torch.isfinite(lf), lf, torch.zeros_like(lf)).double().sum(-1).numpy()
    v16, i16 = torch.topk(logits[0].float(), 16, dim=-1)
    pos = [S0 - 1, S0 + 5, S0 + 20, T - 1]
    rec.update(tf_ids=full_ids.numpy().astype(np.int32), tf_mask=full_mask.numpy(), tf_top2_v=top2_v, tf_top2_i=top2_i,
               tf_sum=sums, tf_text_top16_v=v16.numpy(), tf_text_top16_i=i16.numpy().astype(np.int32),
               tf_positions=np.asarray(pos),
               tf_audio_bits=np.stack([bf16_bits(lg[:, pos, :]) for lg in logits[1:]], axis=2))   # [B, P, n_vq, 1025]
    name = f"lm_wide_nvq{n_vq}_bf16"
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(name, "saved;", "gen rows", out[0][1].shape, "audio row 35:", out[0][1][35, :6].tolist())
    return w


def golden_codec():
    cfg = CodecConfig.released()
    cfg.enc_stages = []
    t0 = time.time()
    w = make_codec_weights(cfg, seed=CODEC_SEED, layer_scale=0.3)
    model = ref.build_codec(cfg, w)
    print(f"[codec] weights + reference model {time.time() - t0:.0f}s")
    T = 130
    codes = torch.randint(0, cfg.codebook_size, (cfg.num_quantizers, 1, T), generator=_gen(CODEC_SEED, "codes"))
    rec = {"versions": _versions(), "cfg": json.dumps(cfg.to_dict()), "seed": CODEC_SEED, "layer_scale": 0.3,
           "codes": codes.numpy().astype(np.int16)}
    # activation magnitudes of the reference's own fp32 run (what the fp16 operands of the CUDA path have to hold)
    peak = {}

    def hook(name):
        def f(mod, inp, out):
            x = inp[0] if isinstance(inp, tuple) else inp
            peak[name] = max(peak.get(name, 0.0), float(x.detach().abs().max()))
        return f
    handles = []
    for n, m in model.named_modules():
        if n.startswith("decoder") and isinstance(m, torch.nn.Linear) and ".layers." in n:
            stage = int(n.split(".")[1]) // 2
            kind = n.split(".")[-1] if "linear" in n else n.split(".")[-2]
            handles.append(m.register_forward_hook(hook(f"stage{stage}.{kind}")))
    t0 = time.time()
    with torch.no_grad():
        full = model.decode(codes, return_dict=True)
        print(f"[codec] reference decode {time.time() - t0:.0f}s")
        for h in handles:
            h.remove()
        chunked = model.decode(codes, return_dict=True, chunk_duration=8)
    rec["wav"] = full.audio.numpy()
    rec["wav_chunked"] = chunked.audio.numpy()
    rec["ref_linear_input_absmax"] = json.dumps(peak)
    np.savez_compressed(os.path.join(OUT, "codec_full.npz"), **rec)
    d = float(np.abs(rec["wav"] - rec["wav_chunked"]).max())
    print("codec_full saved; |wav| max", float(np.abs(rec["wav"]).max()), "chunked-vs-unchunked maxdiff", d)
    print("reference Linear-input |x| maxima:", {k: round(v, 2) for k, v in sorted(peak.items())})


def main(argv):
    os.makedirs(OUT, exist_ok=True)
    what = argv or ["lm32", "lm16", "codec"]
    base = None
    if "lm32" in what:
        base = golden_lm("lm32")
    if "lm16" in what:
        if base is None:
            base = wide_weights(wide_cfg(32))
        golden_lm("lm16", base)
    del base
    if "codec" in what:
        golden_codec()


if __name__ == "__main__":
    main(sys.argv[1:])
