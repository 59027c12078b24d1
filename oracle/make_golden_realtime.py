"""Mint MOSS-TTS-Realtime golden vectors by EXECUTING THE REFERENCE (TEST INFRASTRUCTURE ONLY).

Run in the build container only (needs /root/reference):
    PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden_realtime

The reference's own `MossTTSRealtimeInference` (moss_tts_realtime/mossttsrealtime/streaming_mossttsrealtime.py:35-357) is
driven through prefill() + step() with a real `MossTTSRealtime` model (modeling_mossttsrealtime.py:80-99: Qwen3 backbone +
`MossTTSRealtimeLocalTransformerForCausalLM`) loaded with deterministic synthetic weights.  `generate_local_transformer` is
decorated with `torch.compile(fullgraph=True)`; TorchDynamo is disabled through its documented environment switch so the
decorated function runs as the eager Python it wraps (no source is modified).  Greedy decoding (do_sample=False) WITH the
windowed repetition penalty, which the reference applies before the argmax — so the goldens pin the penalty window too.
Stored per case: prompt, every frame's tokens, the per-channel top-2 raw logits of every frame (near-tie evidence, captured
by a forward hook on the local heads), and the backbone's last hidden state of every step.
"""
from __future__ import annotations

import json
import os
import sys

os.environ.setdefault("TORCHDYNAMO_DISABLE", "1")

import numpy as np
import torch

REF = os.environ.get("MOSS_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

from moss_tts_b200.config import RealtimeConfig                     # noqa: E402
from oracle.weights import make_realtime_weights, _gen               # noqa: E402


class _Tok:
    pad_token_id = None


def build_ref_realtime(cfg: RealtimeConfig, weights, dtype):
    sys.path.insert(0, os.path.join(REF, "moss_tts_realtime"))
    from transformers import Qwen3Config
    from mossttsrealtime.configuration_mossttsrealtime import MossTTSRealtimeConfig, MossTTSRealtimeLocalTransformerConfig
    from mossttsrealtime.modeling_mossttsrealtime import MossTTSRealtime
    lc = Qwen3Config(hidden_size=cfg.hidden_size, intermediate_size=cfg.intermediate_size,
                     num_hidden_layers=cfg.num_layers, num_attention_heads=cfg.num_heads,
                     num_key_value_heads=cfg.num_kv_heads, head_dim=cfg.head_dim, vocab_size=cfg.vocab_size,
                     max_position_embeddings=8192, rope_theta=cfg.rope_theta, rms_norm_eps=cfg.rms_norm_eps,
                     pad_token_id=cfg.pad_token_id)
    loc = MossTTSRealtimeLocalTransformerConfig(
        head_dim=cfg.local_head_dim, hidden_size=cfg.hidden_size, rms_norm_eps=cfg.local_rms_norm_eps,
        num_hidden_layers=cfg.local_num_layers, intermediate_size=cfg.local_intermediate_size,
        num_attention_heads=cfg.local_num_heads, num_key_value_heads=cfg.local_num_kv_heads,
        rope_theta=cfg.local_rope_theta, max_position_embeddings=33)
    loc.rvq = cfg.rvq
    mc = MossTTSRealtimeConfig(language_config=lc, local_config=loc, rvq=cfg.rvq, attn_implementation="eager")
    model = MossTTSRealtime(mc).eval()
    missing, unexpected = model.load_state_dict(weights, strict=False)
    assert not unexpected, unexpected
    assert all("rotary" in m or "language_model.embed_tokens" in m for m in missing), missing
    inv_g = model.language_model.rotary_emb.inv_freq.clone()
    inv_l = model.local_transformer.model.rotary_emb.inv_freq.clone()
    model = model.to(dtype)
    for rot, inv in ((model.language_model.rotary_emb, inv_g), (model.local_transformer.model.rotary_emb, inv_l)):
        rot.inv_freq = inv                                   # from_pretrained keeps the buffer fp32
        rot.original_inv_freq = inv.clone()
    for sub in (model, model.language_model, model.local_transformer, model.local_transformer.model):
        sub.config._attn_implementation = "eager"
    return model


def make_rt_prompt(cfg: RealtimeConfig, batch: int, lens, prefix_lens, seed: int):
    """Per-row system/voice prompt [T, 1+rvq] (text ids in channel 0, audio codes or pad in the others) and text prefix."""
    g = _gen(seed, "rt_prompt")
    prompts, prefixes = [], []
    for b in range(batch):
        T = lens[b]
        ids = np.full((T, cfg.rvq + 1), cfg.audio_pad_code, dtype=np.int64)
        ids[:, 0] = torch.randint(0, cfg.pad_token_id, (T,), generator=g).numpy()
        n_audio = T // 2                                     # a voice-prompt span: audio codes under reference_audio_pad
        ids[:n_audio, 0] = cfg.reference_audio_pad
        ids[:n_audio, 1:] = torch.randint(0, 1024, (n_audio, cfg.rvq), generator=g).numpy()
        prompts.append(ids)
        prefixes.append(torch.randint(0, cfg.pad_token_id, (prefix_lens[b],), generator=g).tolist())
    return prompts, prefixes


@torch.no_grad()
def run_reference(model, cfg, prompts, prefixes, steps, text_tokens, eos_token, penalty, window):
    from mossttsrealtime.streaming_mossttsrealtime import MossTTSRealtimeInference
    inf = MossTTSRealtimeInference(model, _Tok(), max_length=steps + 4, channels=cfg.rvq,
                                   audio_channel_pad=cfg.audio_pad_code, audio_bos_token=cfg.audio_bos_token,
                                   audio_eos_token=eos_token, text_pad_id=cfg.text_pad_id,
                                   aud_pad_id=cfg.reference_audio_pad)
    top2v, top2i, hidden = [], [], []
    frame_v, frame_i = [], []

    def head_hook(_m, _inp, out):
        v, i = torch.topk(out.float(), 2, dim=-1)
        frame_v.append(v)
        frame_i.append(i)

    hooks = [h.register_forward_hook(head_hook) for h in model.local_transformer.local_lm_heads]
    orig_glt = inf.generate_local_transformer

    def spy(hidden_states, **kw):
        hidden.append(hidden_states[:, 0, :].float().clone())
        frame_v.clear(); frame_i.clear()
        out = orig_glt(hidden_states=hidden_states, **kw)
        top2v.append(torch.stack(frame_v)); top2i.append(torch.stack(frame_i))
        return out

    inf.generate_local_transformer = spy
    kw = dict(temperature=0.8, top_p=0.6, top_k=30, do_sample=False, repetition_penalty=penalty, repetition_window=window)
    frames = [inf.prefill(prompts, prefixes, **kw)]
    stopping = [inf._is_stopping.clone()]
    for s in range(steps - 1):
        if inf.is_finished:
            break
        frames.append(inf.step(text_tokens[s] if s < len(text_tokens) else None, **kw))
        stopping.append(inf._is_stopping.clone())
    for h in hooks:
        h.remove()
    return (torch.stack(frames, dim=1), torch.stack(stopping, dim=1), inf.attention_mask.clone(),
            torch.stack(top2v), torch.stack(top2i), torch.stack(hidden))


def golden_rt(name, cfg, dtype, *, batch, lens, prefix_lens, steps, seed, penalty, window, eos_from=None):
    w = make_realtime_weights(cfg, seed=seed)
    model = build_ref_realtime(cfg, w, dtype)
    prompts, prefixes = make_rt_prompt(cfg, batch, lens, prefix_lens, seed)
    g = _gen(seed, "rt_text")
    n_text = steps // 2                                           # text deltas for the first half, then text_pad
    text_tokens = [torch.randint(0, cfg.pad_token_id, (batch,), generator=g).tolist() for _ in range(n_text)]
    eos = cfg.audio_eos_token
    if eos_from is not None:
        # make row `b` stop at frame `f`: declare the channel-0 token the free run emits there to be the eos token
        frames, *_ = run_reference(model, cfg, prompts, prefixes, steps, text_tokens, eos, penalty, window)
        b, f = eos_from
        eos = int(frames[b, f, 0])
    frames, stopping, mask, tv, ti, hid = run_reference(model, cfg, prompts, prefixes, steps, text_tokens, eos, penalty, window)
    # the padded prompt exactly as prefill() builds it (:139-176), restated for the consumers of the golden
    S0 = max(len(p) + len(t) for p, t in zip(prompts, prefixes))
    ids = np.full((batch, S0, cfg.rvq + 1), cfg.audio_pad_code, dtype=np.int64)
    ids[:, :, 0] = cfg.text_pad_id
    pm = np.zeros((batch, S0), dtype=bool)
    for b in range(batch):
        seg = np.full((len(prefixes[b]), cfg.rvq + 1), cfg.audio_pad_code, dtype=np.int64)
        seg[:, 0] = prefixes[b]
        seg[-1, 1] = cfg.audio_bos_token
        row = np.concatenate([prompts[b], seg], axis=0)
        ids[b, S0 - len(row):] = row
        pm[b, S0 - len(row):] = True
    assert np.array_equal(mask[:, :S0].numpy(), pm)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), input_ids=ids, attention_mask=pm,
                        text_tokens=np.asarray(text_tokens, dtype=np.int64).reshape(n_text, batch),
                        frames=frames.numpy(), stopping=stopping.numpy(), final_mask=mask.numpy(),
                        top2_val=tv.numpy(), top2_idx=ti.numpy(), hidden=hid.numpy(),
                        meta=json.dumps({"cfg": cfg.to_dict(), "dtype": str(dtype), "seed": seed, "steps": steps,
                                         "penalty": penalty, "window": window, "eos_token": eos,
                                         "torch": torch.__version__,
                                         "transformers": __import__("transformers").__version__,
                                         "reference": "xiami2019/MOSS-TTS@fe397ee"}))
    print(name, tuple(frames.shape), "eos", eos, "stopping", stopping[:, -1].tolist())


def main():
    os.makedirs(OUT, exist_ok=True)
    only = sys.argv[1:]
    if only:
        global golden_rt
        _all = golden_rt
        golden_rt = lambda name, *a, **k: _all(name, *a, **k) if name in only else None   # noqa: E731
    t4 = RealtimeConfig.tiny(4)
    golden_rt("rt_tiny_f32", t4, torch.float32, batch=2, lens=[9, 6], prefix_lens=[3, 2], steps=10, seed=11,
              penalty=1.3, window=3, eos_from=(1, 5))
    golden_rt("rt_tiny_bf16", t4, torch.bfloat16, batch=2, lens=[9, 6], prefix_lens=[3, 2], steps=10, seed=11,
              penalty=1.3, window=3, eos_from=(1, 5))
    mid = RealtimeConfig.mid(16)
    golden_rt("rt_mid_f32", mid, torch.float32, batch=3, lens=[12, 12, 7], prefix_lens=[4, 4, 2], steps=8, seed=13,
              penalty=1.1, window=50)
    golden_rt("rt_mid_bf16", mid, torch.bfloat16, batch=3, lens=[12, 12, 7], prefix_lens=[4, 4, 2], steps=8, seed=13,
              penalty=1.1, window=50)
    # every per-layer shape of the assumed 1.7 B model (H 2048, 16q/8kv x 128, I 6144, V_t 151936, the full 4-layer local
    # transformer, rvq 16) with 2 of the 28 backbone layers: the widths the decode GEMMs split K at and attn_decode<128> runs
    golden_rt("rt_wide_bf16", RealtimeConfig.wide(2), torch.bfloat16, batch=3, lens=[20, 14, 9], prefix_lens=[4, 3, 2],
              steps=6, seed=17, penalty=1.1, window=50)


if __name__ == "__main__":
    main()
