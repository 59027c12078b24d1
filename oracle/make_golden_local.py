"""Mint MossTTSLocal golden vectors by DRIVING THE REFERENCE'S OWN SUB-MODULES (TEST INFRASTRUCTURE ONLY).

Run in the build container only (needs /root/reference):
    PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden_local

The reference's `generate()` for this family does not run under the installed transformers (its `_sample`
override calls HF internals that were removed).  The loop below therefore follows
`CustomMixin._sample` (reference moss_tts_local/modeling_moss_tts.py:379-450) line by line, but every
tensor operation is executed by the reference's real modules loaded with deterministic synthetic weights:
    model.model(...)                      MosiTTSModel.forward :532-569 with a transformers DynamicCache
    model.speech_embedding_to_local_mlp   MossTTSMLP :47-95
    model.local_transformer               MossTTSLocalTransformer :178-292 (attn_implementation="sdpa": causal)
    model.local_to_speech_embedding_mlps, model.layer_norm_before_lm_heads, model.lm_heads
    model.model.embedding_list
Greedy decoding only (sampled ids depend on torch.multinomial's RNG stream and are not comparable).
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

REF = os.environ.get("MOSS_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

from moss_tts_b200.config import LocalConfig                      # noqa: E402
from oracle.weights import make_local_weights, make_prompt         # noqa: E402


def build_ref_local(cfg: LocalConfig, weights, dtype):
    sys.path.insert(0, REF)
    from transformers import Qwen3Config
    from moss_tts_local.configuration_moss_tts import MossTTSDelayConfig
    from moss_tts_local.modeling_moss_tts import MossTTSDelayModel
    lc = Qwen3Config(hidden_size=cfg.hidden_size, intermediate_size=cfg.intermediate_size,
                     num_hidden_layers=cfg.num_layers, num_attention_heads=cfg.num_heads,
                     num_key_value_heads=cfg.num_kv_heads, head_dim=cfg.head_dim, vocab_size=cfg.vocab_size,
                     max_position_embeddings=8192, rope_theta=cfg.rope_theta, rms_norm_eps=cfg.rms_norm_eps,
                     attn_implementation="eager")
    mc = MossTTSDelayConfig(language_config=lc, n_vq=cfg.n_vq, local_hidden_size=cfg.local_hidden_size,
                            local_ffn_hidden_size=cfg.local_ffn_hidden_size, local_num_layers=cfg.local_num_layers,
                            additional_mlp_ffn_hidden_size=cfg.additional_mlp_ffn_hidden_size)
    model = MossTTSDelayModel(mc).eval()
    missing, unexpected = model.load_state_dict(weights, strict=False)
    # the backbone's own embed_tokens is never used (inputs_embeds come from model.embedding_list)
    assert not unexpected and all("rotary" in m or "language_model.embed_tokens" in m for m in missing), (missing, unexpected)
    inv_freq = model.model.language_model.rotary_emb.inv_freq.clone()
    model = model.to(dtype)
    model.model.language_model.rotary_emb.inv_freq = inv_freq           # from_pretrained keeps it fp32
    model.model.language_model.rotary_emb.original_inv_freq = inv_freq.clone()
    model.model.language_model.config._attn_implementation = "eager"
    model.local_transformer.config._attn_implementation = "sdpa"        # is_causal=True only acts under sdpa
    for layer in model.local_transformer.layers:
        layer.self_attn.config._attn_implementation = "sdpa"
    return model


@torch.no_grad()
def ref_generate(model, input_ids, max_new_tokens, eos_token_id, attention_mask=None):
    """CustomMixin._sample :379-450, greedy, executed by the reference's modules."""
    from transformers import DynamicCache
    cfg = model.config
    B, S0, C = input_ids.shape
    n_inf = C - 1
    cache = DynamicCache(config=cfg.language_config)
    unfinished = torch.ones(B, dtype=torch.long)
    ids = input_ids
    cur = input_ids
    top2 = []           # per frame [C, B, 2] values / indices of the two largest logits (near-tie evidence)
    for _ in range(max_new_tokens):
        past = ids.shape[1] - cur.shape[1]
        if attention_mask is None:
            am = torch.ones(B, ids.shape[1], dtype=torch.long)
            pos = torch.arange(past, ids.shape[1])[None].expand(B, -1)
        else:
            # what HF's generation loop feeds a left-padded batch: the given mask extended with ones
            # (_update_model_kwargs_for_generation) and position_ids = cumsum(mask) - 1, 1 on the padding
            # (prepare_inputs_for_generation)
            am = torch.cat([attention_mask.long(), torch.ones(B, ids.shape[1] - S0, dtype=torch.long)], dim=1)
            pos = (am.cumsum(-1) - 1).masked_fill(am == 0, 1)[:, past:]
        out = model.model(input_ids=cur, attention_mask=am,
                          position_ids=pos, past_key_values=cache, use_cache=True, output_hidden_states=True,
                          return_dict=True, n_vq_for_inference=n_inf)
        g = out.hidden_states[-1][:, -1, :]
        local_in = torch.zeros(B, 0, model.local_transformer_config.hidden_size, dtype=g.dtype)
        cur_in = model.speech_embedding_to_local_mlp(g)
        toks, fv, fi = [], [], []
        for c in range(min(C, 1 + n_inf)):
            local_in = torch.cat([local_in, cur_in.unsqueeze(1)], dim=1)
            lo = model.local_transformer(input_ids=None, attention_mask=None, inputs_embeds=local_in)[0]
            lo = model.layer_norm_before_lm_heads[c](model.local_to_speech_embedding_mlps[c](lo))
            logit = model.lm_heads[c](lo[:, -1, :])
            if c != 0:
                logit[:, cfg.audio_pad_code] = -torch.inf
            tok = torch.argmax(logit, dim=-1)
            v2, i2 = torch.topk(logit.float(), 2, dim=-1)
            fv.append(v2)
            fi.append(i2)
            toks.append(tok)
            cur_in = model.speech_embedding_to_local_mlp(model.model.embedding_list[c](tok))
        nxt = torch.stack(toks, dim=-1)
        top2.append((torch.stack(fv), torch.stack(fi)))
        for i in range(C):
            pd = eos_token_id if i == 0 else cfg.audio_pad_code
            nxt[:, i] = nxt[:, i] * unfinished + pd * (1 - unfinished)
        ids = torch.cat([ids, nxt[:, None, :]], dim=1)
        cur = nxt[:, None, :]
        unfinished = unfinished & (ids[:, -1, 0] != eos_token_id).long()
        if unfinished.max() == 0:
            break
    return ids, torch.stack([t[0] for t in top2]), torch.stack([t[1] for t in top2])


def golden_local(name, cfg, dtype, *, batch, prompt_len, steps, seed, left_pad=None):
    w = make_local_weights(cfg, seed=seed)
    eos = cfg.audio_end_token_id
    model = build_ref_local(cfg, w, dtype)
    ids, mask = make_prompt(cfg, batch, prompt_len, seed=seed, left_pad=left_pad)
    out, tv, ti = ref_generate(model, ids, steps, eos, mask if left_pad is not None else None)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), input_ids=ids.numpy(), gen_ids=out.numpy(),
                        attention_mask=mask.numpy(),
                        top2_val=tv.numpy(), top2_idx=ti.numpy(),
                        meta=json.dumps({"cfg": cfg.to_dict(), "dtype": str(dtype), "seed": seed, "steps": steps,
                                         "torch": torch.__version__,
                                         "transformers": __import__("transformers").__version__,
                                         "reference": "xiami2019/MOSS-TTS@fe397ee"}))
    print(name, tuple(out.shape))


def main():
    os.makedirs(OUT, exist_ok=True)
    only = sys.argv[1:]
    if only:
        global golden_local
        _all = golden_local
        golden_local = lambda name, *a, **k: _all(name, *a, **k) if name in only else None   # noqa: E731
    golden_local("local_tiny_f32", LocalConfig.tiny(4), torch.float32, batch=2, prompt_len=10, steps=8, seed=3)
    golden_local("local_tiny_bf16", LocalConfig.tiny(4), torch.bfloat16, batch=2, prompt_len=10, steps=8, seed=3)
    golden_local("local_mid_f32", LocalConfig.mid(8), torch.float32, batch=2, prompt_len=12, steps=6, seed=5)
    golden_local("local_mid_bf16", LocalConfig.mid(8), torch.bfloat16, batch=2, prompt_len=12, steps=6, seed=5)
    # left-padded prompts (rows of 12, 9 and 7 real tokens)
    golden_local("local_tiny_pad_f32", LocalConfig.tiny(4), torch.float32, batch=3, prompt_len=12, steps=6, seed=9,
                 left_pad=[0, 3, 5])
    golden_local("local_tiny_pad_bf16", LocalConfig.tiny(4), torch.bfloat16, batch=3, prompt_len=12, steps=6, seed=9,
                 left_pad=[0, 3, 5])


if __name__ == "__main__":
    main()
