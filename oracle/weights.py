"""Deterministic synthetic weights shared by the oracle, the golden generator and
the GPU parity tests (TEST INFRASTRUCTURE ONLY).

Names follow the reference modules' state_dict keys so `make_golden.py` can
`load_state_dict` them into the real reference classes:
  LM    : reference moss_tts_delay/modeling_moss_tts.py:170-191 (+ Qwen3Model keys)
  codec : reference moss_audio_tokenizer/modeling_moss_audio_tokenizer.py:563-568,
          716-751,907-914,1046-1053,1217-1236
"""
from __future__ import annotations

import zlib
from typing import Dict

import torch

from moss_tts_b200.config import LMConfig, CodecConfig, LocalConfig, RealtimeConfig


def _gen(seed: int, name: str) -> torch.Generator:
    g = torch.Generator(device="cpu")
    g.manual_seed((seed * 1000003 + zlib.crc32(name.encode())) % (2**63 - 1))
    return g


def _normal(seed, name, shape, std, mean=0.0):
    return torch.empty(shape, dtype=torch.float32).normal_(mean, std, generator=_gen(seed, name))


def make_lm_weights(cfg: LMConfig, seed: int = 0, never_stop: bool = False,
                    std: float = 0.02) -> Dict[str, torch.Tensor]:
    """fp32 master weights.  Norm gains are perturbed around 1 so that a kernel
    that forgets a gain multiply fails parity.  `never_stop` copies text-head row
    151656 into row 151662 (SURVEY.md §8(d) "never-stop trick")."""
    w: Dict[str, torch.Tensor] = {}
    H, I = cfg.hidden_size, cfg.intermediate_size
    w["language_model.embed_tokens.weight"] = _normal(seed, "embed", (cfg.vocab_size, H), std)
    for l in range(cfg.num_layers):
        p = f"language_model.layers.{l}."
        w[p + "self_attn.q_proj.weight"] = _normal(seed, p + "q", (cfg.q_size, H), std)
        w[p + "self_attn.k_proj.weight"] = _normal(seed, p + "k", (cfg.kv_size, H), std)
        w[p + "self_attn.v_proj.weight"] = _normal(seed, p + "v", (cfg.kv_size, H), std)
        w[p + "self_attn.o_proj.weight"] = _normal(seed, p + "o", (H, cfg.q_size), std)
        w[p + "self_attn.q_norm.weight"] = _normal(seed, p + "qn", (cfg.head_dim,), 0.1, 1.0)
        w[p + "self_attn.k_norm.weight"] = _normal(seed, p + "kn", (cfg.head_dim,), 0.1, 1.0)
        w[p + "mlp.gate_proj.weight"] = _normal(seed, p + "g", (I, H), std)
        w[p + "mlp.up_proj.weight"] = _normal(seed, p + "u", (I, H), std)
        w[p + "mlp.down_proj.weight"] = _normal(seed, p + "d", (H, I), std)
        w[p + "input_layernorm.weight"] = _normal(seed, p + "ln1", (H,), 0.1, 1.0)
        w[p + "post_attention_layernorm.weight"] = _normal(seed, p + "ln2", (H,), 0.1, 1.0)
    w["language_model.norm.weight"] = _normal(seed, "norm", (H,), 0.1, 1.0)
    for i in range(cfg.n_vq):
        w[f"emb_ext.{i}.weight"] = _normal(seed, f"emb_ext{i}", (cfg.audio_head_size, H), std)
    # heads keep the HF init scale: |logits| stays below 4, where one bf16 ulp (<= 1.6e-2) is inside the
    # 2e-2 absolute logit tolerance of the north star
    w["lm_heads.0.weight"] = _normal(seed, "head0", (cfg.vocab_size, H), std)
    for i in range(cfg.n_vq):
        w[f"lm_heads.{i + 1}.weight"] = _normal(seed, f"head{i + 1}", (cfg.audio_head_size, H), std)
    if never_stop:
        w["lm_heads.0.weight"][cfg.audio_assistant_delay_slot_token_id] = \
            w["lm_heads.0.weight"][cfg.audio_assistant_gen_slot_token_id]
    return w


def _qwen_layer(w, seed, p, H, I, q_size, kv_size, head_dim, std):
    w[p + "self_attn.q_proj.weight"] = _normal(seed, p + "q", (q_size, H), std)
    w[p + "self_attn.k_proj.weight"] = _normal(seed, p + "k", (kv_size, H), std)
    w[p + "self_attn.v_proj.weight"] = _normal(seed, p + "v", (kv_size, H), std)
    w[p + "self_attn.o_proj.weight"] = _normal(seed, p + "o", (H, q_size), std)
    w[p + "self_attn.q_norm.weight"] = _normal(seed, p + "qn", (head_dim,), 0.1, 1.0)
    w[p + "self_attn.k_norm.weight"] = _normal(seed, p + "kn", (head_dim,), 0.1, 1.0)
    w[p + "mlp.gate_proj.weight"] = _normal(seed, p + "g", (I, H), std)
    w[p + "mlp.up_proj.weight"] = _normal(seed, p + "u", (I, H), std)
    w[p + "mlp.down_proj.weight"] = _normal(seed, p + "d", (H, I), std)
    w[p + "input_layernorm.weight"] = _normal(seed, p + "ln1", (H,), 0.1, 1.0)
    w[p + "post_attention_layernorm.weight"] = _normal(seed, p + "ln2", (H,), 0.1, 1.0)


def make_local_weights(cfg: LocalConfig, seed: int = 0, std: float = 0.02, eos_bias: float = 0.0) -> Dict[str, torch.Tensor]:
    """fp32 master weights of MossTTSLocal under the reference's state_dict keys
    (moss_tts_local/modeling_moss_tts.py:495-507, 573-611).  std 0.05 on the small adapters keeps the
    depth chain's activations O(1) so that rounding-point mistakes show up in the greedy ids."""
    w: Dict[str, torch.Tensor] = {}
    H, I, C = cfg.hidden_size, cfg.intermediate_size, cfg.n_vq + 1
    LH, LF, AF = cfg.local_hidden_size, cfg.local_ffn_hidden_size, cfg.additional_mlp_ffn_hidden_size
    w["model.embedding_list.0.weight"] = _normal(seed, "emb0", (cfg.vocab_size, H), std)
    for i in range(1, C):
        w[f"model.embedding_list.{i}.weight"] = _normal(seed, f"emb{i}", (cfg.audio_head_size, H), std)
    for l in range(cfg.num_layers):
        _qwen_layer(w, seed, f"model.language_model.layers.{l}.", H, I, cfg.q_size, cfg.kv_size, cfg.head_dim, std)
    w["model.language_model.norm.weight"] = _normal(seed, "gnorm", (H,), 0.1, 1.0)
    for l in range(cfg.local_num_layers):
        _qwen_layer(w, seed, f"local_transformer.layers.{l}.", LH, LF, cfg.q_size, cfg.kv_size, cfg.head_dim, 2 * std)
    w["local_transformer.norm.weight"] = _normal(seed, "lnorm", (LH,), 0.1, 1.0)
    a = 3 * std
    w["speech_embedding_to_local_mlp.gate_proj.weight"] = _normal(seed, "in.g", (AF, H), a)
    w["speech_embedding_to_local_mlp.up_proj.weight"] = _normal(seed, "in.u", (AF, H), a)
    w["speech_embedding_to_local_mlp.down_proj.weight"] = _normal(seed, "in.d", (LH, AF), a)
    for c in range(C):
        p = f"local_to_speech_embedding_mlps.{c}."
        w[p + "gate_proj.weight"] = _normal(seed, p + "g", (AF, LH), a)
        w[p + "up_proj.weight"] = _normal(seed, p + "u", (AF, LH), a)
        w[p + "down_proj.weight"] = _normal(seed, p + "d", (H, AF), a)
        w[f"layer_norm_before_lm_heads.{c}.weight"] = _normal(seed, f"hn{c}", (H,), 0.1, 1.0)
        w[f"lm_heads.{c}.weight"] = _normal(seed, f"head{c}", (cfg.vocab_size if c == 0 else cfg.audio_head_size, H), std)
    return w


def make_realtime_weights(cfg: RealtimeConfig, seed: int = 0, std: float = 0.02) -> Dict[str, torch.Tensor]:
    """fp32 master weights of MOSS-TTS-Realtime under the reference's state_dict keys
    (moss_tts_realtime/mossttsrealtime/modeling_mossttsrealtime.py:82-99,
    modeling_mossttsrealtime_local.py:258-268,381-389)."""
    w: Dict[str, torch.Tensor] = {}
    H, A = cfg.hidden_size, cfg.audio_head_size
    w["embed_tokens.0.weight"] = _normal(seed, "emb0", (cfg.vocab_size, H), std)
    for i in range(1, cfg.rvq + 1):
        w[f"embed_tokens.{i}.weight"] = _normal(seed, f"emb{i}", (A, H), std)
    for l in range(cfg.num_layers):
        _qwen_layer(w, seed, f"language_model.layers.{l}.", H, cfg.intermediate_size, cfg.q_size, cfg.kv_size,
                    cfg.head_dim, std)
    w["language_model.norm.weight"] = _normal(seed, "gnorm", (H,), 0.1, 1.0)
    for i in range(cfg.rvq - 1):
        w[f"local_transformer.model.embed_tokens.{i}.weight"] = _normal(seed, f"lemb{i}", (A, H), 50 * std)
    for l in range(cfg.local_num_layers):
        _qwen_layer(w, seed, f"local_transformer.model.layers.{l}.", H, cfg.local_intermediate_size, cfg.local_q_size,
                    cfg.local_kv_size, cfg.local_head_dim, 2 * std)
    w["local_transformer.model.norm.weight"] = _normal(seed, "lnorm", (H,), 0.1, 1.0)
    for i in range(cfg.rvq):
        w[f"local_transformer.local_lm_heads.{i}.weight"] = _normal(seed, f"lhead{i}", (A, H), std)
    return w


def make_codec_weights(cfg: CodecConfig, seed: int = 0, layer_scale: float = 0.3) -> Dict[str, torch.Tensor]:
    """fp32 codec decoder-side weights.  `layer_scale` defaults to 0.3 rather than the
    reference's init value 0.01 (config.json "layer_scale") so that attention/FFN errors
    are not hidden by a tiny residual gain."""
    w: Dict[str, torch.Tensor] = {}
    for q in range(cfg.num_quantizers):
        p = f"quantizer.quantizers.{q}."
        w[p + "codebook.weight"] = _normal(seed, p + "cb", (cfg.codebook_size, cfg.codebook_dim), 1.0)
        w[p + "out_proj.parametrizations.weight.original0"] = _normal(seed, p + "g", (cfg.rvq_dim, 1, 1), 0.2, 1.0)
        w[p + "out_proj.parametrizations.weight.original1"] = _normal(seed, p + "v", (cfg.rvq_dim, cfg.codebook_dim, 1), 0.3)
        w[p + "out_proj.bias"] = _normal(seed, p + "b", (cfg.rvq_dim,), 0.05)
    p = "quantizer.output_proj."
    w[p + "parametrizations.weight.original0"] = _normal(seed, p + "g", (cfg.code_dim, 1, 1), 0.2, 1.0)
    w[p + "parametrizations.weight.original1"] = _normal(seed, p + "v", (cfg.code_dim, cfg.rvq_dim, 1), 0.05)
    w[p + "bias"] = _normal(seed, p + "b", (cfg.code_dim,), 0.05)
    if cfg.enc_stages:
        # encode-side quantiser parameters (ResidualLFQ.input_proj :1217-1219, LFQ.in_proj :1047)
        p = "quantizer.input_proj."
        w[p + "parametrizations.weight.original0"] = _normal(seed, p + "g", (cfg.rvq_dim, 1, 1), 0.2, 1.0)
        w[p + "parametrizations.weight.original1"] = _normal(seed, p + "v", (cfg.rvq_dim, cfg.code_dim, 1), 0.05)
        w[p + "bias"] = _normal(seed, p + "b", (cfg.rvq_dim,), 0.05)
        for q in range(cfg.num_quantizers):
            p = f"quantizer.quantizers.{q}.in_proj."
            w[p + "parametrizations.weight.original0"] = _normal(seed, p + "g", (cfg.codebook_dim, 1, 1), 0.2, 1.0)
            w[p + "parametrizations.weight.original1"] = _normal(seed, p + "v", (cfg.codebook_dim, cfg.rvq_dim, 1), 0.1)
            w[p + "bias"] = _normal(seed, p + "b", (cfg.codebook_dim,), 0.05)
    stage_list = [(f"decoder.{2 * s}.", st) for s, st in enumerate(cfg.stages)] + \
                 [(f"encoder.{2 * s + 1}.", st) for s, st in enumerate(cfg.enc_stages)]
    for base, st in stage_list:
        d = st.d_model
        if st.input_dim != d:
            w[base + "input_proj.weight"] = _normal(seed, base + "ip", (d, st.input_dim), st.input_dim ** -0.5)
        if st.output_dim != d:
            w[base + "output_proj.weight"] = _normal(seed, base + "op", (st.output_dim, d), d ** -0.5)
        for l in range(st.num_layers):
            p = base + f"transformer.layers.{l}."
            w[p + "self_attn.in_projs.0.weight"] = _normal(seed, p + "in", (3 * d, d), d ** -0.5)
            w[p + "self_attn.out_projs.0.weight"] = _normal(seed, p + "out", (d, d), d ** -0.5)
            w[p + "norm1.weight"] = _normal(seed, p + "n1w", (d,), 0.1, 1.0)
            w[p + "norm1.bias"] = _normal(seed, p + "n1b", (d,), 0.1)
            w[p + "norm2.weight"] = _normal(seed, p + "n2w", (d,), 0.1, 1.0)
            w[p + "norm2.bias"] = _normal(seed, p + "n2b", (d,), 0.1)
            w[p + "linear1.weight"] = _normal(seed, p + "l1", (st.dim_feedforward, d), d ** -0.5)
            w[p + "linear2.weight"] = _normal(seed, p + "l2", (d, st.dim_feedforward), st.dim_feedforward ** -0.5)
            w[p + "layer_scale_1.scale"] = _normal(seed, p + "ls1", (d,), 0.1 * layer_scale, layer_scale)
            w[p + "layer_scale_2.scale"] = _normal(seed, p + "ls2", (d,), 0.1 * layer_scale, layer_scale)
    return w


def make_prompt(cfg: LMConfig, batch: int, prompt_len: int, seed: int = 0, left_pad=None):
    """Synthetic continuation-mode prompt (SURVEY.md §8(d) "Synthetic inputs"):
    channel 0 random text ids with im_start at S-4 and audio_start last, audio channels = pad
    code.  `left_pad[b]` pads row b on the left like processor._pad (reference
    processing_moss_tts.py:410-431)."""
    g = _gen(seed, "prompt")
    ids = torch.full((batch, prompt_len, 1 + cfg.n_vq), cfg.audio_pad_code, dtype=torch.int64)
    ids[:, :, 0] = torch.randint(0, cfg.pad_token_id, (batch, prompt_len), generator=g)
    ids[:, prompt_len - 4, 0] = cfg.im_start_token_id
    ids[:, prompt_len - 1, 0] = cfg.audio_start_token_id
    mask = torch.ones(batch, prompt_len, dtype=torch.bool)
    if left_pad is not None:
        for b, n in enumerate(left_pad):
            if n > 0:
                ids[b, :n, 0] = cfg.pad_token_id
                ids[b, :n, 1:] = cfg.audio_pad_code
                mask[b, :n] = False
    return ids, mask
