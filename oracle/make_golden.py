"""Mint golden vectors by EXECUTING THE REFERENCE (TEST INFRASTRUCTURE ONLY).

Run in the build container only (needs /root/reference, which does not exist on the
GPU box):   PYTHONDONTWRITEBYTECODE=1 python -m oracle.make_golden

The reference ships no golden vectors, so these fixtures are what pins
the oracle: deterministic synthetic weights (oracle/weights.py) are loaded into the
reference's own classes
  moss_tts_delay.modeling_moss_tts.MossTTSDelayModel          (generate :392-525, forward :225-300)
  moss_tts_delay.inference_utils                              (:19-165)
  moss_tts_delay.processing_moss_tts.MossTTSDelayProcessor    (:516-537, :668-709, :880-930)
  moss_audio_tokenizer...MossAudioTokenizerModel              (decode :1650-1755, decode_codes :1276-1283)
and their outputs are stored under tests/golden/*.npz together with the library versions.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

REF = os.environ.get("MOSS_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

from moss_tts_b200.config import LMConfig, CodecConfig          # noqa: E402
from oracle.weights import make_lm_weights, make_codec_weights, make_prompt, _gen   # noqa: E402


def _versions():
    import transformers
    return json.dumps({"torch": torch.__version__, "transformers": transformers.__version__,
                       "reference": "xiami2019/MOSS-TTS@fe397ee"})


def build_ref_lm(cfg: LMConfig, weights, dtype):
    sys.path.insert(0, REF)
    from transformers import Qwen3Config
    from moss_tts_delay.configuration_moss_tts import MossTTSDelayConfig
    from moss_tts_delay.modeling_moss_tts import MossTTSDelayModel
    lc = Qwen3Config(hidden_size=cfg.hidden_size, intermediate_size=cfg.intermediate_size,
                     num_hidden_layers=cfg.num_layers, num_attention_heads=cfg.num_heads,
                     num_key_value_heads=cfg.num_kv_heads, head_dim=cfg.head_dim, vocab_size=cfg.vocab_size,
                     max_position_embeddings=8192, rope_theta=cfg.rope_theta, rms_norm_eps=cfg.rms_norm_eps,
                     attn_implementation="eager")
    mc = MossTTSDelayConfig(language_config=lc, n_vq=cfg.n_vq)
    model = MossTTSDelayModel(mc).eval()
    missing, unexpected = model.load_state_dict(weights, strict=False)
    assert not unexpected and all("rotary" in m for m in missing), (missing, unexpected)
    inv_freq = model.language_model.rotary_emb.inv_freq.clone()
    model = model.to(dtype)
    # from_pretrained(dtype=bf16) keeps the (non-persistent, explicitly fp32) inv_freq buffer in fp32;
    # a blanket .to(bf16) would round it, which no deployment of the reference does.
    model.language_model.rotary_emb.inv_freq = inv_freq
    model.language_model.rotary_emb.original_inv_freq = inv_freq.clone()
    model.config._attn_implementation = "eager"
    model.language_model.config._attn_implementation = "eager"
    return model


def build_ref_codec(cfg: CodecConfig, weights):
    sys.path.insert(0, REF)
    import transformers.configuration_utils as cu
    orig = cu.PreTrainedConfig.__dict__.get("__init_subclass__")
    cu.PreTrainedConfig.__init_subclass__ = classmethod(lambda cls, **kw: None)   # SURVEY.md §8(c) import shim
    try:
        from moss_audio_tokenizer.configuration_moss_audio_tokenizer import MossAudioTokenizerConfig
    finally:
        if orig is not None:
            cu.PreTrainedConfig.__init_subclass__ = orig
        else:
            del cu.PreTrainedConfig.__init_subclass__
    from moss_audio_tokenizer.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
    dec = []
    for st in cfg.stages:
        dec.append(dict(causal=True, conv_layout=True, d_model=st.d_model, dim_feedforward=st.dim_feedforward,
                        gating="none", input_dimension=st.input_dim, layer_scale=0.01, max_period=cfg.max_period,
                        module_type="Transformer", norm="layer_norm", num_heads=st.num_heads,
                        num_layers=st.num_layers, output_dimension=st.output_dim, positional_embedding="rope"))
        dec.append(dict(module_type="PatchedPretransform", patch_size=st.patch_size))
    enc = [dict(module_type="PatchedPretransform", patch_size=cfg.downsample_rate)]   # encoder is out of scope
    q = dict(codebook_dim=cfg.codebook_dim, codebook_size=cfg.codebook_size, input_dim=cfg.code_dim,
             num_quantizers=cfg.num_quantizers, output_dim=cfg.code_dim, quantizer_type="rlfq", rvq_dim=cfg.rvq_dim)
    conf = MossAudioTokenizerConfig(sampling_rate=cfg.sampling_rate, downsample_rate=cfg.downsample_rate,
                                    causal_transformer_context_duration=cfg.context_duration,
                                    encoder_kwargs=enc, decoder_kwargs=dec, quantizer_type="rlfq",
                                    quantizer_kwargs=q)
    model = MossAudioTokenizerModel(conf).eval()
    missing, unexpected = model.load_state_dict(weights, strict=False)
    assert not unexpected, unexpected
    # only encode-side parameters (quantizer input projections, encoder) may be absent
    assert all(("in_proj." in m or "input_proj." in m or "encoder" in m) and "decoder" not in m
               for m in missing), missing
    return model


def head_digest(logits, positions):
    """Compact per-head summary: full audio rows, text top-16 + fp64 row sum over finite entries."""
    out = {}
    for i, lg in enumerate(logits):
        sel = lg[:, positions, :].float()
        if i == 0:
            v, ix = torch.topk(sel, 16, dim=-1)
            out["text_top_v"] = v.numpy()
            out["text_top_i"] = ix.numpy()
            out["text_sum"] = torch.where(torch.isfinite(sel), sel, torch.zeros_like(sel)).double().sum(-1).numpy()
        else:
            out[f"audio{i}"] = sel.numpy()
    return out


def golden_lm(name, cfg, dtype, *, batch, prompt_len, left_pad, steps, never_stop, seed, penalty=1.0):
    w = make_lm_weights(cfg, seed=seed, never_stop=never_stop)
    model = build_ref_lm(cfg, w, dtype)
    ids, mask = make_prompt(cfg, batch, prompt_len, seed=seed, left_pad=left_pad)
    out = model.generate(ids, mask, max_new_tokens=steps, text_temperature=0, audio_temperature=0,
                         audio_repetition_penalty=penalty)
    rec = {"versions": _versions(), "cfg": json.dumps(cfg.to_dict()), "dtype": str(dtype), "seed": seed,
           "never_stop": never_stop, "batch": batch, "prompt_len": prompt_len, "steps": steps,
           "left_pad": np.asarray(left_pad if left_pad else [0] * batch), "penalty": penalty,
           "prompt_ids": ids.numpy(), "prompt_mask": mask.numpy()}
    for b, (start_len, g) in enumerate(out):
        rec[f"start_length{b}"] = int(start_len)
        rec[f"gen{b}"] = g.numpy()
    # teacher-forced logits on prompt + generated ids (all rows have equal length here)
    new_rows = torch.stack([g[int(sl):] for sl, g in out])          # [B, n_new, C]
    full_ids = torch.cat([ids, new_rows], dim=1)
    full_mask = torch.cat([mask, torch.ones(batch, new_rows.shape[1], dtype=torch.bool)], dim=1)
    with torch.inference_mode():
        fo = model(input_ids=full_ids, attention_mask=full_mask)
    pos = [prompt_len - 1, prompt_len + 1, full_ids.shape[1] - 1]
    rec["tf_positions"] = np.asarray(pos)
    rec["tf_ids"] = full_ids.numpy()
    rec["tf_mask"] = full_mask.numpy()
    rec.update({"tf_" + k: v for k, v in head_digest(fo.logits, pos).items()})
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(name, "gen rows", out[0][1].shape, "first rows", out[0][1][:3, :3].tolist())


def golden_sampling():
    sys.path.insert(0, REF)
    from moss_tts_delay import inference_utils as R
    g = _gen(7, "sampling")
    rec = {"versions": _versions()}
    for tag, dtype, V in (("f32", torch.float32, 1025), ("bf16", torch.bfloat16, 1025), ("f32big", torch.float32, 4000)):
        x = (torch.randn(6, V, generator=g) * 2.0).to(dtype)
        x[:, -1] = float("-inf")
        rec[f"{tag}_logits"] = x.float().numpy()
        for k, p in ((25, 0.8), (50, 0.9), (5, 0.5), (25, 1.0)):
            y = R.apply_top_k(x.clone(), k)
            if p < 1.0:
                y = R.apply_top_p_optimized(y, p)
            pr = torch.softmax(y, dim=-1)
            rec[f"{tag}_k{k}_p{p}_probs"] = pr.float().numpy()
        prev = torch.randint(0, V, (3, 17, 2), generator=g)
        z = R.apply_repetition_penalty_delay_pattern(x.clone(), prev, 1.3)
        rec[f"{tag}_prev"] = prev.numpy()
        rec[f"{tag}_penalized"] = z.float().numpy()
        rec[f"{tag}_argmax"] = R.sample_token(x.clone(), do_sample=False).numpy()
    t = torch.randint(0, 5, (7, 23), generator=g)
    rec["fle_in"] = t.numpy()
    rec["fle_out3"] = R.find_last_equal_C(t, 3).numpy()
    rec["fle_out9"] = R.find_last_equal_C(t, 9).numpy()
    np.savez_compressed(os.path.join(OUT, "sampling.npz"), **rec)
    print("sampling ok")


def golden_sampling3d():
    """sample_token on [B, H, V] logits with per-head repetition penalty (inference_utils.py:88-106, "case 2")."""
    sys.path.insert(0, REF)
    from moss_tts_delay import inference_utils as R
    g = _gen(9, "sampling3d")
    rec = {"versions": _versions()}
    x = (torch.randn(3, 4, 1025, generator=g) * 2.0).to(torch.bfloat16)
    prev = torch.randint(0, 1025, (3, 40, 4), generator=g)
    prev[:, :, 2] = torch.randint(0, 64, (3, 40), generator=g)          # head 2 has a small id set: differs from the union
    rec["logits"] = x.float().numpy()
    rec["prev"] = prev.numpy()
    rec["penalized"] = R.apply_repetition_penalty_delay_pattern(x.clone(), prev, 1.3).float().numpy()
    rec["argmax"] = R.sample_token(x.clone(), prev_tokens=prev, repetition_penalty=1.3, do_sample=False).numpy()
    rec["argmax_union"] = R.sample_token(x.clone().view(12, 1025), prev_tokens=prev, repetition_penalty=1.3, do_sample=False).numpy()
    np.savez_compressed(os.path.join(OUT, "sampling3d.npz"), **rec)
    print("sampling3d ok; rows where per-head and union penalties pick different tokens:",
          int((rec["argmax"].reshape(-1) != rec["argmax_union"]).sum()))


def golden_codec():
    cfg = CodecConfig.mini()
    w = make_codec_weights(cfg, seed=3)
    model = build_ref_codec(cfg, w)
    g = _gen(3, "codes")
    rec = {"versions": _versions(), "cfg": json.dumps(cfg.to_dict()), "seed": 3}
    codes = torch.randint(0, cfg.codebook_size, (cfg.num_quantizers, 2, 40), generator=g)
    rec["codes"] = codes.numpy()
    with torch.no_grad():
        rec["emb"] = model.quantizer.decode_codes(codes).numpy()
        rec["emb_nq3"] = model.quantizer.decode_codes(codes[:3]).numpy()
        full = model.decode(codes, return_dict=True)
        rec["wav"] = full.audio.numpy()
        rec["wav_lengths"] = full.audio_lengths.numpy()
        pm = torch.ones(2, 40, dtype=torch.bool)
        pm[1, 29:] = False
        part = model.decode(codes, padding_mask=pm, return_dict=True, num_quantizers=5)
        rec["wav_nq5"] = part.audio.numpy()
        rec["wav_nq5_lengths"] = part.audio_lengths.numpy()
        # chunked streaming decode (B must be 1): 0.8 s chunks = 10 frames, 33 frames -> 4 chunks
        c1 = codes[:, :1, :33]
        ch = model.decode(c1, return_dict=True, chunk_duration=0.8)
        rec["wav_chunked"] = ch.audio.numpy()
        rec["wav_unchunked33"] = model.decode(c1, return_dict=True).audio.numpy()
        bd = model.batch_decode([codes[:, 0, :17], codes[:, 1, :40]])
        rec["wav_batch"] = bd.audio.numpy()
        rec["wav_batch_lengths"] = bd.audio_lengths.numpy()
    # processor glue
    from moss_tts_delay.processing_moss_tts import MossTTSDelayProcessor
    lmcfg = LMConfig.tiny(n_vq=cfg.num_quantizers)
    raw = torch.randint(0, 1024, (21, cfg.num_quantizers), generator=g)
    delayed = MossTTSDelayProcessor.apply_delay_pattern(raw, 1024)
    rec["dp_raw"] = raw.numpy()
    rec["dp_delayed"] = delayed.numpy()
    rec["dp_dedelayed"] = MossTTSDelayProcessor.apply_de_delay_pattern(delayed).numpy()

    class _MC:
        audio_pad_code = 1024
    # decode_audio_codes hard-codes chunk_duration=8 (:913-915) -> needs a >= 8 s context codec
    cfg10 = CodecConfig.mini(context_duration=10.0)
    model10 = build_ref_codec(cfg10, make_codec_weights(cfg10, seed=5))
    rec["cfg10"] = json.dumps(cfg10.to_dict())
    proc = object.__new__(MossTTSDelayProcessor)
    proc.model_config = _MC()
    proc.audio_tokenizer = model10
    # generation-shaped channel matrix: [audio_start row | delayed codes | trailing pad rows]
    gen_audio = torch.cat([torch.full((1, cfg.num_quantizers), 1024), delayed,
                           torch.full((2, cfg.num_quantizers), 1024)])
    wavs = proc._parse_audio_codes(1, gen_audio)
    rec["parse_in"] = gen_audio.numpy()
    rec["parse_n"] = len(wavs)
    for i, wv in enumerate(wavs):
        rec[f"parse_wav{i}"] = wv.numpy()
    np.savez_compressed(os.path.join(OUT, "codec_mini.npz"), **rec)
    print("codec ok", rec["wav"].shape, "chunked-vs-unchunked maxdiff",
          float(np.abs(rec["wav_chunked"] - rec["wav_unchunked33"]).max()))


def main():
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(0)
    golden_sampling()
    golden_sampling3d()
    golden_codec()
    tiny = LMConfig.tiny(n_vq=4)
    golden_lm("lm_tiny_f32", tiny, torch.float32, batch=2, prompt_len=12, left_pad=[0, 3], steps=16,
              never_stop=False, seed=1)
    golden_lm("lm_tiny_bf16", tiny, torch.bfloat16, batch=2, prompt_len=12, left_pad=[0, 3], steps=16,
              never_stop=False, seed=1)
    golden_lm("lm_tiny_pen_f32", tiny, torch.float32, batch=3, prompt_len=10, left_pad=None, steps=14,
              never_stop=True, seed=2, penalty=1.3)
    mid = LMConfig.mid(n_vq=32)
    golden_lm("lm_mid_f32", mid, torch.float32, batch=2, prompt_len=16, left_pad=[2, 0], steps=40,
              never_stop=True, seed=4)
    golden_lm("lm_mid_bf16", mid, torch.bfloat16, batch=2, prompt_len=16, left_pad=[2, 0], steps=40,
              never_stop=True, seed=4)


if __name__ == "__main__":
    main()
