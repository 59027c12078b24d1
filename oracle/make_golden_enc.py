"""Mints tests/golden/codec_enc_mini.npz by EXECUTING THE REFERENCE's encode side:
MossAudioTokenizerModel._encode_frame / encode / batch_encode and quantizer.forward of
/root/reference/moss_audio_tokenizer/modeling_moss_audio_tokenizer.py, loaded with the deterministic synthetic
weights of oracle/weights.py.  Runs only in the build container (needs /root/reference); the GPU box reads the
committed .npz.  TEST INFRASTRUCTURE ONLY.

  PYTHONDONTWRITEBYTECODE=1 python oracle/make_golden_enc.py
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")

from moss_tts_b200.config import CodecConfig          # noqa: E402
from oracle.weights import make_codec_weights, _gen   # noqa: E402
from oracle.make_golden import _versions              # noqa: E402


def _tkw(st, max_period):
    return dict(causal=True, conv_layout=True, d_model=st.d_model, dim_feedforward=st.dim_feedforward, gating="none",
                input_dimension=st.input_dim, layer_scale=0.01, max_period=max_period, module_type="Transformer",
                norm="layer_norm", num_heads=st.num_heads, num_layers=st.num_layers, output_dimension=st.output_dim,
                positional_embedding="rope")


def build_ref_codec_full(cfg: CodecConfig, weights):
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
    dec, enc = [], []
    for st in cfg.stages:
        dec.append(_tkw(st, cfg.max_period))
        dec.append(dict(module_type="PatchedPretransform", patch_size=st.patch_size))
    for st in cfg.enc_stages:
        enc.append(dict(module_type="PatchedPretransform", patch_size=st.patch_size))
        enc.append(_tkw(st, cfg.max_period))
    q = dict(codebook_dim=cfg.codebook_dim, codebook_size=cfg.codebook_size, input_dim=cfg.code_dim,
             num_quantizers=cfg.num_quantizers, output_dim=cfg.code_dim, quantizer_type="rlfq", rvq_dim=cfg.rvq_dim)
    conf = MossAudioTokenizerConfig(sampling_rate=cfg.sampling_rate, downsample_rate=cfg.downsample_rate,
                                    causal_transformer_context_duration=cfg.context_duration,
                                    encoder_kwargs=enc, decoder_kwargs=dec, quantizer_type="rlfq", quantizer_kwargs=q)
    model = MossAudioTokenizerModel(conf).eval()
    missing, unexpected = model.load_state_dict(weights, strict=False)
    assert not unexpected and not missing, (missing, unexpected)
    return model


def main():
    os.makedirs(OUT, exist_ok=True)
    cfg = CodecConfig.mini()
    w = make_codec_weights(cfg, seed=11)
    model = build_ref_codec_full(cfg, w)
    g = _gen(11, "wav")
    rec = {"versions": _versions(), "cfg": json.dumps(cfg.to_dict()), "seed": 11}
    D = cfg.downsample_rate
    T = 24 * D + 700                                       # 24 full frames + a ragged tail (padded to 25)
    wav = torch.randn(2, 1, T, generator=g) * 0.3
    lens = torch.tensor([T, 17 * D + 5])
    rec["wav"] = wav.numpy()
    rec["wav_lengths"] = lens.numpy()
    with torch.no_grad():
        out = model._encode_frame(wav, lens)
        rec["codes"] = out.audio_codes.numpy()
        rec["codes_lengths"] = out.audio_codes_lengths.numpy()
        rec["hidden"] = out.encoder_hidden_states.numpy()
        # the quantiser alone on the reference's own hidden states
        zq, idx, _ = model.quantizer(out.encoder_hidden_states, out.audio_codes_lengths, 5)
        rec["codes_nq5"] = idx.numpy()
        rec["zq_nq5"] = zq.numpy()
        # public encode(): padding mask + num_quantizers
        pm = torch.zeros(2, T, dtype=torch.bool)
        pm[0, :] = True
        pm[1, : int(lens[1])] = True
        e2 = model.encode(wav, padding_mask=pm, num_quantizers=3, return_dict=True)
        rec["enc_nq3_codes"] = e2.audio_codes.numpy()
        rec["enc_nq3_lengths"] = e2.audio_codes_lengths.numpy()
        # batch_encode of ragged waveforms
        be = model.batch_encode([wav[0, 0, : 9 * D + 100], wav[1, 0, : 20 * D]])
        rec["batch_codes"] = be.audio_codes.numpy()
        rec["batch_lengths"] = be.audio_codes_lengths.numpy()
        # chunked streaming encode (B must be 1): 0.4 s chunks = 5 frames; 24 frames + tail -> 5 chunks
        ch = model.encode(wav[:1], return_dict=True, chunk_duration=0.4)
        rec["chunk_codes"] = ch.audio_codes.numpy()
        rec["chunk_lengths"] = ch.audio_codes_lengths.numpy()
        rec["chunk_hidden"] = ch.encoder_hidden_states.numpy()
        # encode -> decode round trip of the reference (size-independent property used at full size on the GPU)
        rt = model.decode(out.audio_codes, return_dict=True)
        rec["roundtrip_wav"] = rt.audio.numpy()
    np.savez_compressed(os.path.join(OUT, "codec_enc_mini.npz"), **rec)
    print("codec encode golden ok", rec["codes"].shape, rec["codes_lengths"], rec["chunk_codes"].shape,
          "chunked-vs-unchunked code agreement",
          float((rec["chunk_codes"][:, 0] == rec["codes"][:, 0, : rec["chunk_codes"].shape[-1]]).mean()))


if __name__ == "__main__":
    main()
