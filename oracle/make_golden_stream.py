"""Mints tests/golden/codec_stream_mini.npz by EXECUTING THE REFERENCE's incremental streaming decode:
`with model.streaming(batch_size): model._decode_frame(chunk)` over uneven pushes (reference
moss_audio_tokenizer/modeling_moss_audio_tokenizer.py:1393-1400, StreamingModule :121-177, RingKVCache :413-480,
MHAState :488-499) - the way moss_tts_realtime's AudioStreamDecoder drives the codec.  Build container only.
TEST INFRASTRUCTURE ONLY.

  PYTHONDONTWRITEBYTECODE=1 python oracle/make_golden_stream.py
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "tests", "golden")

from moss_tts_b200.config import CodecConfig          # noqa: E402
from oracle.weights import make_codec_weights, _gen   # noqa: E402
from oracle.make_golden import _versions              # noqa: E402
from oracle.make_golden_enc import build_ref_codec_full   # noqa: E402

PUSHES = [3, 10, 1, 7, 12, 2, 5]          # frames per push; context = 12 frames, so the ring wraps several times


def main():
    cfg = CodecConfig.mini()
    w = make_codec_weights(cfg, seed=11)
    model = build_ref_codec_full(cfg, w)
    g = _gen(31, "stream")
    T = sum(PUSHES)
    codes = torch.randint(0, cfg.codebook_size, (cfg.num_quantizers, 2, T), generator=g)
    rec = {"versions": _versions(), "cfg": json.dumps(cfg.to_dict()), "seed": 11, "codes": codes.numpy(),
           "pushes": np.asarray(PUSHES)}
    outs = []
    with torch.no_grad(), model.streaming(batch_size=2):
        pos = 0
        for n in PUSHES:
            o = model._decode_frame(codes[:, :, pos: pos + n])
            outs.append(o.audio)
            pos += n
    rec["wav_stream"] = torch.cat(outs, dim=-1).numpy()
    with torch.no_grad():
        rec["wav_full"] = model.decode(codes, return_dict=True).audio.numpy()
    np.savez_compressed(os.path.join(OUT, "codec_stream_mini.npz"), **rec)
    print("stream golden ok", rec["wav_stream"].shape, "stream-vs-full maxdiff",
          float(np.abs(rec["wav_stream"] - rec["wav_full"]).max()))

    # ---- incremental streaming ENCODE: `with model.streaming(B): model._encode_frame(chunk)` over the same uneven pushes
    # (the encoder's transformers keep their ring caches between pushes; the quantiser is per frame)
    wav = torch.randn(2, 1, T * cfg.downsample_rate, generator=_gen(32, "stream_enc")) * 0.3
    enc = {"versions": _versions(), "cfg": json.dumps(cfg.to_dict()), "seed": 11, "wav": wav.numpy(), "pushes": np.asarray(PUSHES)}
    codes_s, hid_s = [], []
    with torch.no_grad(), model.streaming(batch_size=2):
        pos = 0
        for n in PUSHES:
            o = model._encode_frame(wav[:, :, pos * cfg.downsample_rate: (pos + n) * cfg.downsample_rate])
            codes_s.append(o.audio_codes)
            hid_s.append(o.encoder_hidden_states)
            pos += n
    enc["codes_stream"] = torch.cat(codes_s, dim=-1).numpy()
    enc["hidden_stream"] = torch.cat(hid_s, dim=-1).numpy()
    with torch.no_grad():
        full = model._encode_frame(wav)
    enc["codes_full"] = full.audio_codes.numpy()
    np.savez_compressed(os.path.join(OUT, "codec_enc_stream_mini.npz"), **enc)
    print("encode stream golden ok", enc["codes_stream"].shape, "codes differing from the one-pass encode:",
          int((enc["codes_stream"] != enc["codes_full"]).sum()), "of", enc["codes_full"].size)


if __name__ == "__main__":
    main()
