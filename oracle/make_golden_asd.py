"""Mints tests/golden/stream_decoder_mini.npz by EXECUTING THE REFERENCE's AudioStreamDecoder
(moss_tts_realtime/mossttsrealtime/streaming_mossttsrealtime.py:682-799) over the reference codec (mini widths,
seeded weights): tokens pushed in uneven pieces, `audio_chunks()` drained after every push, `flush()` at the end.
Build container only.  TEST INFRASTRUCTURE ONLY.

  PYTHONDONTWRITEBYTECODE=1 python oracle/make_golden_asd.py
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
REF_RT = "/root/reference/moss_tts_realtime"

from moss_tts_b200.config import CodecConfig          # noqa: E402
from oracle.weights import make_codec_weights, _gen   # noqa: E402
from oracle.make_golden import _versions              # noqa: E402
from oracle.make_golden_enc import build_ref_codec_full   # noqa: E402

PUSHES = [5, 9, 13, 6]        # frames per push_tokens call
CHUNK, OVERLAP = 10, 2        # 10-frame chunks (> the 4-frame chunk_duration=0.32 ring rule), 2 frames of crossfade


def main():
    cfg = CodecConfig.mini()
    w = make_codec_weights(cfg, seed=11)
    model = build_ref_codec_full(cfg, w)
    sys.path.insert(0, REF_RT)
    from mossttsrealtime.streaming_mossttsrealtime import AudioStreamDecoder
    g = _gen(41, "asd")
    T = sum(PUSHES)
    tokens = torch.randint(0, cfg.codebook_size, (T, cfg.num_quantizers), generator=g)
    dec = AudioStreamDecoder(model, chunk_frames=CHUNK, overlap_frames=OVERLAP, device=torch.device("cpu"))
    rec = {"versions": _versions(), "cfg": json.dumps(cfg.to_dict()), "seed": 11, "tokens": tokens.numpy(),
           "pushes": np.asarray(PUSHES), "chunk_frames": CHUNK, "overlap_frames": OVERLAP}
    chunks, after_push = [], []
    with torch.no_grad():
        pos = 0
        for n in PUSHES:
            dec.push_tokens(tokens[pos: pos + n].numpy())
            pos += n
            for wav in dec.audio_chunks():
                chunks.append(wav.numpy())
            after_push.append(len(chunks))
        last = dec.flush()
        if last is not None:
            chunks.append(last.numpy())
    rec["n_chunks"] = len(chunks)
    rec["chunks_after_push"] = np.asarray(after_push)
    for i, ch in enumerate(chunks):
        rec[f"chunk{i}"] = ch
    # crossfade alone on seeded vectors (kernel known-answer: prev_tail, head -> merged)
    pt = torch.randn(3840, generator=g)
    hd = torch.randn(3840, generator=g)
    fo = torch.linspace(1.0, 0.0, 3840)
    rec["xf_prev"], rec["xf_head"], rec["xf_out"] = pt.numpy(), hd.numpy(), (pt * fo + hd * (1.0 - fo)).numpy()
    np.savez_compressed(os.path.join(OUT, "stream_decoder_mini.npz"), **rec)
    print("asd golden ok", [c.shape for c in chunks], after_push)


if __name__ == "__main__":
    main()
