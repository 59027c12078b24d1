"""oracle/stream.py (restatement of the reference's AudioStreamDecoder) vs the chunks the reference class itself
emitted (tests/golden/stream_decoder_mini.npz, minted by oracle/make_golden_asd.py)."""
import json
import os

import numpy as np
import torch

from conftest import GOLDEN
from moss_tts_b200.config import CodecConfig
from oracle.codec import OracleCodec
from oracle.stream import OracleStreamDecoder, crossfade, linspace_fade_out
from oracle.weights import make_codec_weights


def _g():
    return np.load(os.path.join(GOLDEN, "stream_decoder_mini.npz"), allow_pickle=False)


def test_crossfade_known_answer():
    g = _g()
    np.testing.assert_allclose(crossfade(g["xf_prev"], g["xf_head"]), g["xf_out"], rtol=0, atol=1e-6)
    for n in (1, 2, 3, 7, 1152, 3840):
        np.testing.assert_array_equal(linspace_fade_out(n), torch.linspace(1.0, 0.0, n).numpy())


def test_stream_decoder_matches_reference_chunks():
    g = _g()
    cfg = CodecConfig.from_dict(json.loads(str(g["cfg"])))
    dec = OracleStreamDecoder(OracleCodec(cfg, make_codec_weights(cfg, seed=int(g["seed"]))),
                              chunk_frames=int(g["chunk_frames"]), overlap_frames=int(g["overlap_frames"]))
    chunks, after, pos = [], [], 0
    for n in g["pushes"]:
        dec.push_tokens(g["tokens"][pos: pos + int(n)])
        pos += int(n)
        chunks += dec.audio_chunks()
        after.append(len(chunks))
    last = dec.flush()
    if last is not None:
        chunks.append(last)
    assert after == g["chunks_after_push"].tolist() and len(chunks) == int(g["n_chunks"])
    for i, ch in enumerate(chunks):
        ref = g[f"chunk{i}"]
        assert ch.shape == ref.shape
        np.testing.assert_allclose(ch, ref, rtol=0, atol=2e-5 * max(1.0, float(np.abs(ref).max())))
    assert dec.flush() is None
