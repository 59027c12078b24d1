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


class _FakeCodec:
    """codec.decode stand-in for the host-logic test: one sample per frame = the frame's first code."""
    device = None

    def __init__(self):
        self.calls = []

    def decode(self, tokens_t, chunk_duration=None, **kw):
        self.calls.append((tuple(tokens_t.shape), chunk_duration, kw))
        return {"audio": [tokens_t[0].float()[None, :]]}


def test_product_stream_decoder_host_logic_on_cpu():
    """Buffering / chunking / decode_kwargs handling of the product AudioStreamDecoder (no crossfade -> no kernel):
    same consumption order and chunk boundaries as the oracle's restatement; the crossfade path refuses CPU tensors."""
    import pytest
    from moss_tts_b200 import _lib as L
    from moss_tts_b200.streaming import AudioStreamDecoder
    rng = np.random.default_rng(0)
    tokens = rng.integers(0, 1024, size=(23, 4))
    fake = _FakeCodec()
    dec = AudioStreamDecoder(fake, chunk_frames=5, overlap_frames=0, decode_kwargs={"chunk_duration": -1, "num_quantizers": 4})
    got, pos = [], 0
    for n in (2, 9, 1, 11):
        dec.push_tokens(tokens[pos: pos + n])
        pos += n
        got += [w.numpy() for w in dec.audio_chunks()]
    got.append(dec.flush().numpy())
    assert [g.shape[0] for g in got] == [5, 5, 5, 5, 3]
    np.testing.assert_array_equal(np.concatenate(got), tokens[:, 0].astype(np.float32))
    assert all(c[1] is None and c[2] == {"num_quantizers": 4} for c in fake.calls)      # chunk_duration <= 0 -> chunking off
    assert dec.flush() is None
    dec2 = AudioStreamDecoder(_FakeCodec(), chunk_frames=4, overlap_frames=2)
    dec2.push_tokens(tokens[:8])
    it = dec2.audio_chunks()
    assert next(it).shape[0] == 4                      # first chunk: nothing to fade against
    with pytest.raises(L.MossTTSNativeError):
        next(it)                                       # CPU waveforms: there is no CPU crossfade


def test_stream_planner_against_the_oracle_dedelay():
    """Host logic of generate_stream on CPU: feed the planner the text channel of a synthetic generation slice by slice;
    the frames it schedules, read off the delay diagonal, must be the oracle's de-delay of every row's audio span, each
    frame exactly once and only after its last code exists."""
    from moss_tts_b200.streaming import plan_push, scan_text_tokens
    from oracle.codec import apply_de_delay_pattern
    rng = np.random.default_rng(3)
    GEN, DELAY, END, IM_END, PAD, nq = 7, 8, 9, 10, 0, 4
    B, steps = 3, 40
    n_frames = [17, 5, 30]                       # row 2 is cut by max_new_tokens before its span ends
    lead = [1, 3, 0]                             # text tokens before the first gen slot
    gen = np.full((B, steps, 1 + nq), 1024, dtype=np.int64)
    for b in range(B):
        seq = [PAD + 1] * lead[b] + [GEN] * n_frames[b] + [DELAY] * (nq - 1) + [END, IM_END]
        seq = (seq + [PAD] * steps)[:steps]
        gen[b, :, 0] = seq
        for f in range(n_frames[b]):
            for q in range(nq):
                if lead[b] + f + q < steps:
                    gen[b, lead[b] + f + q, 1 + q] = rng.integers(0, 1024)
    frame_rows, finished = [[] for _ in range(B)], [False] * B
    scanned = pushed = 0
    got = [[] for _ in range(B)]
    pushes = []
    for upto in [1, 6, 11, 19, 23, 31, 40]:      # uneven slices
        scan_text_tokens(gen[:, scanned:upto, 0], scanned, frame_rows, finished, GEN, IM_END)
        scanned = upto
        if upto == steps:
            finished = [True] * B
        while True:
            plan = plan_push(frame_rows, finished, scanned, pushed, nq, chunk_frames=3, max_push=6)
            if plan is None:
                break
            T, counts = plan
            assert 0 < T <= 6 and (all(finished) or T >= 3)
            for b, n in enumerate(counts):
                for r in frame_rows[b][pushed:pushed + n]:
                    assert r + nq - 1 < scanned                      # every code of the frame exists
                    got[b].append([gen[b, r + q, 1 + q] for q in range(nq)])
            pushes.append(T)
            pushed += T
    for b in range(B):
        span = gen[b, lead[b]:, 1:]
        n_real = min(n_frames[b], steps - lead[b] - (nq - 1))        # frames whose last code was generated
        ref = apply_de_delay_pattern(span[: n_real + nq - 1])
        np.testing.assert_array_equal(np.asarray(got[b], dtype=np.int64).reshape(-1, nq), ref)
    assert finished == [True, True, True] and sum(pushes) == max(len(g) for g in got)
    assert plan_push(frame_rows, finished, scanned, pushed, nq, 3, 6) is None
