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
    """Host logic of generate_stream on CPU: feed the planner a synthetic generation slice by slice, laid out the way the
    reference's loop lays it out (SURVEY.md 8c KAT: k gen-slot rows, then n_vq delay-slot rows of which the FIRST still
    carries every channel, then audio_end), with a prompt that already holds a few frames.  The frames it schedules must be
    the ones processor.decode finds -- the non-pad rows of the oracle's de-delay of the returned rows -- each exactly once,
    in order, and only after its last code exists."""
    from moss_tts_b200.streaming import plan_push, scan_finished, scan_frames
    from oracle.codec import apply_de_delay_pattern, split_audio_segments
    rng = np.random.default_rng(3)
    GEN, DELAY, END, IM_END, TXT, P, nq = 7, 8, 9, 10, 1, 1024, 4
    B, S0, steps = 3, 9, 40
    start = [2, 0, 5]                            # first returned row (im_start + 3)
    prompt_frames = [3, 0, 0]                    # row 0: the prompt ends inside an audio span (continuation)
    n_gen = [16, 5, 40]                          # gen-slot rows generated; row 2 is cut by max_new_tokens before its span ends
    lead = [0, 3, 0]                             # free text rows before the generated span
    ids = np.full((B, S0 + steps, 1 + nq), P, dtype=np.int64)
    ids[:, :, 0] = TXT

    def put_span(b, row0, k, delay_rows):        # k gen-slot rows from row0, then delay-slot rows; codes on the diagonals
        n_fr = k + (1 if delay_rows else 0)      # the first delay-slot row is a frame as well
        for f in range(n_fr):
            for q in range(nq):
                if row0 + f + q < ids.shape[1]:
                    ids[b, row0 + f + q, 1 + q] = rng.integers(0, 1024)
        for r in range(row0, min(row0 + k, ids.shape[1])):
            ids[b, r, 0] = GEN
        for r in range(row0 + k, min(row0 + k + delay_rows, ids.shape[1])):
            ids[b, r, 0] = DELAY
        return row0 + k + delay_rows

    for b in range(B):
        r = S0 - prompt_frames[b] if prompt_frames[b] else S0 + lead[b]
        r = put_span(b, r, prompt_frames[b] + n_gen[b], nq)
        for tok in (END, IM_END):
            if r < ids.shape[1]:
                ids[b, r, 0] = tok
                r += 1
    codes = ids[:, :, 1:].astype(np.int16)
    frame_rows, finished, next_row = [[] for _ in range(B)], [False] * B, list(start)
    scanned = pushed = 0
    got = [[] for _ in range(B)]
    pushes = []
    for upto in [1, 6, 11, 19, 23, 31, 40]:      # uneven slices
        scan_finished(ids[:, S0 + scanned:S0 + upto, 0], finished, IM_END)
        scanned = upto
        scan_frames(codes, next_row, S0 + scanned, frame_rows, P)
        if upto == steps:
            finished = [True] * B
        while True:
            plan = plan_push(frame_rows, finished, pushed, chunk_frames=3, max_push=6)
            if plan is None:
                break
            T, counts = plan
            assert 0 < T <= 6 and (all(finished) or T >= 3)
            for b, n in enumerate(counts):
                for r in frame_rows[b][pushed:pushed + n]:
                    assert r + nq - 1 < S0 + scanned                 # every code of the frame exists
                    got[b].append([ids[b, r + q, 1 + q] for q in range(nq)])
            pushes.append(T)
            pushed += T
    for b in range(B):
        segs = split_audio_segments(apply_de_delay_pattern(ids[b, start[b]:, 1:]), P)
        ref = np.concatenate(segs, 0)
        np.testing.assert_array_equal(np.asarray(got[b], dtype=np.int64).reshape(-1, nq), ref)
    assert len(got[0]) == 3 + 16 + 1 and len(got[1]) == 5 + 1            # prompt frames + gen-slot rows + the first delay-slot row
    assert finished == [True, True, True] and sum(pushes) == max(len(g) for g in got)
    assert plan_push(frame_rows, finished, pushed, 3, 6) is None
