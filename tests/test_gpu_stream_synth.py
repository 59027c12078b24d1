"""Streaming synthesis: mt_lm_stream_* + mt_stream_gather_frames + mt_codec_stream_push behind
moss_tts_b200.streaming.generate_stream, and the reference-shaped AudioStreamDecoder (mt_crossfade).  Parity:
* the reference AudioStreamDecoder's own chunks (tests/golden/stream_decoder_mini.npz, reference-executed),
* generate_stream ids == generate ids (bit-exact, same seed), streamed audio == the codec's streaming decode of the
  finished, de-delayed codes with the same pushes (bit-exact) and the oracle's closed form (SNR)."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from moss_tts_b200 import _lib as L
from moss_tts_b200.config import CodecConfig, LMConfig
from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
from moss_tts_b200.modeling_moss_tts import MossTTSDelayModel
from moss_tts_b200.streaming import AudioStreamDecoder, generate_stream
from oracle.codec import OracleCodec, apply_de_delay_pattern, split_audio_segments
from oracle.weights import make_codec_weights, make_lm_weights, make_prompt
from gpu_util import DEV, snr_db

pytestmark = pytest.mark.gpu
SNR_DB = 40.0


@pytest.fixture(scope="module")
def codec():
    cfg = CodecConfig.mini()
    w = make_codec_weights(cfg, seed=11)
    return cfg, MossAudioTokenizerModel(cfg, w, device=DEV), OracleCodec(cfg, w)


def test_crossfade_kernel_known_answer():
    g = np.load(os.path.join(GOLDEN, "stream_decoder_mini.npz"), allow_pickle=False)
    pt, hd = torch.from_numpy(g["xf_prev"]).to(DEV), torch.from_numpy(g["xf_head"]).to(DEV)
    out = torch.empty_like(pt)
    L.check(L.lib().mt_crossfade(L.ptr(pt), L.ptr(hd), pt.numel(), L.ptr(out), L.stream_ptr(DEV)), "mt_crossfade")
    np.testing.assert_allclose(out.cpu().numpy(), g["xf_out"], rtol=0, atol=1e-6)
    one = torch.empty(1, device=DEV)
    L.check(L.lib().mt_crossfade(L.ptr(pt), L.ptr(hd), 1, L.ptr(one), L.stream_ptr(DEV)), "mt_crossfade")
    assert one.item() == pt[0].item()                      # linspace(1, 0, 1) == [1]


def test_audio_stream_decoder_matches_reference_chunks(codec):
    g = np.load(os.path.join(GOLDEN, "stream_decoder_mini.npz"), allow_pickle=False)
    cfg, model, _ = codec
    dec = AudioStreamDecoder(model, chunk_frames=int(g["chunk_frames"]), overlap_frames=int(g["overlap_frames"]))
    chunks, after, pos = [], [], 0
    for n in g["pushes"]:
        dec.push_tokens(g["tokens"][pos: pos + int(n)])
        pos += int(n)
        chunks += list(dec.audio_chunks())
        after.append(len(chunks))
    last = dec.flush()
    if last is not None:
        chunks.append(last)
    assert after == g["chunks_after_push"].tolist() and len(chunks) == int(g["n_chunks"])
    for i, ch in enumerate(chunks):
        ref = torch.from_numpy(g[f"chunk{i}"])
        assert tuple(ch.shape) == tuple(ref.shape)
        s = snr_db(ref, ch.cpu())
        print(f"[AudioStreamDecoder chunk {i}] SNR vs the reference's chunk {s:.1f} dB")
        assert s >= SNR_DB
    assert dec.flush() is None
    with pytest.raises(ValueError):
        dec.push_tokens(torch.zeros(3, dtype=torch.long))


def _stream_case(codec, never_stop, B, steps, chunk_frames, sample, left_pad=None, lm_slice=None, text_greedy=False,
                 first_chunk_frames=None):
    ccfg, cmodel, oracle = codec
    cfg = LMConfig.tiny(n_vq=ccfg.num_quantizers)
    w = make_lm_weights(cfg, seed=5, never_stop=never_stop)
    lm = MossTTSDelayModel(cfg, w, device=DEV)
    ids, mask = make_prompt(cfg, B, 12, seed=3, left_pad=left_pad)
    kw = dict(max_new_tokens=steps, audio_repetition_penalty=1.1 if sample else 1.0, seed=7)
    if not sample:
        kw.update(text_temperature=0, audio_temperature=0)
    elif text_greedy:
        kw.update(text_temperature=0)                       # sampled audio codes under a never-stopping text channel
    ref = lm.generate(ids, mask, **kw)
    chunks = list(generate_stream(lm, cmodel, ids, mask, chunk_frames=chunk_frames, lm_slice=lm_slice,
                                  first_chunk_frames=first_chunk_frames, **kw))
    out = lm.last_stream_output
    for (a, x), (b_, y) in zip(ref, out):                   # the sliced loop is the same loop
        assert a == b_ and torch.equal(x, y)
    # frames of every row from the finished ids, the way processor.decode finds them (processing_moss_tts.py:668-685):
    # de-delay the returned ids, every row that is not all pad is a frame, runs of them are the audio segments
    nq, ds = cfg.n_vq, cmodel.downsample_rate
    frames = []
    for sl, x in out:
        segs = split_audio_segments(apply_de_delay_pattern(x.cpu().numpy()[:, 1:]), cfg.audio_pad_code) if x.shape[0] >= nq else []
        frames.append(np.concatenate(segs, 0) if segs else np.zeros((0, nq), np.int64))
    pushes = [w_.shape[1] // ds for w_, _ in chunks]
    total = [sum(v[b] for _, v in chunks) // ds for b in range(B)]
    assert total == [len(f) for f in frames]
    if not pushes:
        return chunks, frames
    Tmax = sum(pushes)
    codes = torch.zeros(nq, B, Tmax, dtype=torch.int64)
    for b in range(B):
        codes[:, b, :len(frames[b])] = torch.from_numpy(frames[b].T)
    outs, pos = [], 0
    with cmodel.streaming(batch_size=B):
        for n in pushes:
            outs.append(cmodel._decode_frame(codes[:, :, pos: pos + n].to(DEV)).audio[:, 0])
            pos += n
    direct = torch.cat(outs, -1)
    streamed = torch.cat([w_ for w_, _ in chunks], -1)
    orc = oracle.decode_frames(codes, chunk_frames=pushes)[:, 0]
    for b in range(B):
        n = len(frames[b]) * ds
        assert torch.equal(streamed[b, :n], direct[b, :n])
        if n:
            assert snr_db(orc[b, :n], streamed[b, :n].cpu()) >= SNR_DB
    # valid_samples are prefixes that only shrink once a row has finished
    for b in range(B):
        v = [val[b] for _, val in chunks]
        full = [w_.shape[1] for w_, _ in chunks]
        short = [i for i in range(len(v)) if v[i] < full[i]]
        assert all(v[i] == 0 for i in range(short[0] + 1, len(v))) if short else True
    return chunks, frames


@pytest.mark.parametrize("sample", [False, True])
def test_generate_stream_never_stop(codec, sample):
    chunks, frames = _stream_case(codec, True, 3, 45, 4, sample, left_pad=[0, 2, 5], text_greedy=True)
    assert len(chunks) >= 5 and all(len(f) in (37, 38) for f in frames)
    assert all(w.shape[1] // 1920 >= 4 for w, _ in chunks[:-1])


def test_generate_stream_slices_and_single_row(codec):
    _stream_case(codec, True, 1, 30, 3, False, lm_slice=5)
    _stream_case(codec, True, 2, 9, 4, False)               # too short for one complete frame... or exactly one
    _stream_case(codec, True, 2, 1, 4, False)               # max_new_tokens = 1: time_step 0 only


def test_generate_stream_small_first_chunk(codec):
    chunks, frames = _stream_case(codec, True, 2, 45, 6, False, first_chunk_frames=2)
    sizes = [w.shape[1] // 1920 for w, _ in chunks]
    assert sizes[0] == 2 and all(s >= 6 for s in sizes[1:-1]) and sum(sizes) == len(frames[0])


def test_generate_stream_rows_that_stop(codec):
    chunks, frames = _stream_case(codec, False, 4, 60, 2, True, left_pad=[0, 1, 0, 3])
    print("[generate_stream, stopping rows] frames per row", [len(f) for f in frames], "chunks", len(chunks))


def test_generate_stream_total_samples_equal_processor_decode(codec):
    """Rows whose audio span ends naturally (ramp-down, audio_end): the frame on the first delay-slot row is streamed too, so
    every row's streamed samples total the samples of the segments processor.decode(generate()) decodes (before its
    start_length waveform trim).  The audio itself is checked inside _stream_case (bit-exact vs the codec's streaming decode
    of those frames with the same pushes, SNR vs the oracle)."""
    ccfg, cmodel, _ = codec
    cfg = LMConfig.tiny(n_vq=ccfg.num_quantizers)
    ds = cmodel.downsample_rate
    natural_ends = 0
    for seed in (5, 6, 9, 12):
        lm = MossTTSDelayModel(cfg, make_lm_weights(cfg, seed=seed, never_stop=False), device=DEV)
        ids, mask = make_prompt(cfg, 3, 12, seed=seed, left_pad=[0, 2, 1])
        kw = dict(max_new_tokens=70, audio_repetition_penalty=1.0, seed=seed, text_temperature=1.0, text_top_k=50)
        chunks = list(generate_stream(lm, cmodel, ids, mask, chunk_frames=3, **kw))
        for b, (sl, x) in enumerate(lm.last_stream_output):
            xs = x.cpu().numpy()
            segs = split_audio_segments(apply_de_delay_pattern(xs[:, 1:]), cfg.audio_pad_code) if xs.shape[0] >= cfg.n_vq else []
            want = sum(s_.shape[0] for s_ in segs) * ds
            got = sum(v[b] for _, v in chunks)
            assert got == want, (seed, b, got, want)
            ends = np.nonzero(xs[:, 0] == cfg.audio_end_token_id)[0]
            natural_ends += len(ends)
            for e in ends:                   # the n_vq rows before audio_end are delay slots; the first of them is a frame
                assert e >= cfg.n_vq and (xs[e - cfg.n_vq: e, 0] == cfg.audio_assistant_delay_slot_token_id).all()
                assert (xs[e - cfg.n_vq, 1:] != cfg.audio_pad_code).any()
    print(f"[generate_stream] audio spans that ended naturally: {natural_ends}")
    assert natural_ends > 0, "no row ended its audio span naturally: the case under test did not occur"


def test_open_session_owns_the_engine(codec):
    ccfg, cmodel, _ = codec
    cfg = LMConfig.tiny(n_vq=ccfg.num_quantizers)
    lm = MossTTSDelayModel(cfg, make_lm_weights(cfg, seed=5, never_stop=True), device=DEV)
    ids, mask = make_prompt(cfg, 2, 12, seed=3)
    kw = dict(max_new_tokens=30, text_temperature=0, audio_temperature=0)
    it = generate_stream(lm, cmodel, ids, mask, chunk_frames=2, **kw)
    first = next(it)                                         # the session is open and has steps queued
    with pytest.raises(ValueError):
        lm.generate(ids, mask, **kw)
    with pytest.raises(ValueError):
        lm.forward(ids, mask)
    with pytest.raises(RuntimeError):
        with cmodel.streaming(2):                            # the codec is streaming for the session, too
            pass
    rest = list(it)
    assert first[0].shape[0] == 2 and len(rest) >= 1
    ref = lm.generate(ids, mask, **kw)                       # closed: the handle is free again and unchanged
    for (a, x), (b_, y) in zip(ref, lm.last_stream_output):
        assert a == b_ and torch.equal(x, y)
    it2 = generate_stream(lm, cmodel, ids, mask, chunk_frames=2, **kw)
    next(it2)
    it2.close()                                              # abandoning a session releases engine and codec stream
    assert torch.equal(lm.generate(ids, mask, **kw)[0][1], ref[0][1])
    with cmodel.streaming(2):
        pass
