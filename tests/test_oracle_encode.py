"""Pins the oracle's encode-side restatement (encoder stack, ResidualLFQ quantise, chunked encode) against
outputs of the reference itself (tests/golden/codec_enc_mini.npz, minted by oracle/make_golden_enc.py)."""
import json
import os

import numpy as np
import torch

from conftest import GOLDEN
from moss_tts_b200.config import CodecConfig
from oracle.codec import OracleCodec
from oracle.weights import make_codec_weights


def _load():
    g = np.load(os.path.join(GOLDEN, "codec_enc_mini.npz"), allow_pickle=False)
    cfg = CodecConfig.from_dict(json.loads(str(g["cfg"])))
    return g, cfg, OracleCodec(cfg, make_codec_weights(cfg, seed=int(g["seed"])))


def test_encode_frame_matches_reference():
    g, cfg, oc = _load()
    wav, lens = torch.from_numpy(g["wav"]), torch.from_numpy(g["wav_lengths"])
    codes, cl, hidden = oc.encode_frame(wav, lens)
    assert np.array_equal(cl.numpy(), g["codes_lengths"])
    assert np.abs(hidden.numpy() - g["hidden"]).max() <= 2e-5 * max(1.0, np.abs(g["hidden"]).max())
    assert np.array_equal(codes.numpy(), g["codes"])            # integer work: bit-exact


def test_quantizer_alone_matches_reference():
    g, cfg, oc = _load()
    codes = oc.quantize(torch.from_numpy(g["hidden"]), torch.from_numpy(g["codes_lengths"]), 5)
    assert np.array_equal(codes.numpy(), g["codes_nq5"])
    # quantized_out of the reference == decode_codes(codes) on the valid frames (the masked tail stays zero + bias)
    emb = oc.decode_codes(codes).numpy()
    for b, n in enumerate(g["codes_lengths"]):
        assert np.abs(emb[b, :, :n] - g["zq_nq5"][b, :, :n]).max() <= 1e-4


def test_encode_public_surface_matches_reference():
    g, cfg, oc = _load()
    wav, lens = torch.from_numpy(g["wav"]), torch.from_numpy(g["wav_lengths"])
    codes, cl, _ = oc.encode_frame(wav, lens, nq=3)
    assert np.array_equal(codes.numpy(), g["enc_nq3_codes"]) and np.array_equal(cl.numpy(), g["enc_nq3_lengths"])
    # batch_encode :1403-1429: zero-pad to the longest, lengths = samples
    D = cfg.downsample_rate
    a, b = wav[0, 0, : 9 * D + 100], wav[1, 0, : 20 * D]
    x = torch.zeros(2, 1, 20 * D)
    x[0, 0, : a.numel()] = a
    x[1, 0, : b.numel()] = b
    codes, cl, _ = oc.encode_frame(x, torch.tensor([a.numel(), b.numel()]))
    assert np.array_equal(codes.numpy(), g["batch_codes"]) and np.array_equal(cl.numpy(), g["batch_lengths"])


def test_chunked_encode_closed_form_matches_reference_streaming():
    g, cfg, oc = _load()
    wav = torch.from_numpy(g["wav"])[:1]
    cf = int(round(0.4 * cfg.sampling_rate)) // cfg.downsample_rate
    codes, cl, hidden = oc.encode_frame(wav, None, chunk_frames=cf)
    n = int(g["chunk_lengths"][0])                       # the reference drops the ragged tail frame of the last chunk
    assert n == wav.shape[-1] // cfg.downsample_rate
    assert np.abs(hidden.numpy()[:, :, :n] - g["chunk_hidden"]).max() <= 2e-5 * max(1.0, np.abs(g["chunk_hidden"]).max())
    assert np.array_equal(codes.numpy()[:, :, :n], g["chunk_codes"])
    assert not np.array_equal(g["chunk_codes"], g["codes"][:, :1, :n])      # and it genuinely differs from unchunked


def test_streaming_encode_pushes_closed_form_matches_reference_streaming():
    """`with model.streaming(B): model._encode_frame(chunk)` over uneven pushes (3, 10, 1, 7, 12, 2, 5 frames; the
    12-frame rings wrap several times): the oracle's per-push closed form (a query sees keys >= end-of-its-push - context)
    against the reference's own streaming run — hidden states <= 2e-5, codes equal except where the reference's own
    top-2 code distances are within 1e-5 (argmax near-ties), and genuinely different from the one-pass encode."""
    g = np.load(os.path.join(GOLDEN, "codec_enc_stream_mini.npz"), allow_pickle=False)
    cfg = CodecConfig.from_dict(json.loads(str(g["cfg"])))
    o = OracleCodec(cfg, make_codec_weights(cfg, seed=int(g["seed"])))
    pushes = [int(n) for n in g["pushes"]]
    codes, lens, hidden, margins = o.encode_frame(torch.from_numpy(g["wav"]), chunk_frames=pushes, return_margins=True)
    assert np.abs(hidden.numpy() - g["hidden_stream"]).max() <= 2e-5
    diff = codes.numpy() != g["codes_stream"]
    assert diff.sum() <= 2 and (margins.numpy()[diff] <= 1e-5).all(), (int(diff.sum()), margins.numpy()[diff])
    assert (g["codes_stream"] != g["codes_full"]).mean() > 0.3
