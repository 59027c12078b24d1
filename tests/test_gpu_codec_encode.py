"""Codec encode side (patch downsample, encoder transformer stages, residual LFQ quantise) through
the C ABI vs the reference-executed golden (tests/golden/codec_enc_mini.npz) and the oracle."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from moss_tts_b200.config import CodecConfig
from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
from oracle.codec import OracleCodec
from oracle.weights import make_codec_weights
from gpu_util import DEV, snr_db

pytestmark = pytest.mark.gpu
SNR_DB = 40.0            # same gate as the decoder's waveform: encoder_hidden_states SNR vs the reference (TF32 GEMMs)
TIE = 1e-5               # a code may differ from the reference only where the reference's own top-2 distance margin <= TIE


@pytest.fixture(scope="module")
def enc():
    g = np.load(os.path.join(GOLDEN, "codec_enc_mini.npz"), allow_pickle=False)
    cfg = CodecConfig.from_dict(json.loads(str(g["cfg"])))
    w = make_codec_weights(cfg, seed=int(g["seed"]))
    return g, cfg, MossAudioTokenizerModel(cfg, w, device=DEV), OracleCodec(cfg, w)


def _assert_codes(got, ref, margins, lengths, what):
    """Bit-exact, or the first differing quantiser of a frame is a near-tie of the reference arithmetic (later
    quantisers of that frame then see a different residual and are not comparable)."""
    nq, B, T = ref.shape
    bad = 0
    for b in range(B):
        for t in range(int(lengths[b])):
            d = np.nonzero(got[:, b, t] != ref[:, b, t])[0]
            if len(d):
                bad += 1
                assert margins[d[0], b, t] <= TIE, (what, b, t, int(d[0]), float(margins[d[0], b, t]))
    print(f"[{what}] frames with a near-tie divergence: {bad} / {int(np.sum(lengths))}")


def test_quantizer_on_reference_hidden_states(enc):
    g, cfg, model, oracle = enc
    hidden = torch.from_numpy(g["hidden"])
    lens = torch.from_numpy(g["codes_lengths"])
    _, margins = oracle.quantize(hidden, lens, return_margins=True)
    _, codes, _ = model.quantizer(hidden.to(DEV), lens.to(DEV))
    _assert_codes(codes.cpu().numpy(), g["codes"], margins.numpy(), g["codes_lengths"], "quantiser, all nq")
    _, codes5, _ = model.quantizer(hidden.to(DEV), lens.to(DEV), 5)
    _assert_codes(codes5.cpu().numpy(), g["codes_nq5"], margins.numpy(), g["codes_lengths"], "quantiser, nq=5")
    # frames past the length quantise the bias alone (mask = 0): identical codes for every masked frame
    b = 1
    n = int(g["codes_lengths"][b])
    assert np.array_equal(codes.cpu().numpy()[:, b, n:], g["codes"][:, b, n:])


def test_encode_frame_matches_reference(enc):
    g, cfg, model, oracle = enc
    wav, lens = torch.from_numpy(g["wav"]), torch.from_numpy(g["wav_lengths"])
    out = model._encode_frame(wav.to(DEV), lens.to(DEV))
    assert np.array_equal(out.audio_codes_lengths.cpu().numpy(), g["codes_lengths"])
    hidden = out.encoder_hidden_states.cpu()
    assert hidden.shape == g["hidden"].shape
    s = snr_db(torch.from_numpy(g["hidden"]), hidden)
    print(f"[encoder hidden states] SNR {s:.1f} dB")
    assert s >= SNR_DB
    # the quantiser on OUR hidden states must agree with the oracle's quantiser on the same states
    codes_o, margins = oracle.quantize(hidden.contiguous(), out.audio_codes_lengths.cpu(), return_margins=True)
    _assert_codes(out.audio_codes.cpu().numpy(), codes_o.numpy(), margins.numpy(), g["codes_lengths"], "encode, own hidden")
    agree = float((out.audio_codes.cpu().numpy()[0] == g["codes"][0]).mean())
    print(f"[encode] first-quantiser agreement with the fp32 reference codes: {agree:.3f}")
    assert agree >= 0.8


def test_encode_public_surface(enc):
    g, cfg, model, oracle = enc
    wav = torch.from_numpy(g["wav"]).to(DEV)
    T = wav.shape[-1]
    pm = torch.zeros(2, T, dtype=torch.bool, device=DEV)
    pm[0] = True
    pm[1, : int(g["wav_lengths"][1])] = True
    codes, lens = model.encode(wav, padding_mask=pm, num_quantizers=3, return_dict=False)
    assert codes.shape == g["enc_nq3_codes"].shape and np.array_equal(lens.cpu().numpy(), g["enc_nq3_lengths"])
    D = cfg.downsample_rate
    be = model.batch_encode([wav[0, 0, : 9 * D + 100], wav[1, 0, : 20 * D]])
    assert be.audio_codes.shape == g["batch_codes"].shape
    assert np.array_equal(be.audio_codes_lengths.cpu().numpy(), g["batch_lengths"])
    assert float((be.audio_codes.cpu().numpy()[0] == g["batch_codes"][0]).mean()) >= 0.8
    with pytest.raises(ValueError):
        model.encode(wav, chunk_duration=0.4)                 # B != 1
    with pytest.raises(ValueError):
        model.encode(wav[:1], chunk_duration=0.41)            # not a whole number of frames
    with pytest.raises(ValueError):
        model.encode(wav[:1], chunk_duration=5.0)             # > context duration
    with pytest.raises(ValueError):
        model.batch_encode([])


def test_chunked_encode_matches_reference_streaming(enc):
    g, cfg, model, oracle = enc
    wav = torch.from_numpy(g["wav"])[:1].to(DEV)
    out = model.encode(wav, chunk_duration=0.4)
    assert np.array_equal(out.audio_codes_lengths.cpu().numpy(), g["chunk_lengths"])
    assert out.audio_codes.shape == g["chunk_codes"].shape
    s = snr_db(torch.from_numpy(g["chunk_hidden"]), out.encoder_hidden_states.cpu())
    print(f"[chunked encoder hidden states] SNR {s:.1f} dB")
    assert s >= SNR_DB
    plain = model.encode(wav).encoder_hidden_states.cpu()[:, :, : g["chunk_hidden"].shape[-1]]
    assert snr_db(torch.from_numpy(g["chunk_hidden"]), plain) < 30.0      # the ring rule genuinely matters here


@pytest.mark.parametrize("B,T", [(1, 1920), (3, 1920 * 7 + 1), (5, 1920 * 30)])
def test_encode_sweep_vs_oracle(enc, B, T):
    g, cfg, model, oracle = enc
    wav = torch.randn(B, 1, T, generator=torch.Generator().manual_seed(T)) * 0.3
    out = model._encode_frame(wav.to(DEV))
    codes_o, lens_o, hidden_o = oracle.encode_frame(wav)
    assert np.array_equal(out.audio_codes_lengths.cpu().numpy(), lens_o.numpy())
    assert snr_db(hidden_o, out.encoder_hidden_states.cpu()) >= SNR_DB


def test_encode_decode_round_trip_is_idempotent(enc):
    """Size-independent property: decode(codes) of the folded-table path equals the quantiser's own reconstruction, so
    re-quantising the dequantised embedding with ONE quantiser returns the same first code."""
    g, cfg, model, oracle = enc
    codes = torch.from_numpy(g["codes"]).to(DEV)
    wav = model.decode(codes).audio
    assert snr_db(torch.from_numpy(g["roundtrip_wav"]), wav.cpu()) >= SNR_DB


def test_released_encoder_shapes_one_layer():
    """Released encoder widths (240 -> 768 -> 384 | 768 | 1280 -> 768, windows 1000..125), 1 layer per stage, nq = 32."""
    cfg = CodecConfig.released()
    for st in cfg.stages + cfg.enc_stages:
        st.num_layers = 1
    w = make_codec_weights(cfg, seed=9)
    model = MossAudioTokenizerModel(cfg, w, device=DEV)
    oracle = OracleCodec(cfg, w)
    wav = torch.randn(2, 1, 1920 * 40 + 333, generator=torch.Generator().manual_seed(2)) * 0.3
    out = model._encode_frame(wav.to(DEV))
    codes_o, lens_o, hidden_o = oracle.encode_frame(wav)
    s = snr_db(hidden_o, out.encoder_hidden_states.cpu())
    print(f"[encoder released-width, 1 layer/stage] hidden SNR {s:.1f} dB")
    assert s >= SNR_DB and np.array_equal(out.audio_codes_lengths.cpu().numpy(), lens_o.numpy())
    co, margins = oracle.quantize(out.encoder_hidden_states.cpu().contiguous(), lens_o, return_margins=True)
    _assert_codes(out.audio_codes.cpu().numpy(), co.numpy(), margins.numpy(), lens_o.numpy(), "released-width encode")


def test_streaming_encode_pushes_match_reference_streaming():
    """`with model.streaming(B): model._encode_frame(push)` over uneven pushes (3, 10, 1, 7, 12, 2, 5 frames; the
    12-token rings wrap) through mt_codec_stream_push_encode, against the reference's own streaming run
    (tests/golden/codec_enc_stream_mini.npz): hidden states >= 40 dB, our codes = the oracle's quantiser on our hidden
    states up to near-ties, and the stream genuinely differs from the one-pass encode."""
    g = np.load(os.path.join(GOLDEN, "codec_enc_stream_mini.npz"), allow_pickle=False)
    cfg = CodecConfig.from_dict(json.loads(str(g["cfg"])))
    w = make_codec_weights(cfg, seed=int(g["seed"]))
    model, oracle = MossAudioTokenizerModel(cfg, w, device=DEV), OracleCodec(cfg, w)
    wav = torch.from_numpy(g["wav"]).to(DEV)
    B, D = wav.shape[0], cfg.downsample_rate
    hs, cs, f = [], [], 0
    with model.streaming(B):
        for n in [int(x) for x in g["pushes"]]:
            out = model._encode_frame(wav[..., f * D:(f + n) * D])
            assert out.audio_codes.shape == (cfg.num_quantizers, B, n)
            hs.append(out.encoder_hidden_states.cpu()); cs.append(out.audio_codes.cpu()); f += n
        with pytest.raises(ValueError):
            model._encode_frame(wav[:1, ..., :D])                                   # batch size fixed by streaming()
        with pytest.raises(Exception, match="ring"):
            model._encode_frame(wav[..., : 13 * D])                                 # 13 frames > the 12-token ring
    hidden, codes = torch.cat(hs, dim=-1), torch.cat(cs, dim=-1)
    s = snr_db(torch.from_numpy(g["hidden_stream"]), hidden)
    one = model._encode_frame(wav)                                                  # the rings are gone: one-pass again
    s_one = snr_db(torch.from_numpy(g["hidden_stream"]), one.encoder_hidden_states.cpu())
    print(f"[streaming encode] hidden SNR vs reference streaming {s:.1f} dB (one-pass encode vs the same: {s_one:.1f} dB)")
    assert s >= SNR_DB and s_one < 30.0
    lens = torch.full((B,), f, dtype=torch.long)
    codes_o, margins = oracle.quantize(hidden.contiguous(), lens, return_margins=True)
    _assert_codes(codes.numpy(), codes_o.numpy(), margins.numpy(), lens.numpy(), "streaming encode, own hidden")
    agree = float((codes.numpy()[0] == g["codes_stream"][0]).mean())
    print(f"[streaming encode] first-quantiser agreement with the reference's streamed codes: {agree:.3f}")
    assert agree >= 0.8
    # encode and decode streams live side by side under one streaming() (reference: every StreamingModule starts)
    with model.streaming(B):
        e = model._encode_frame(wav[..., : 4 * D])
        d = model._decode_frame(e.audio_codes)
        assert d.audio.shape == (B, 1, 4 * D) and torch.isfinite(d.audio).all()


def test_stream_push_entry_points_refuse_the_wrong_half(enc):
    """mt_codec_stream_push on an encoder stream / mt_codec_stream_push_encode on a decoder stream fail loudly (C ABI)."""
    import ctypes as C
    from moss_tts_b200 import _lib as L
    g, cfg, model, oracle = enc
    B, D = 1, cfg.downsample_rate
    wav = torch.zeros(B, 2 * D, device=DEV)
    codes = torch.zeros(cfg.num_quantizers, B, 2, dtype=torch.int64, device=DEV)
    hidden = torch.empty(B, 2, cfg.code_dim, device=DEV)
    out_wav = torch.empty(B, 2 * D, device=DEV)
    ws = torch.empty(64 << 20, dtype=torch.uint8, device=DEV)
    with model.streaming(B):
        rc = L.lib().mt_codec_stream_push(model._enc_stream, L.ptr(codes), cfg.num_quantizers, 2, L.ptr(out_wav), L.ptr(ws),
                                          C.c_longlong(ws.numel()), L.stream_ptr(DEV))
        assert rc != 0 and b"encoder" in L.lib().mt_last_error()
        rc = L.lib().mt_codec_stream_push_encode(model._stream, L.ptr(wav), 2, None, cfg.num_quantizers, L.ptr(codes), L.ptr(hidden),
                                                 L.ptr(ws), C.c_longlong(ws.numel()), L.stream_ptr(DEV))
        assert rc != 0 and b"decoder" in L.lib().mt_last_error()
        rc = L.lib().mt_codec_stream_push_encode(model._enc_stream, L.ptr(wav), 0, None, cfg.num_quantizers, L.ptr(codes), L.ptr(hidden),
                                                 L.ptr(ws), C.c_longlong(ws.numel()), L.stream_ptr(DEV))
        assert rc != 0 and b"empty" in L.lib().mt_last_error()
        out = model._encode_frame(wav.unsqueeze(1))          # the stream is still usable after the refused calls
        assert out.audio_codes.shape == (cfg.num_quantizers, B, 2)
