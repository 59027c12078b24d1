"""Codec decode side (RVQ dequantise, transformer decoder, patch upsample, processor glue) vs the oracle."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from moss_tts_b200.config import CodecConfig, CodecStage, LMConfig
from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
from moss_tts_b200.processing_moss_tts import MossTTSDelayProcessor
from oracle.codec import OracleCodec, apply_delay_pattern
from oracle.weights import make_codec_weights
from gpu_util import DEV, snr_db

pytestmark = pytest.mark.gpu
SNR_DB = 40.0          # north_star: waveform SNR >= 40 dB vs the reference decoder
RVQ_TOL = 1e-5         # RVQ emb max-abs <= 1e-5


def _cfg(js):
    d = json.loads(js)
    stages = [CodecStage(**s) for s in d.pop("stages")]
    return CodecConfig(stages=stages, **d)


@pytest.fixture(scope="module")
def mini():
    g = np.load(os.path.join(GOLDEN, "codec_mini.npz"), allow_pickle=False)
    cfg = _cfg(str(g["cfg"]))
    w = make_codec_weights(cfg, seed=int(g["seed"]))
    return g, cfg, w, MossAudioTokenizerModel(cfg, w, device=DEV), OracleCodec(cfg, w)


def test_rvq_dequant(mini):
    g, cfg, w, model, oracle = mini
    codes = torch.from_numpy(g["codes"])
    emb = model.quantizer.decode_codes(codes.to(DEV)).cpu().numpy()
    scale = np.abs(g["emb"]).max()
    assert np.abs(emb - g["emb"]).max() <= RVQ_TOL * max(1.0, scale)
    emb3 = model.quantizer.decode_codes(codes[:3].to(DEV)).cpu().numpy()       # variable bitrate prefix
    assert np.abs(emb3 - g["emb_nq3"]).max() <= RVQ_TOL * max(1.0, scale)


@pytest.mark.parametrize("B,T", [(1, 1), (3, 13), (2, 125), (1, 300), (2, 64), (5, 6)])
def test_rvq_dequant_shapes(mini, B, T):
    g, cfg, w, model, oracle = mini
    codes = torch.randint(0, cfg.codebook_size, (cfg.num_quantizers, B, T), generator=torch.Generator().manual_seed(T))
    got = model.quantizer.decode_codes(codes.to(DEV)).cpu()
    ref = oracle.decode_codes(codes)
    assert got.shape == ref.shape
    assert (got - ref).abs().max().item() <= RVQ_TOL * max(1.0, ref.abs().max().item())


def test_decode_tf32_layer_gemms_match_reference_golden(mini):
    """layer_gemm="tf32" (fp32 storage, TF32 math) stays selectable next to the default fp16-operand GEMMs."""
    g, cfg, w, model, oracle = mini
    m32 = MossAudioTokenizerModel(cfg, w, device=DEV, layer_gemm="tf32")
    assert model.layer_gemm == "fp16" and m32.layer_gemm == "tf32"
    codes = torch.from_numpy(g["codes"]).to(DEV)
    s16 = snr_db(torch.from_numpy(g["wav"]), model.decode(codes).audio.cpu())
    s32 = snr_db(torch.from_numpy(g["wav"]), m32.decode(codes).audio.cpu())
    print(f"[codec mini] SNR vs reference: fp16-operand layer GEMMs {s16:.1f} dB, TF32 {s32:.1f} dB")
    assert s16 >= SNR_DB and s32 >= SNR_DB


def test_decode_matches_reference_golden(mini):
    g, cfg, w, model, oracle = mini
    codes = torch.from_numpy(g["codes"]).to(DEV)
    out = model.decode(codes)
    s = snr_db(torch.from_numpy(g["wav"]), out.audio.cpu())
    print(f"[codec mini] unchunked SNR vs reference {s:.1f} dB")
    assert s >= SNR_DB
    np.testing.assert_array_equal(out.audio_lengths.cpu().numpy(), g["wav_lengths"])
    pm = torch.ones(2, 40, dtype=torch.bool); pm[1, 29:] = False
    out5 = model.decode(codes, padding_mask=pm.to(DEV), num_quantizers=5)
    assert snr_db(torch.from_numpy(g["wav_nq5"]), out5.audio.cpu()) >= SNR_DB
    np.testing.assert_array_equal(out5.audio_lengths.cpu().numpy(), g["wav_nq5_lengths"])
    bd = model.batch_decode([codes[:, 0, :17], codes[:, 1, :40]])
    ref = torch.from_numpy(g["wav_batch"])
    assert snr_db(ref[0, :, : 17 * cfg.downsample_rate], bd.audio.cpu()[0, :, : 17 * cfg.downsample_rate]) >= SNR_DB
    assert snr_db(ref[1], bd.audio.cpu()[1]) >= SNR_DB
    np.testing.assert_array_equal(bd.audio_lengths.cpu().numpy(), g["wav_batch_lengths"])


def test_chunked_decode_matches_reference_streaming(mini):
    g, cfg, w, model, oracle = mini
    c1 = torch.from_numpy(g["codes"])[:, :1, :33].to(DEV)
    out = model.decode(c1, chunk_duration=0.8)
    s = snr_db(torch.from_numpy(g["wav_chunked"]), out.audio.cpu())
    print(f"[codec mini] chunked SNR vs reference streaming {s:.1f} dB")
    assert s >= SNR_DB
    assert snr_db(torch.from_numpy(g["wav_unchunked33"]), out.audio.cpu()) < 30.0     # genuinely different semantics
    with pytest.raises(ValueError):
        model.decode(torch.from_numpy(g["codes"]).to(DEV), chunk_duration=0.8)          # B != 1
    with pytest.raises(ValueError):
        model.decode(c1, chunk_duration=2.0)
    with pytest.raises(ValueError):
        model.decode(c1, num_quantizers=99)


@pytest.mark.parametrize("B,T,chunk", [(1, 1, None), (5, 7, None), (2, 61, None), (1, 47, 0.4), (3, 100, None)])
def test_decode_sweep_vs_oracle(mini, B, T, chunk):
    g, cfg, w, model, oracle = mini
    codes = torch.randint(0, cfg.codebook_size, (cfg.num_quantizers, B, T), generator=torch.Generator().manual_seed(B * 100 + T))
    got = model.decode(codes.to(DEV), chunk_duration=chunk).audio.cpu()
    ref, _ = oracle.decode(codes, chunk_duration=chunk)
    assert got.shape == ref.shape
    assert snr_db(ref, got) >= SNR_DB


def test_processor_glue_matches_reference(mini):
    g = mini[0]
    cfg10 = _cfg(str(g["cfg10"]))
    model10 = MossAudioTokenizerModel(cfg10, make_codec_weights(cfg10, seed=5), device=DEV)
    proc = MossTTSDelayProcessor(LMConfig.tiny(n_vq=cfg10.num_quantizers), audio_tokenizer=model10)
    wavs = proc._parse_audio_codes(1, torch.from_numpy(g["parse_in"]).to(DEV))
    assert len(wavs) == int(g["parse_n"])
    for i, wv in enumerate(wavs):
        ref = torch.from_numpy(g[f"parse_wav{i}"])
        assert wv.shape == ref.shape and wv.device.type == "cpu" and wv.dtype == torch.float32
        assert snr_db(ref, wv) >= SNR_DB
    assert proc._parse_audio_codes(0, torch.full((12, cfg10.num_quantizers), 1024, device=DEV)) == []


def test_released_stage_shapes_one_layer():
    """The released widths (1280/768, ff 5120/3072, windows 125..1000) with 1 layer per stage."""
    cfg = CodecConfig.released()
    cfg.enc_stages = []                      # encode side: tests/test_gpu_codec_encode.py
    for st in cfg.stages:
        st.num_layers = 1
    w = make_codec_weights(cfg, seed=9)
    model = MossAudioTokenizerModel(cfg, w, device=DEV)
    oracle = OracleCodec(cfg, w)
    codes = torch.randint(0, 1024, (32, 2, 130), generator=torch.Generator().manual_seed(1))
    got = model.decode(codes.to(DEV)).audio.cpu()
    ref, _ = oracle.decode(codes)
    s = snr_db(ref, got)
    print(f"[codec released-width, 1 layer/stage] SNR {s:.1f} dB")
    assert s >= SNR_DB
    one = model.decode(codes[:, :1].to(DEV), chunk_duration=8).audio.cpu()
    refc, _ = oracle.decode(codes[:, :1], chunk_duration=8)
    assert snr_db(refc, one) >= SNR_DB
    m32 = MossAudioTokenizerModel(cfg, w, device=DEV, layer_gemm="tf32")
    s32 = snr_db(ref, m32.decode(codes.to(DEV)).audio.cpu())
    print(f"[codec released-width, 1 layer/stage] TF32 layer GEMMs: SNR {s32:.1f} dB")
    assert s32 >= SNR_DB


@pytest.mark.parametrize("B,T,chunk", [(1, 9, None), (2, 33, None), (3, 70, None), (2, 131, None), (1, 300, None),
                                       (1, 131, 0.8), (1, 300, 0.48)])
def test_tcgen05_attention_matches_mma_sync_kernel_and_oracle(mini, B, T, chunk):
    """The tcgen05 / TMEM attention (codec_attn_tc5.cu, the product path) against the mma.sync FlashAttention-2 kernel it
    replaces on the same inputs, and against the oracle: one-tile and multi-tile sequences, ragged last tiles, windows
    shorter than the sequence (mini: 12/25/50/100-token windows), chunked decode masks, several samples per launch."""
    from moss_tts_b200 import _lib as L
    g, cfg, w, model, oracle = mini
    gen = torch.Generator().manual_seed(100 + T)
    codes = torch.randint(0, cfg.codebook_size, (cfg.num_quantizers, B, T), generator=gen)
    kw = {} if chunk is None else {"chunk_duration": chunk}
    try:
        L.lib().mt_codec_debug_set_attention(1)
        old = model.decode(codes.to(DEV), **kw).audio.cpu()
    finally:
        L.lib().mt_codec_debug_set_attention(0)
    new = model.decode(codes.to(DEV), **kw).audio.cpu()
    ref, _ = oracle.decode(codes, **kw)
    s_old, s_new, s_pair = snr_db(ref, old), snr_db(ref, new), snr_db(old, new)
    print(f"[attention B={B} T={T} chunk={chunk}] vs oracle: mma.sync {s_old:.1f} dB, tcgen05 {s_new:.1f} dB; "
          f"tcgen05 vs mma.sync {s_pair:.1f} dB")
    assert torch.isfinite(new).all()
    assert s_new >= SNR_DB and s_new >= s_old - 1.0
    assert s_pair >= 45.0
