"""Incremental streaming decode (`with model.streaming(B): model._decode_frame(chunk)`) through mt_codec_stream_* vs the
reference's own streaming run (tests/golden/codec_stream_mini.npz) and the oracle's closed form."""
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
SNR_DB = 40.0


@pytest.fixture(scope="module")
def st():
    g = np.load(os.path.join(GOLDEN, "codec_stream_mini.npz"), allow_pickle=False)
    cfg = CodecConfig.from_dict(json.loads(str(g["cfg"])))
    w = make_codec_weights(cfg, seed=int(g["seed"]))
    return g, cfg, MossAudioTokenizerModel(cfg, w, device=DEV), OracleCodec(cfg, w)


def _stream(model, codes, pushes):
    outs, pos = [], 0
    with model.streaming(batch_size=codes.shape[1]):
        for n in pushes:
            o = model._decode_frame(codes[:, :, pos: pos + n])
            assert o.audio.shape[-1] == n * model.downsample_rate
            outs.append(o.audio)
            pos += n
    return torch.cat(outs, dim=-1).cpu()


def test_streaming_pushes_match_reference_streaming(st):
    g, cfg, model, oracle = st
    codes = torch.from_numpy(g["codes"]).to(DEV)
    wav = _stream(model, codes, [int(n) for n in g["pushes"]])
    s = snr_db(torch.from_numpy(g["wav_stream"]), wav)
    print(f"[codec streaming, uneven pushes] SNR vs the reference's streaming run {s:.1f} dB")
    assert s >= SNR_DB
    assert snr_db(torch.from_numpy(g["wav_full"]), wav) < 30.0          # streaming is not one-shot decoding
    # a second session starts from a clean state (stale ring contents are never visible)
    wav2 = _stream(model, codes, [int(n) for n in g["pushes"]])
    assert torch.equal(wav, wav2)


def test_streaming_equals_chunked_decode_for_uniform_pushes(st):
    g, cfg, model, oracle = st
    codes = torch.from_numpy(g["codes"])[:, :1, :33].to(DEV)
    wav = _stream(model, codes, [10, 10, 10, 3])
    ref = model.decode(codes, chunk_duration=0.8).audio.cpu()           # the one-pass form of the same ring rule
    assert snr_db(ref, wav) >= 60.0


@pytest.mark.parametrize("pushes", [[1] * 30, [12, 12, 12], [5, 1, 11, 2, 9]])
def test_streaming_vs_oracle(st, pushes):
    g, cfg, model, oracle = st
    T = sum(pushes)
    codes = torch.randint(0, cfg.codebook_size, (cfg.num_quantizers, 3, T), generator=torch.Generator().manual_seed(T))
    wav = _stream(model, codes.to(DEV), pushes)
    ref = oracle.decode_frames(codes, chunk_frames=pushes)
    assert snr_db(ref, wav) >= SNR_DB


def test_streaming_argument_errors(st):
    g, cfg, model, oracle = st
    codes = torch.from_numpy(g["codes"]).to(DEV)
    with model.streaming(batch_size=2):
        with pytest.raises(RuntimeError):
            with model.streaming(batch_size=2):
                pass
        with pytest.raises(ValueError):
            model._decode_frame(codes[:, :1, :4])                       # wrong batch
        with pytest.raises(ValueError):
            model._decode_frame(codes[:, :, :13])                       # push longer than the 12-frame context
        with pytest.raises(RuntimeError):
            model.decode(codes[:, :1, :30], chunk_duration=0.8)         # nested streaming, as in the reference
    tf32 = MossAudioTokenizerModel(cfg, make_codec_weights(cfg, seed=int(g["seed"])), device=DEV, layer_gemm="tf32")
    with pytest.raises(NotImplementedError):
        with tf32.streaming(batch_size=1):
            pass
