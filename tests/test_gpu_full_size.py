"""Size-independent properties at BASELINE.json's full sizes (configs[1]: assumed released 8B shape, B = 64, 10 s;
released codec shape).  The oracle cannot run these sizes in seconds, so the checks are structural: the delay-pattern
state machine's row pattern (SURVEY.md §8c observed KAT), run-to-run and CUDA-graph on/off bit-identity, de-delay /
delay round trips, decode(batch) == decode(single) for the per-sample-independent codec, chunked == streaming pushes,
and an encode -> decode -> encode length / shape round trip."""
import pytest
import torch

from moss_tts_b200.config import CodecConfig, LMConfig
from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
from moss_tts_b200.modeling_moss_tts import MossTTSDelayModel
from moss_tts_b200.processing_moss_tts import MossTTSDelayProcessor
from gpu_util import DEV, snr_db

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lm8b():
    import bench
    cfg = LMConfig.delay_8b()
    model = MossTTSDelayModel.random_init(cfg, device=DEV, seed=1234, never_stop=True)
    ids, mask = bench.make_prompts(cfg, 64, seed=0)
    return cfg, model, ids.to(DEV), mask.to(DEV)


def test_8b_generate_row_pattern_and_determinism(lm8b):
    cfg, model, ids, mask = lm8b
    steps = 48
    kw = dict(max_new_tokens=steps, text_temperature=0, audio_temperature=0)
    out = model.generate(ids, mask, **kw)
    P = cfg.audio_pad_code
    for sl, g in out:
        assert sl == 1 and tuple(g.shape) == (steps + 1, 1 + cfg.n_vq)
        g = g.cpu()
        assert int(g[0, 0]) == cfg.audio_start_token_id and bool((g[0, 1:] == P).all())
        # never-stop tie: every generated text token is the gen slot (argmax picks the lower index of the tie)
        assert bool((g[1:, 0] == cfg.audio_assistant_gen_slot_token_id).all())
        for r in range(1, steps + 1):                       # delay pattern: channel i starts at generated row i + 1
            live = g[r, 1:1 + min(r, cfg.n_vq)]
            assert bool(((live >= 0) & (live < P)).all()), r
            assert bool((g[r, 1 + min(r, cfg.n_vq):] == P).all()), r
    again = model.generate(ids, mask, **kw)
    eager = model.generate(ids, mask, use_cuda_graph=False, **kw)
    for (_, a), (_, b), (_, c) in zip(out, again, eager):
        assert torch.equal(a, b) and torch.equal(a, c)
    # rows are independent under greedy decoding with penalty 1.0: a sub-batch reproduces its rows
    # (different batch -> different GEMM column padding only; accumulation order per row is unchanged)
    sub = model.generate(ids[:8], mask[:8], **kw)
    for (_, a), (_, b) in zip(out[:8], sub):
        assert torch.equal(a, b)


def test_8b_dedelay_roundtrip_and_frame_count(lm8b):
    cfg, model, ids, mask = lm8b
    out = model.generate(ids, mask, max_new_tokens=40, text_temperature=0, audio_temperature=0)
    g = out[0][1][:, 1:]
    dd = MossTTSDelayProcessor.apply_de_delay_pattern(g)
    assert dd.shape[0] == g.shape[0] - cfg.n_vq + 1
    back = MossTTSDelayProcessor.apply_delay_pattern(dd, cfg.audio_pad_code)
    # delay(de-delay(x)) restores every cell the delay pattern carries
    T = dd.shape[0]
    for i in range(cfg.n_vq):
        assert torch.equal(back[i: i + T, i], g[i: i + T, i])


@pytest.fixture(scope="module")
def codec_full():
    ccfg = CodecConfig.released()
    return ccfg, MossAudioTokenizerModel.random_init(ccfg, device=DEV, seed=99, workspace_bytes=8 << 30, encoder=True,
                                                     layer_scale=0.3)


def test_released_codec_batch_independence_and_streaming(codec_full):
    ccfg, codec = codec_full
    codes = torch.randint(0, 1024, (32, 4, 150), device=DEV, generator=torch.Generator(device=DEV).manual_seed(5))
    full = codec.decode(codes).audio
    assert tuple(full.shape) == (4, 1, 150 * 1920) and bool(torch.isfinite(full).all())
    one = codec.decode(codes[:, 2:3]).audio
    assert snr_db(one[0].cpu(), full[2].cpu()) >= 60.0                  # per-sample independent
    nq8 = codec.decode(codes, num_quantizers=8).audio                   # variable bitrate prefix really changes the output
    assert snr_db(full.cpu(), nq8.cpu()) < 20.0
    # 150 frames > 125-frame context: chunked decode (8 s chunks) == two streaming pushes of 100 + 50 frames
    ch = codec.decode(codes[:, :1], chunk_duration=8).audio
    with codec.streaming(batch_size=1):
        a = codec._decode_frame(codes[:, :1, :100]).audio
        b = codec._decode_frame(codes[:, :1, 100:]).audio
    assert snr_db(ch.cpu(), torch.cat([a, b], -1).cpu()) >= 60.0
    assert snr_db(full[:1].cpu(), ch.cpu()) < 60.0                      # and the ring rule matters at this length


def test_released_codec_encode_shapes_and_determinism(codec_full):
    ccfg, codec = codec_full
    wav = torch.randn(3, 1, 24000 * 4 + 777, device=DEV, generator=torch.Generator(device=DEV).manual_seed(6)) * 0.1
    lens = torch.tensor([wav.shape[-1], 24000 * 2, 1920 * 5 + 1], device=DEV)
    pm = torch.arange(wav.shape[-1], device=DEV)[None, :] < lens[:, None]
    enc = codec.encode(wav, padding_mask=pm)
    Tf = (wav.shape[-1] + 1919) // 1920
    assert tuple(enc.audio_codes.shape) == (32, 3, Tf)
    assert enc.audio_codes_lengths.tolist() == [wav.shape[-1] // 1920, 25, 5]
    assert int(enc.audio_codes.min()) >= 0 and int(enc.audio_codes.max()) < 1024
    again = codec.encode(wav, padding_mask=pm)
    assert torch.equal(enc.audio_codes, again.audio_codes)
    # causal encoder: a sample's valid codes do not depend on what follows its length or on its batch neighbours
    solo = codec.encode(wav[1:2, :, : 24000 * 2])
    assert torch.equal(solo.audio_codes[:, 0, :25], enc.audio_codes[:, 1, :25])
    # decode what we encoded: shapes / lengths line up
    dec = codec.decode(enc.audio_codes, padding_mask=torch.arange(Tf, device=DEV)[None, :] < enc.audio_codes_lengths[:, None])
    assert tuple(dec.audio.shape) == (3, 1, Tf * 1920)
    assert dec.audio_lengths.tolist() == [n * 1920 for n in enc.audio_codes_lengths.tolist()]


def test_8b_streaming_synthesis_equals_generate_plus_streaming_decode(lm8b, codec_full):
    """BASELINE shapes (8B LM, released codec): the sliced loop reproduces generate()'s ids bit-for-bit and the chunks it
    yields are the codec's streaming decode of the frames those ids carry (same pushes -> bit-exact)."""
    from moss_tts_b200.streaming import generate_stream
    cfg, model, ids, mask = lm8b
    ccfg, codec = codec_full
    B, steps, cf = 8, 72, 8
    kw = dict(max_new_tokens=steps, text_temperature=0, audio_temperature=0)
    ref = model.generate(ids[:B], mask[:B], **kw)
    chunks = list(generate_stream(model, codec, ids[:B], mask[:B], chunk_frames=cf, **kw))
    for (a, x), (b, y) in zip(ref, model.last_stream_output):
        assert a == b and torch.equal(x, y)
    nq, ds = cfg.n_vq, codec.downsample_rate
    pushes = [w.shape[1] // ds for w, _ in chunks]
    frames = steps - (nq - 1)                               # never-stop weights: every generated row opens a frame
    assert sum(pushes) == frames and all(p >= cf for p in pushes[:-1]) and len(pushes) >= 4
    assert all(v == [p * ds] * B for (_, v), p in zip(chunks, pushes))
    codes = torch.stack([MossTTSDelayProcessor.apply_de_delay_pattern(g[sl:, 1:]).T for sl, g in ref], 1)   # [nq, B, T]
    assert codes.shape[-1] == frames
    outs, pos = [], 0
    with codec.streaming(batch_size=B):
        for n in pushes:
            outs.append(codec._decode_frame(codes[:, :, pos: pos + n].contiguous()).audio[:, 0])
            pos += n
    assert torch.equal(torch.cat(outs, -1), torch.cat([w for w, _ in chunks], -1))
    # and one un-chunked decode of the same codes agrees while everything is inside the 10 s context
    full = codec.decode(codes).audio[:, 0]
    assert snr_db(full.cpu(), torch.cat(outs, -1).cpu()) >= 45.0
