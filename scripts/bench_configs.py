"""Measurements of the BASELINE.json configurations that are not the bench.py headline (configs[1]):
  C3  codec decode sweep                     B in {1,4,16,64,256} x T in {13,63,125,375,750} frames
  C4  MossTTSLocal (assumed 1.7B shape)      B = 128, 96-token prompt, 125 frames, n_vq_for_inference = 32
  C5  MossTTSDelay n_vq = 16 long context    B = 4, ~512-token prompt, 3765 steps (>= 5 min of audio)
  stream  incremental streaming decode: per-push latency (SURVEY §8f rank 3 building block)
  synth  streaming synthesis: generate_stream at the 8B / released-codec shapes, time to first audio
         and total time vs generate + one-shot decode
  rt  MOSS-TTS-Realtime step-wise decode: ms per frame at the assumed 1.7B shape
  enc codec ENCODE sweep B in {1,16,64} x {1,10,30} s of 24 kHz audio -> 32 codebooks
One JSON line per measurement; device-timed with CUDA events after a warm-up call.

  python scripts/bench_configs.py [--what c3,c4,c5] [--quick]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


RANK = int(os.environ.get("RANK", 0))
WORLD = int(os.environ.get("WORLD_SIZE", 1))
LOCAL = int(os.environ.get("LOCAL_RANK", 0))


def _barrier():
    if WORLD > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn, reps=1):
    """Device time of `reps` calls; under torchrun (request-sharded runs of C4 / C5): barrier on both sides, max over ranks."""
    _barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        r = fn()
    e1.record()
    _barrier()
    ms = e0.elapsed_time(e1) / reps
    if WORLD > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms, r


def c3(args, dev):
    from moss_tts_b200.config import CodecConfig
    from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
    ccfg = CodecConfig.released()
    codec = MossAudioTokenizerModel.random_init(ccfg, device=dev, seed=99, workspace_bytes=24 << 30)
    Bs = [1, 16, 256] if args.quick else [1, 4, 16, 64, 256]
    Ts = [13, 125, 750] if args.quick else [13, 63, 125, 375, 750]
    for B in Bs:
        for T in Ts:
            codes = torch.randint(0, 1024, (ccfg.num_quantizers, B, T), device=dev)
            codec.decode(codes)                                     # warm-up (workspace, attributes)
            ms, _ = timed(lambda: codec.decode(codes), reps=2 if B * T > 20000 else 5)
            ms_c, _ = timed(lambda: codec.decode(codes, chunk_duration=8, allow_batched_chunks=True),
                            reps=2 if B * T > 20000 else 5)
            print(json.dumps({"config": "C3 codec decode", "B": B, "T_frames": T, "audio_s": B * T / 12.5, "ms": ms,
                              "audio_s_per_s": B * T / 12.5 / (ms * 1e-3), "ms_chunked_8s": ms_c,
                              "audio_s_per_s_chunked": B * T / 12.5 / (ms_c * 1e-3)}), flush=True)


def enc(args, dev):
    from moss_tts_b200.config import CodecConfig
    from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
    ccfg = CodecConfig.released()
    codec = MossAudioTokenizerModel.random_init(ccfg, device=dev, seed=99, workspace_bytes=24 << 30, encoder=True)
    for B in ([1, 16] if args.quick else [1, 16, 64]):
        for sec in ([1, 10] if args.quick else [1, 10, 30]):
            wav = torch.randn(B, 1, sec * ccfg.sampling_rate, device=dev) * 0.1
            codec.encode(wav)
            ms, out = timed(lambda: codec.encode(wav), reps=3)
            ms_q, _ = timed(lambda: codec.quantizer(out.encoder_hidden_states, out.audio_codes_lengths), reps=3)
            print(json.dumps({"config": "codec encode (released shape, 68 layers + 32-codebook RLFQ)", "B": B, "audio_s": B * sec,
                              "frames": int(out.audio_codes.shape[-1]), "ms": ms, "audio_s_per_s": B * sec / (ms * 1e-3),
                              "ms_quantiser_alone": ms_q}), flush=True)


def stream(args, dev):
    """Incremental streaming decode at the released shape: per-push latency for 0.32 s pushes (4 frames, what
    moss_tts_realtime's AudioStreamDecoder uses), steady state = ring caches full (10 s of context)."""
    from moss_tts_b200.config import CodecConfig
    from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
    ccfg = CodecConfig.released()
    codec = MossAudioTokenizerModel.random_init(ccfg, device=dev, seed=99, workspace_bytes=8 << 30)
    for B, n in ((1, 4), (1, 1), (16, 4), (64, 4)):
        pushes = 60 if n == 4 else 160
        codes = torch.randint(0, 1024, (ccfg.num_quantizers, B, n * pushes), device=dev)
        ms = []
        with codec.streaming(batch_size=B):
            for i in range(pushes):
                t, _ = timed(lambda: codec._decode_frame(codes[:, :, i * n: (i + 1) * n]))
                ms.append(t)
        tail = sorted(ms[len(ms) // 2:])
        print(json.dumps({"config": "codec streaming decode (released shape, ring KV caches)", "B": B, "frames_per_push": n,
                          "audio_ms_per_push": n * 80, "first_push_ms": ms[0], "steady_push_ms_median": tail[len(tail) // 2],
                          "realtime_factor": B * n * 80 / tail[len(tail) // 2]}), flush=True)


def synth(args, dev):
    """Streaming synthesis at the headline shapes: B rows x 10 s, chunks of `cf` frames.  Wall-clock (host) timing:
    time to first audio is a host-visible latency, the chunk is on the device and synchronised when it is yielded."""
    import time
    import bench
    from moss_tts_b200.config import CodecConfig, LMConfig
    from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
    from moss_tts_b200.modeling_moss_tts import MossTTSDelayModel
    from moss_tts_b200.streaming import generate_stream
    cfg = LMConfig.delay_8b()
    lm = MossTTSDelayModel.random_init(cfg, device=dev, seed=1234, never_stop=True)
    ccfg = CodecConfig.released()
    codec = MossAudioTokenizerModel.random_init(ccfg, device=dev, seed=99, workspace_bytes=8 << 30)
    steps = 60 if args.quick else 156
    for B, cf, fc in ((1, 4, None), (1, 8, None), (8, 8, None), (64, 8, None), (64, 25, None), (64, 25, 8), (64, 25, 4)):
        ids, mask = bench.make_prompts(cfg, B, seed=0)
        kw = dict(max_new_tokens=steps, text_temperature=0, audio_temperature=0)
        for rep in range(2):                                   # rep 0 warms graphs / lazy attributes
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            first, n_chunks, frames = None, 0, 0
            for wav, valid in generate_stream(lm, codec, ids, mask, chunk_frames=cf, first_chunk_frames=fc, **kw):
                if first is None:
                    first = time.perf_counter() - t0
                n_chunks += 1
                frames += wav.shape[1] // codec.downsample_rate
            torch.cuda.synchronize()
            total = time.perf_counter() - t0
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = lm.generate(ids, mask, **kw)
        torch.cuda.synchronize()
        t_gen = time.perf_counter() - t0
        print(json.dumps({"config": "streaming synthesis (8B shape + released codec, greedy)", "B": B, "chunk_frames": cf, "first_chunk_frames": fc,
                          "steps": steps, "chunks": n_chunks, "frames_per_row": frames,
                          "first_audio_ms": None if first is None else first * 1e3, "total_ms": total * 1e3,
                          "generate_only_ms": t_gen * 1e3,
                          "audio_s_per_s": B * frames / 12.5 / total}), flush=True)


def c4(args, dev):
    from moss_tts_b200.config import LocalConfig
    from moss_tts_b200.modeling_moss_tts_local import MossTTSLocalModel
    import bench
    cfg = LocalConfig.local_1p7b(32)
    model = MossTTSLocalModel.random_init(cfg, device=dev, seed=4321 + RANK)
    B_total = 32 if args.quick else 128
    B = B_total // WORLD                                  # request sharding: one group of utterances per GPU
    frames = 16 if args.quick else 125
    bench.PROMPT_LEN = 96
    ids, _ = bench.make_prompts(cfg, B, seed=0)
    ids = ids.to(dev)
    kw = dict(max_new_tokens=frames, n_vq_for_inference=cfg.n_vq, eos_token_id=None, record_step_times=True)
    model.generate(ids, **kw)
    ms, out = timed(lambda: model.generate(ids, **kw))
    steps = sorted(model.last_step_ms[2:])
    if RANK == 0:
        print(json.dumps({"config": "C4 MossTTSLocal greedy", "shape": "assumed Qwen3-1.7B backbone + local 4L/1536/8960/2048",
                          "n_gpus": WORLD, "B_total": B_total, "B_per_gpu": B, "frames": frames, "ms_total": ms,
                          "prefill_plus_first_frame_ms": model.last_step_ms[0], "ms_per_frame_median": steps[len(steps) // 2],
                          "audio_s_per_s": B * WORLD * frames / 12.5 / (ms * 1e-3), "weight_GB": model.weight_bytes() / 1e9,
                          "sharding": "request-sharded, no in-step collective; timing = max over ranks"}), flush=True)


def c5(args, dev):
    from moss_tts_b200.config import LMConfig
    from moss_tts_b200.modeling_moss_tts import MossTTSDelayModel
    import bench
    cfg = LMConfig.delay_8b(n_vq=16)
    model = MossTTSDelayModel.random_init(cfg, device=dev, seed=77 + RANK, never_stop=True)
    B = 4                                                 # per GPU; BASELINE config 5 shards 32 utterances over 8 GPUs
    steps = 400 if args.quick else 3765
    bench.PROMPT_LEN = 512
    ids, mask = bench.make_prompts(cfg, B, seed=0)
    g = torch.Generator().manual_seed(1)
    ids[:, 100:500, 1:] = torch.randint(0, 1024, (B, 400, cfg.n_vq), generator=g)      # reference codes in the prompt
    ids, mask = ids.to(dev), mask.to(dev)
    for name, kw in (("greedy pen=1.0", dict(text_temperature=0, audio_temperature=0)),
                     ("sampled T=1.1 k=50 p=0.9 pen=1.1", dict(text_temperature=1.1, text_top_k=50, text_top_p=0.9,
                                                              audio_temperature=1.1, audio_top_k=50, audio_top_p=0.9,
                                                              audio_repetition_penalty=1.1))):
        kw = dict(max_new_tokens=steps, record_step_times=True, **kw)
        model.generate(ids, mask, **dict(kw, max_new_tokens=8))
        ms, out = timed(lambda: model.generate(ids, mask, **kw))
        st = model.last_step_ms
        n = model.last_n_steps
        tail = sorted(st[-200:])
        if RANK == 0:
          print(json.dumps({"config": "C5 MossTTSDelay n_vq=16 long context", "mode": name, "n_gpus": WORLD, "B_per_gpu": B, "B_total": B * WORLD, "prompt": 512,
                          "steps_run": n, "ms_total": ms, "prefill_ms": st[0], "ms_per_step_first_200": sorted(st[2:202])[100],
                          "ms_per_step_last_200": tail[len(tail) // 2], "final_kv_len": 512 + n,
                          "audio_s_per_s": B * WORLD * max(n - cfg.n_vq, 0) / 12.5 / (ms * 1e-3)}), flush=True)


def rt(args, dev):
    """MOSS-TTS-Realtime step-wise decode at the assumed released shape (Qwen3-1.7B backbone + the
    reference config's 4-layer local transformer, rvq 16): prefill of a 96-position prompt, then finish() for 125 frames
    with the reference's default sampling (T 0.8, top-p 0.6, top-k 30, penalty 1.1 over a 50-frame window)."""
    from moss_tts_b200.config import RealtimeConfig
    from moss_tts_b200.modeling_mossttsrealtime import MossTTSRealtimeInference, MossTTSRealtimeModel
    import numpy as np
    cfg = RealtimeConfig.realtime_1p7b()
    model = MossTTSRealtimeModel.random_init(cfg, device=dev, seed=2468)
    frames = 32 if args.quick else 125
    for B in ((1, 16) if args.quick else (1, 16, 64, 128)):
        rng = np.random.default_rng(B)
        prompts = []
        for _ in range(B):
            p = np.full((92, cfg.rvq + 1), cfg.audio_pad_code, dtype=np.int64)
            p[:, 0] = rng.integers(0, cfg.pad_token_id, 92)
            p[:40, 0] = cfg.reference_audio_pad
            p[:40, 1:] = rng.integers(0, 1024, (40, cfg.rvq))
            prompts.append(p)
        prefixes = [rng.integers(0, cfg.pad_token_id, 4).tolist() for _ in range(B)]
        for mode, kw in (("sampled (reference defaults)", {}), ("greedy + penalty", {"do_sample": False})):
            inf = MossTTSRealtimeInference(model, None, max_length=frames + 4, channels=cfg.rvq, max_batch=B, seed=1,
                                           audio_eos_token=4095)            # never emitted: every run makes all frames
            inf.prefill(prompts, prefixes, **kw)
            inf.finish(max_steps=4, **kw)                                    # graph capture / warm-up
            inf = MossTTSRealtimeInference(model, None, max_length=frames + 4, channels=cfg.rvq, max_batch=B, seed=1,
                                           audio_eos_token=4095)
            ms_pre, _ = timed(lambda: inf.prefill(prompts, prefixes, **kw))
            ms, outs = timed(lambda: inf.finish(max_steps=frames, **kw))
            n = len(outs)
            print(json.dumps({"config": "MOSS-TTS-Realtime step-wise decode", "shape": "assumed Qwen3-1.7B backbone + local 4L (rvq 16)",
                              "mode": mode, "B": B, "frames": n, "prefill_plus_first_frame_ms": ms_pre, "ms_total": ms,
                              "ms_per_frame": ms / max(n, 1), "audio_s_per_s": B * n / 12.5 / (ms * 1e-3),
                              "x_realtime_per_stream": n / 12.5 / (ms * 1e-3), "weight_GB": model.weight_bytes() / 1e9}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="c3,c4,c5")
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    dev = torch.device("cuda", LOCAL)
    torch.cuda.set_device(LOCAL)
    if WORLD > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    for w in args.what.split(","):
        {"c3": c3, "c4": c4, "c5": c5, "enc": enc, "stream": stream, "synth": synth, "rt": rt}[w](args, dev)
        torch.cuda.empty_cache()
    if WORLD > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
