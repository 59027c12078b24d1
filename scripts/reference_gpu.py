"""SURVEY.md §8(d) row "Reference GPU timing (the bar)": the UNMODIFIED reference (oracle/_ref, built by oracle/ref.py) on
1 x B200 through its own public path -- MossTTSDelayModel.generate (bf16, attn_implementation="sdpa": cuBLAS GEMMs + PyTorch
SDPA + ATen) followed by MossTTSDelayProcessor._parse_audio_codes per utterance (fp32 codec, chunk_duration=8, B=1 per
segment as the reference requires) -- on bench.py's workload (configs[1]: B=64, 96-token prompt, 156 greedy steps, 125 frames).
Measurement script, not product code: it times the competitor SURVEY §2.4 names.  Prints one JSON line.

    python scripts/reference_gpu.py [--batch 64] [--steps 156] [--runs 2]
"""
from __future__ import annotations

import argparse
import contextlib
import io
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch                                                    # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=156)
    ap.add_argument("--runs", type=int, default=2)
    args = ap.parse_args()
    import bench
    from moss_tts_b200.config import LMConfig
    from oracle import ref
    from transformers.models.qwen3.modeling_qwen3 import Qwen3DecoderLayer      # noqa: F401
    assert ref.available() and torch.cuda.is_available()
    dev = torch.device("cuda:0")
    cfg = LMConfig.delay_8b()
    MossTTSDelayConfig, MossTTSDelayModel, _, _ = ref.import_delay()
    torch.manual_seed(0)
    t0 = time.perf_counter()
    lc = ref.qwen3_config(cfg, attn_implementation="sdpa")
    mc = MossTTSDelayConfig(language_config=lc, n_vq=cfg.n_vq)
    with torch.device(dev):                                       # all 36 real, distinct layers, initialised on the GPU
        lm = MossTTSDelayModel(mc)
    lm = lm.to(torch.bfloat16).eval()
    lm.config._attn_implementation = "sdpa"
    lm.language_model.config._attn_implementation = "sdpa"
    with torch.no_grad():
        w = lm.lm_heads[0].weight                                  # never-stop tie so every run is 156 steps
        w[mc.audio_assistant_delay_slot_token_id] = w[mc.audio_assistant_gen_slot_token_id]
    with torch.device(dev):
        codec = ref.build_codec_bench()
    codec = codec.to(dev).eval()                                   # fp32: config.json dtype, the reference never casts it
    proc = ref.bare_processor(codec)
    t_build = time.perf_counter() - t0
    ids, mask = bench.make_prompts(cfg, args.batch, 0)
    ids, mask = ids.to(dev), mask.to(dev)

    def one(steps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        with contextlib.redirect_stderr(io.StringIO()), torch.inference_mode():
            out = lm.generate(ids, mask, max_new_tokens=steps, text_temperature=0, audio_temperature=0)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        samples = 0
        with torch.inference_mode():
            for sl, g in out:
                for wv in proc._parse_audio_codes(int(sl), g[:, 1:]):
                    samples += wv.numel()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        return t1 - t0, t2 - t1, samples

    one(40)                                                        # warm-up (cuBLAS heuristics, allocator)
    runs = [one(args.steps) for _ in range(args.runs)]
    gen_s = min(r[0] for r in runs)
    codec_s = min(r[1] for r in runs)
    frames = args.batch * (args.steps + 1 - cfg.n_vq)
    line = {"what": "reference (unmodified, oracle/_ref) on 1 x B200: the competitor of SURVEY 2.4 / 8(d)",
            "metric": bench.METRIC, "value": frames / 12.5 / (gen_s + codec_s), "unit": "audio-s/s",
            "batch": args.batch, "steps": args.steps, "frames": frames,
            "generate_s": gen_s, "ms_per_step_incl_prefill": gen_s / args.steps * 1e3, "codec_s": codec_s,
            "samples": runs[0][2], "lm": "bf16, sdpa, 36 distinct layers, assumed 8B shape", "codec": "fp32, released config.json, "
            "processor path (one decode per utterance, chunk_duration=8)", "runs": [[r[0], r[1]] for r in runs],
            "build_s": t_build, "torch": torch.__version__, "gpu": torch.cuda.get_device_name(0)}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
