"""Short pass over the hot path at the bench shapes for ncu launch lists / full captures.

  python scripts/profile_path.py [--steps 6] [--batch 64] [--frames 125] [--codec-batch 64] [--what lm,codec]

Same shapes as bench.py (8B-shape LM, released codec) but only a handful of decode steps, so the
whole launch list is a few thousand kernels and ncu needs no skip window.  Prints the launch count
of every phase so a window can be picked with -s/-c.
"""
from __future__ import annotations

import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--prompt", type=int, default=96)
    ap.add_argument("--frames", type=int, default=125)
    ap.add_argument("--codec-batch", type=int, default=64)
    ap.add_argument("--what", default="lm,codec")
    ap.add_argument("--repeat", type=int, default=1)
    ap.add_argument("--no-graph", action="store_true")
    args = ap.parse_args()
    from moss_tts_b200 import _lib as L
    from moss_tts_b200.config import LMConfig, CodecConfig
    from moss_tts_b200.modeling_moss_tts import MossTTSDelayModel
    from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
    import bench

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    what = args.what.split(",")
    if "lm" in what:
        cfg = LMConfig.delay_8b()
        model = MossTTSDelayModel.random_init(cfg, device=dev, seed=1234, never_stop=True)
        bench.PROMPT_LEN = args.prompt
        ids, mask = bench.make_prompts(cfg, args.batch, seed=0)
        ids, mask = ids.to(dev), mask.to(dev)
        for r in range(args.repeat):
            n0 = L.launch_count()
            model.generate(ids, mask, max_new_tokens=args.steps, text_temperature=0, audio_temperature=0,
                           record_step_times=True, use_cuda_graph=not args.no_graph)
            torch.cuda.synchronize()
            ms = model.last_step_ms
            print(f"lm: {L.launch_count() - n0} launches; prefill {ms[0]:.2f} ms; decode steps "
                  + " ".join(f"{m:.3f}" for m in ms[1:]), flush=True)
    if "codec" in what:
        ccfg = CodecConfig.released()
        codec = MossAudioTokenizerModel.random_init(ccfg, device=dev, seed=99, workspace_bytes=24 << 30)
        codes = torch.randint(0, 1024, (ccfg.num_quantizers, args.codec_batch, args.frames), device=dev)
        for r in range(args.repeat):
            n0 = L.launch_count()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            codec.decode(codes, chunk_duration=8, allow_batched_chunks=True)
            e1.record()
            torch.cuda.synchronize()
            print(f"codec: {L.launch_count() - n0} launches; {e0.elapsed_time(e1):.2f} ms", flush=True)


if __name__ == "__main__":
    main()
