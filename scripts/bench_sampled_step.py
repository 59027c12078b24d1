"""Decode-step cost of the sampler at the headline shape (8B LM, B = 64, 96-token prompt): greedy vs the reference's
default audio sampling (T = 1.7, top-p 0.8, top-k 25) with and without the repetition penalty.  Text stays greedy so the
never-stop weights keep every row inside its audio span (same work per step in all modes).  One JSON line per mode.

  python scripts/bench_sampled_step.py [--batch 64] [--steps 100]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--steps", type=int, default=100)
    args = ap.parse_args()
    import bench
    from moss_tts_b200.config import LMConfig
    from moss_tts_b200.modeling_moss_tts import MossTTSDelayModel
    dev = torch.device("cuda", 0)
    cfg = LMConfig.delay_8b()
    model = MossTTSDelayModel.random_init(cfg, device=dev, seed=1234, never_stop=True)
    ids, mask = bench.make_prompts(cfg, args.batch, seed=0)
    modes = {
        "greedy": dict(text_temperature=0, audio_temperature=0),
        "audio sampled T=1.7 p=0.8 k=25": dict(text_temperature=0, audio_temperature=1.7, audio_top_p=0.8, audio_top_k=25),
        "audio sampled + penalty 1.1": dict(text_temperature=0, audio_temperature=1.7, audio_top_p=0.8, audio_top_k=25,
                                            audio_repetition_penalty=1.1),
        "audio sampled, no top-k (1024 codes) p=0.8": dict(text_temperature=0, audio_temperature=1.7, audio_top_p=0.8,
                                                            audio_top_k=None),
    }
    for name, kw in modes.items():
        for _ in range(2):
            model.generate(ids, mask, max_new_tokens=args.steps, record_step_times=True, **kw)
        ms = sorted(model.last_step_ms[2:])
        print(json.dumps({"mode": name, "B": args.batch, "steps_run": model.last_n_steps,
                          "ms_per_step_median": ms[len(ms) // 2], "ms_per_step_p90": ms[int(len(ms) * 0.9)]}), flush=True)


if __name__ == "__main__":
    main()
