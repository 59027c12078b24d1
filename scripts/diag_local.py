"""Diagnostic: per-channel teacher-forced logit drift of the MossTTSLocal CUDA path vs the oracle (GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from test_oracle_local import load_local_golden
from moss_tts_b200.modeling_moss_tts_local import MossTTSLocalModel
from oracle.local import OracleLocal
from oracle.weights import make_local_weights

for name in sys.argv[1:] or ["local_mid_bf16"]:
    g, meta, cfg = load_local_golden(name)
    w = make_local_weights(cfg, seed=meta["seed"])
    model = MossTTSLocalModel(cfg, w)
    oracle = OracleLocal(cfg, w, torch.bfloat16)
    truth = OracleLocal(cfg, {k: v.to(torch.bfloat16).float() for k, v in w.items()}, torch.float32)
    gen = torch.from_numpy(g["gen_ids"]); S0 = g["input_ids"].shape[1]
    for f in (0, 3):
        ctx, frame = gen[:, :S0 + f], gen[:, S0 + f]
        mine = model.frame_logits(ctx, frame); ref = oracle.frame_logits(ctx, frame); tr = truth.frame_logits(ctx, frame)
        for c, (m, r, t) in enumerate(zip(mine, ref, tr)):
            m = m.float().cpu(); r = r.float(); fin = torch.isfinite(r)
            print(f"   vs fp32 truth: cuda mean|e| {(m - t).abs()[fin].mean():.5f} max {(m - t).abs()[fin].max():.4f} | "
                  f"oracle-bf16 mean|e| {(r - t).abs()[fin].mean():.5f} max {(r - t).abs()[fin].max():.4f}")
            d = (m - r).abs()[fin]
            rel = ((m - r) / r.abs().clamp_min(1e-3))[fin]
            print(f"{name} f{f} c{c}: max|d| {d.max():.4f} mean|d| {d.mean():.5f} frac>0 {float((d>0).float().mean()):.3f} "
                  f"max|logit| {r[fin].abs().max():.3f} median rel {rel.median():+.5f} per-row median rel "
                  + " ".join(f"{((m[b]-r[b])/r[b].abs().clamp_min(1e-3))[torch.isfinite(r[b])].median():+.4f}" for b in range(m.shape[0])))
