"""Standalone rooflines of the HBM-bound small kernels (SURVEY.md §8d "sampling", "RVQ dequant"):
mt_sample_tokens over many rows so that the logits (one read, the algorithmic bytes) exceed L2, and mt_rvq_dequant at
the released codec shape over config 3's largest case (B = 256, T = 750).  Prints one JSON line per case.

  python scripts/bench_sampling.py [--rows 2048] [--iters 5]
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
    ap.add_argument("--rows", type=int, default=2048)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--what", default="all", choices=["all", "sampling", "rvq"])
    args = ap.parse_args()
    if args.what in ("all", "rvq"):
        rvq(args)
    if args.what == "rvq":
        return
    from moss_tts_b200 import _lib as L
    from moss_tts_b200.inference_utils import sample_token
    import bench
    hbm, _, src = bench.measured_peaks()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    cases = [("text greedy", args.rows, 151936, dict(do_sample=False)),
             ("text top-k 50", args.rows, 151936, dict(do_sample=True, top_k=50, top_p=1.0)),
             ("audio greedy", args.rows * 64, 1025, dict(do_sample=False)),
             ("audio top-k 25 top-p 0.8", args.rows * 64, 1025, dict(do_sample=True, top_k=25, top_p=0.8)),
             ("audio top-k 25 top-p 0.8 penalty 1.1", args.rows * 64, 1025,
              dict(do_sample=True, top_k=25, top_p=0.8, repetition_penalty=1.1))]
    for name, n, v, kw in cases:
        logits = (torch.randn(n, v, device=dev) * 2).to(torch.bfloat16)
        prev = torch.randint(0, v, (4096,), device=dev) if "repetition_penalty" in kw else None
        u = torch.rand(n, device=dev)
        import ctypes as C
        from moss_tts_b200.inference_utils import _bitmap_from_tokens
        bitmap = _bitmap_from_tokens(prev, v).contiguous() if prev is not None else None
        out = torch.empty(n, dtype=torch.int64, device=dev)

        def call():          # the C-ABI entry point itself: bitmap / uniforms are inputs prepared once
            L.check(L.lib().mt_sample_tokens(L.ptr(logits), C.c_longlong(v), n, v, C.c_float(1.0), L.ptr(bitmap),
                                             C.c_float(kw.get("repetition_penalty", 1.0)), C.c_float(kw.get("top_p", 1.0)),
                                             int(kw.get("top_k", 0)), int(kw["do_sample"]), L.ptr(u), L.ptr(out),
                                             L.stream_ptr(dev)), "mt_sample_tokens")
        ref = sample_token(logits, prev, uniforms=u, **kw)
        for _ in range(2):
            call()
        assert torch.equal(out, ref.reshape(-1))
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = L.launch_count()
        e0.record()
        for _ in range(args.iters):
            call()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.iters
        nbytes = logits.numel() * 2 + n * 8
        print(json.dumps({"case": name, "rows": n, "vocab": v, "ms": ms, "algorithmic_bytes": nbytes,
                          "achieved_gbs": nbytes / (ms * 1e-3) / 1e9, "peak_gbs": hbm, "frac": nbytes / (ms * 1e-3) / 1e9 / hbm,
                          "launches_per_call": (L.launch_count() - l0) / args.iters, "peak_source": src,
                          "note": "logits larger than L2 (%.0f MB); mt_sample_tokens called through the C ABI, bitmap / uniforms prepared once" % (logits.numel() * 2 / 1e6)}),
              flush=True)
        del logits


def rvq(args):
    """a10: gather-sum over the folded [nq, 1024, 768] fp32 table (folded-table formulation: nq*768*4 B gathered +
    768*4 B written + nq*8 B of codes per frame)."""
    import ctypes as C
    from moss_tts_b200 import _lib as L
    from moss_tts_b200.config import CodecConfig
    from moss_tts_b200.modeling_moss_audio_tokenizer import MossAudioTokenizerModel
    import bench
    hbm, _, src = bench.measured_peaks()
    dev = torch.device("cuda", 0)
    ccfg = CodecConfig.released()
    codec = MossAudioTokenizerModel.random_init(ccfg, device=dev, seed=99, workspace_bytes=1 << 30)
    nq, D = ccfg.num_quantizers, ccfg.code_dim
    for B, T in ((64, 125), (256, 750)):
        codes = torch.randint(0, 1024, (nq, B, T), device=dev)
        out = torch.empty(B * T * D, dtype=torch.float32, device=dev)
        for cm, name in ((0, "token-major [B*T,768] (the decoder's input)"), (1, "channel-major [B,768,T] (decode_codes)")):
            def call():
                L.check(L.lib().mt_rvq_dequant(codec._engine, L.ptr(codes), nq, B, T, cm, L.ptr(out), L.stream_ptr(dev)),
                        "mt_rvq_dequant")
            for _ in range(2):
                call()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(args.iters):
                call()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.iters
            # The gathers hit the 100.7 MB folded table, which stays in the 126 MB L2 (outputs leave with streaming stores):
            # the kernel is bound by L2 GATHER bandwidth, not by HBM.  Reported: the gather rate next to what crosses HBM
            # (the fp32 output + the codes).  No HBM roofline fraction is claimed for the gather bytes.
            gather = B * T * nq * D * 4
            hbm_side = B * T * (D * 4 + nq * 8)
            print(json.dumps({"case": "rvq dequant " + name, "B": B, "T": T, "nq": nq, "ms": ms,
                              "bound": "L2 gather (folded table L2-resident)", "l2_gather_bytes": gather,
                              "l2_gather_gbs": gather / (ms * 1e-3) / 1e9,
                              "hbm_side_bytes": hbm_side, "hbm_side_gbs": hbm_side / (ms * 1e-3) / 1e9,
                              "hbm_peak_gbs": hbm, "hbm_frac": hbm_side / (ms * 1e-3) / 1e9 / hbm, "peak_source": src,
                              "minimal_formulation_bytes_per_frame": 3328,      # codes + one 768-float output
                              "note": "L2 peak is not in MEASURED_PEAKS.json; B300_MICROARCH.md quotes ~20 TB/s aggregate L2 read "
                                      "bandwidth for this generation"}), flush=True)


if __name__ == "__main__":
    main()
