"""In-situ timeline of the decode step (no profiler): every gemm_dec_kernel CTA and every attn_decode_kernel
block records %globaltimer stamps (mt_debug_set_trace); the stamps of the LAST graph replay are summarised.

  python scripts/trace_decode.py [--steps 40] [--batch 64]

GEMM stamps per CTA: 0 entry, 1 setup done (barriers, TMEM), 2 exit, 3 dependency wait returned (producer),
4 first stage landed (first MMA), 5 last MMA committed, 6 last accumulator ready (epilogue warp), 7 epilogue done.
Attention stamps per block: 0 entry, 1 dependency wait returned, 2 q/k/v prepared + appended, 3 scores done,
4 softmax done, 5 P.V done, 6 output written.
"""
from __future__ import annotations

import argparse
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--layer", type=int, default=17)
    args = ap.parse_args()
    from moss_tts_b200 import _lib as L
    from moss_tts_b200.config import LMConfig
    from moss_tts_b200.modeling_moss_tts import MossTTSDelayModel
    import bench

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    cfg = LMConfig.delay_8b()
    B = args.batch
    model = MossTTSDelayModel.random_init(cfg, device=dev, seed=1234, never_stop=True)
    ids, mask = bench.make_prompts(cfg, B, seed=0)
    ids, mask = ids.to(dev), mask.to(dev)
    kw = dict(max_new_tokens=args.steps, text_temperature=0, audio_temperature=0, record_step_times=True)
    model.generate(ids, mask, **kw)                      # warm
    GW, AW, NW = 148 * 8, B * cfg.num_kv_heads * 8, B * 8
    # launch order of a step: add_rmsnorm | per layer: qkv, attention, o, add_rmsnorm, gate|up, (silu), down, add_rmsnorm | heads
    step_words = NW + cfg.num_layers * (4 * GW + AW + 2 * NW) + 2 * GW
    buf = torch.zeros(2 * GW + 2 * step_words + 1024, dtype=torch.int64, device=dev)
    lib = L.lib()
    lib.mt_debug_set_trace.restype = C.c_longlong
    lib.mt_debug_set_trace(C.c_void_p(buf.data_ptr()), C.c_longlong(buf.numel()))
    model.generate(ids, mask, **kw)
    used = lib.mt_debug_set_trace(None, C.c_longlong(0))
    torch.cuda.synchronize()
    ms = model.last_step_ms
    print(f"decode ms/step (median of last 20): {sorted(ms[-20:])[10]:.3f}; trace words used {used} (expected {2 * GW + 2 * step_words})")
    raw = buf[:used].cpu().numpy().astype(np.int64)[-step_words:]      # the captured step = last replay
    gemm, attn, norm = [], [], []
    off = NW                                              # the step's first add_rmsnorm (after the embedding)

    def take(n, rows):
        nonlocal off
        a = raw[off:off + n].reshape(rows, 8)
        off += n
        return a

    for l in range(cfg.num_layers):
        gemm.append(take(GW, 148))                        # qkv
        attn.append(take(AW, -1))
        gemm.append(take(GW, 148))                        # o
        norm.append(take(NW, B))                          # + residual, post-attention norm
        gemm.append(take(GW, 148))                        # gate|up
        gemm.append(take(GW, 148))                        # down
        norm.append(take(NW, B))                          # + residual, next input norm
    names = ["qkv", "o", "gate|up", "down"]
    nbytes = {"qkv": 50.3e6, "o": 33.6e6, "gate|up": 201.3e6, "down": 100.7e6}
    EXIT = 2

    def ev(a, e, fn):
        v = a[a[:, 0] > 0][:, e]
        return float(fn(v))

    rows = {n: [] for n in names}
    arows = []
    for l in range(1, cfg.num_layers):                    # layer 0's qkv follows the embedding, skip it
        for k, n in enumerate(names):
            a = gemm[4 * l + k]
            prev_end = ev(gemm[4 * l + k - 1], EXIT, np.max)
            live = a[a[:, 0] > 0]
            rows[n].append(dict(
                ctas=len(live),
                entry=(ev(a, 0, np.median) - prev_end) / 1e3,
                wait_ret=(ev(a, 3, np.median) - prev_end) / 1e3,
                first_mma=np.median(live[:, 4] - live[:, 3]) / 1e3,
                mainloop=(ev(a, 5, np.max) - ev(a, 4, np.min)) / 1e3,
                acc_ready=np.median(live[:, 6] - live[:, 5]) / 1e3,
                epilogue=np.median(live[:, 7] - live[:, 6]) / 1e3,
                exit=np.median(live[:, EXIT] - live[:, 7]) / 1e3,
                tail=(ev(a, EXIT, np.max) - ev(a, 5, np.max)) / 1e3,
                total=(ev(a, EXIT, np.max) - prev_end) / 1e3))
        at = attn[l]
        q_end = ev(gemm[4 * l], EXIT, np.max)
        arows.append(dict(
            entry_med=(np.median(at[:, 0]) - q_end) / 1e3, entry_max=(at[:, 0].max() - q_end) / 1e3,
            wait_med=(np.median(at[:, 1]) - q_end) / 1e3, wait_max=(at[:, 1].max() - q_end) / 1e3,
            inputs=np.median(at[:, 7] - at[:, 1]) / 1e3, prep=np.median(at[:, 2] - at[:, 1]) / 1e3, scores=np.median(at[:, 3] - at[:, 2]) / 1e3,
            softmax=np.median(at[:, 4] - at[:, 3]) / 1e3, pv=np.median(at[:, 5] - at[:, 4]) / 1e3,
            write=np.median(at[:, 6] - at[:, 5]) / 1e3, block_total=np.median(at[:, 6] - at[:, 1]) / 1e3,
            end_max=(at[:, 6].max() - q_end) / 1e3, end_med=(np.median(at[:, 6]) - q_end) / 1e3))
    keys = ["ctas", "entry", "wait_ret", "first_mma", "mainloop", "acc_ready", "epilogue", "exit", "tail", "total"]
    print("\nGEMMs: mean over layers 1..L-1, microseconds; entry / wait_ret / total are measured from the last CTA exit of the "
          "previous GEMM, so they include the kernels in between:")
    print("| kind | ideal at 6.5 TB/s | " + " | ".join(keys) + " |")
    print("|---|---:|" + "---:|" * len(keys))
    for n in names:
        r = rows[n]
        print(f"| {n} | {nbytes[n] / 6.536e12 * 1e6:.1f} | " + " | ".join(f"{np.mean([x[k] for x in r]):.2f}" for k in keys) + " |")
    layer_total = sum(np.mean([x["total"] for x in rows[n]]) for n in names)
    print(f"\nsum of the four per-layer totals: {layer_total:.1f} us -> x{cfg.num_layers} = {layer_total * cfg.num_layers / 1e3:.2f} ms/step")
    print("\nattention blocks (microseconds; entry/wait/end measured from the last CTA exit of the qkv GEMM):")
    print("| " + " | ".join(arows[0].keys()) + " |")
    print("|" + "---:|" * len(arows[0]))
    print("| " + " | ".join(f"{np.mean([x[k] for x in arows]):.2f}" for k in arows[0]) + " |")
    # add_rmsnorm after the o projection (index 2l) and after the down projection (2l + 1)
    nrows = []
    for l in range(1, cfg.num_layers):
        for which, gi in ((0, 4 * l + 1), (1, 4 * l + 3)):
            nt = norm[2 * l + which]
            g_end = ev(gemm[gi], EXIT, np.max)
            nrows.append(dict(entry_med=(np.median(nt[:, 0]) - g_end) / 1e3, wait_med=(np.median(nt[:, 1]) - g_end) / 1e3,
                              probe1=np.median(nt[:, 5] - nt[:, 1]) / 1e3, probe2=np.median(nt[:, 6] - nt[:, 5]) / 1e3,
                              loads=np.median(nt[:, 2] - nt[:, 1]) / 1e3, block_sum=np.median(nt[:, 3] - nt[:, 2]) / 1e3,
                              normalise=np.median(nt[:, 4] - nt[:, 3]) / 1e3, end_max=(nt[:, 4].max() - g_end) / 1e3))
    print("\nadd_rmsnorm blocks (microseconds; entry/wait/end measured from the last CTA exit of the producing GEMM):")
    print("| " + " | ".join(nrows[0].keys()) + " |")
    print("|" + "---:|" * len(nrows[0]))
    print("| " + " | ".join(f"{np.mean([x[k] for x in nrows]):.2f}" for k in nrows[0]) + " |")


if __name__ == "__main__":
    main()
