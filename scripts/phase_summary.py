"""Splits an ncu launch list of scripts/profile_path.py (or a bench.py window) into phases and prints one table per
phase (markdown).  Phases: prefill = first embed_sum .. second embed_sum, decode step k = k-th .. (k+1)-th embed_sum,
codec = last rvq_dequant .. end.   usage: phase_summary.py launches.csv [--step 4]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import summarize_launches as S  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--step", type=int, default=4)
    a = ap.parse_args()
    rows = S.load(a.csv)
    emb = [i for i, r in enumerate(rows) if "embed_sum_kernel" in r[1]]
    rvq = [i for i, r in enumerate(rows) if "rvq_dequant_kernel" in r[1]]
    if len(emb) >= 2:
        S.summarize(rows[emb[0]:emb[1]], "prefill + first sampling")
    if len(emb) > a.step + 1:
        S.summarize(rows[emb[a.step]:emb[a.step + 1]], f"decode step {a.step} (graph replay)")
    elif len(emb) >= 2 and not rvq:
        S.summarize(rows[emb[-2]:emb[-1]], "one decode step inside the window")
    if rvq:
        S.summarize([r for r in rows[rvq[-1]:] if "at::" not in r[1]], "codec decode (64 utterances x 125 frames, chunk_duration=8)")


if __name__ == "__main__":
    main()
