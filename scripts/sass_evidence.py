"""Per-kernel SASS evidence: which kernels of libmosstts_sm100.so use the 5th-generation tensor cores (UTC*MMA), tensor memory
(LDTM / STTM), TMA (UTMALDG / UTMAPF) and which still run on legacy mma.sync (HMMA.16816 / HMMA.1688).

  python scripts/sass_evidence.py > profiles/r02_sass_tensor_kernels.md      (needs cuobjdump; no GPU)
"""
from __future__ import annotations

import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "moss_tts_b200", "libmosstts_sm100.so")
PAT = collections.OrderedDict([
    ("UTCHMMA (tcgen05.mma f16/bf16/tf32)", re.compile(r"\bUTC[A-Z]*MMA")),
    ("LDTM (tcgen05.ld)", re.compile(r"\bLDTM")),
    ("STTM (tcgen05.st)", re.compile(r"\bSTTM")),
    ("UTMALDG (TMA load)", re.compile(r"\bUTMALDG")),
    ("UTMAPF / UBLKPF (TMA / bulk prefetch)", re.compile(r"\bUTMAPF|\bUBLKPF")),
    ("HMMA (mma.sync)", re.compile(r"\bHMMA")),
    ("MUFU.EX2", re.compile(r"\bMUFU\.EX2")),
])


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], check=True, capture_output=True, text=True).stdout
    demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
    counts, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            counts[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        for k, p in PAT.items():
            if p.search(line):
                counts[cur][k] += 1
    print("# SASS evidence per kernel of libmosstts_sm100.so (sm_100a)\n")
    print("`cuobjdump -sass moss_tts_b200/libmosstts_sm100.so`, instruction counts per kernel symbol "
          "(`scripts/sass_evidence.py`).  Kernels with none of these instructions are omitted.\n")
    keys = list(PAT)
    print("| kernel | " + " | ".join(keys) + " |")
    print("|---|" + "---:|" * len(keys))
    for fn, c in counts.items():
        if not any(c[k] for k in keys[:6]):
            continue
        name = demangle(fn)
        name = re.sub(r"\((int|bool)\)", "", name).replace("mt::(anonymous namespace)::", "").replace("mt::<unnamed>::", "")
        name = re.sub(r">\(.*", ">", name) if ">(" in name else re.sub(r"\(.*", "", name)
        name = name.replace("void ", "")
        print(f"| `{name[:90]}` | " + " | ".join(str(c[k]) if c[k] else "" for k in keys) + " |")
    tc5 = [fn for fn, c in counts.items() if c[keys[0]]]
    legacy = [fn for fn, c in counts.items() if c[keys[5]] and not c[keys[0]]]
    print(f"\n{len(tc5)} kernel instantiations issue tcgen05 MMAs; {len(legacy)} use mma.sync only "
          "(decode / prefill LM attention, the TF32 and ring-cache codec attention).")


if __name__ == "__main__":
    sys.exit(main())
