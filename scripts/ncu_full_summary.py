"""Summarise an `ncu --set full` report (.ncu-rep) as a markdown table: one column per captured launch.
usage: ncu_full_summary.py report.ncu-rep [title]"""
import csv, io, subprocess, sys

WANT = ["Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]

def main():
    rep = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 else rep
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    names = [r[hdr.index("Kernel Name")][:60] for r in data]
    print(f"# {title}\n")
    print("| metric | " + " | ".join(f"launch {i}" for i in range(len(data))) + " |")
    print("|---|" + "---:|" * len(data))
    print("| kernel | " + " | ".join(f"`{n}`" for n in names) + " |")
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            print(f"| {w} [{units[i]}] | " + " | ".join(r[i] for r in data) + " |")

if __name__ == "__main__":
    main()
