#!/usr/bin/env python
"""Summarise .ncu-rep files (ncu -i ... --page raw --csv) into a markdown table for profiles/."""
import csv, io, subprocess, sys

WANT = [("gpu__time_duration.sum", "time_us"), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor_pipe_%"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_%"), ("dram__bytes_read.sum", "dram_rd"),
        ("dram__bytes_write.sum", "dram_wr"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_%"),
        ("launch__registers_per_thread", "regs"), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_%"),
        ("launch__grid_size", "grid"), ("launch__block_size", "block"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_%")]


def rows_of(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rd = list(csv.reader(io.StringIO(out)))
    hdr, units = rd[0], rd[1]
    res = []
    for r in rd[2:]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        item = {"kernel": d.get("Kernel Name", "?")[:70]}
        for k, short in WANT:
            if k in d:
                item[short] = d[k] + (" " + u[k] if u.get(k) and short in ("time_us", "dram_rd", "dram_wr") else "")
        res.append(item)
    return res


if __name__ == "__main__":
    print("| kernel | " + " | ".join(s for _, s in WANT) + " |")
    print("|---|" + "---|" * len(WANT))
    for rep in sys.argv[1:]:
        for it in rows_of(rep):
            print("| `" + it["kernel"] + "` | " + " | ".join(str(it.get(s, "")) for _, s in WANT) + " |")
