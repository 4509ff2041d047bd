#!/usr/bin/env python
"""Summarise .ncu-rep captures (taken on the GPU box) into profiles/ncu_*.md on the CPU box.

    python benchmarks/ncu_summary.py gpurun_out/prof_gemm2sm.ncu-rep [...]
"""
import csv
import io
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
    "smsp__cycles_active.avg", "sm__cycles_elapsed.max", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
    "smsp__warp_issue_stalled_barrier_per_warp_active.pct", "smsp__warp_issue_stalled_wait_per_warp_active.pct",
]


def main():
    peaks = {"hbm_gbs": 6551.4}
    for rep in sys.argv[1:]:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        if len(rows) < 3:
            print("no data in", rep)
            continue
        hdr, units = rows[0], rows[1]
        name_i = hdr.index("Kernel Name")
        out = [f"# ncu --set full summary: {os.path.basename(rep)}", "",
               "Captured on a B200 with `ncu --set full --clock-control none --import-source on` after the same command exited 0",
               "without ncu (benchmarks/ncu_target.py).  Durations under ncu are for shares only, never bench values.", ""]
        for r in rows[2:]:
            out.append(f"## {r[name_i][:110]}")
            vals = {}
            for k in KEYS:
                cands = [i for i, h in enumerate(hdr) if h == k or h.startswith(k)]
                if cands:
                    i = cands[0]
                    vals[k] = (r[i], units[i])
                    out.append(f"- {hdr[i]}: {r[i]} {units[i]}")
            try:
                dur_ns = float(vals["gpu__time_duration.sum"][0].replace(",", ""))
                unit = vals["gpu__time_duration.sum"][1]
                dur_s = dur_ns * {"ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "nsecond": 1e-9, "ms": 1e-3, "msecond": 1e-3}.get(unit, 1e-9)
                def tobytes(k):
                    v, u = vals[k]
                    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
                traffic = tobytes("dram__bytes_read.sum") + tobytes("dram__bytes_write.sum")
                out.append(f"- derived: DRAM traffic {traffic/1e6:.1f} MB in {dur_s*1e6:.1f} us = {traffic/dur_s/1e9:.0f} GB/s "
                           f"= {traffic/dur_s/1e9/peaks['hbm_gbs']:.2f} of measured copy bandwidth ({peaks['hbm_gbs']} GB/s)")
            except Exception:
                pass
            out.append("")
        dst = os.path.join(ROOT, "profiles", "ncu_" + os.path.basename(rep).replace(".ncu-rep", ".md"))
        with open(dst, "w") as f:
            f.write("\n".join(out))
        print("wrote", dst, len(rows) - 2, "kernels")


if __name__ == "__main__":
    main()
