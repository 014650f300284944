"""Roofline table from the ncu reports of scripts/ncu_all.sh (runs on the CPU box: `ncu -i ... --page raw --csv`).

    python scripts/ncu_table.py > profiles/ncu_kernels_r2.txt

Per kernel: duration, DRAM bytes read + written, achieved DRAM bandwidth against the MEASURED copy bandwidth
(MEASURED_PEAKS.json), tensor-pipe activity, registers, occupancy, and the top stall reason.  Cold-cache single
launches under the profiler: compare shares and fractions, not absolute times (B200_PROFILING.md)."""
import csv
import glob
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}


_SCALE = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6,
          "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "Tbyte": 1e6}


def raw(path):
    """{metric: (value, unit)}; times are normalised to microseconds, sizes to megabytes."""
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    if len(rows) < 3:
        return None
    hdr, units, vals = rows[0], rows[1], rows[-1]
    d = {}
    for h, u, v in zip(hdr, units, vals):
        try:
            x = float(v.replace(",", ""))
        except ValueError:
            continue
        d[h] = x * _SCALE.get(u, 1.0)
    return d


def f(d, key, default=0.0):
    for k, v in d.items():
        if k == key or k.endswith(key):
            return v
    return default


print(f"# ncu --set full, one launch per kernel (2-layer slice of config 3 at full width; flash kernels: reference-as-shipped shape)")
print(f"# roofline denominators (MEASURED_PEAKS.json): HBM {peaks['hbm_gbs']:.0f} GB/s, bf16 {peaks['bf16_tflops']:.0f} TFLOP/s burst")
print("kernel | duration us | DRAM read MB | DRAM write MB | DRAM GB/s (of measured) | tensor pipe active % | regs/thread | achieved occupancy % | dominant stall")
for path in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "ncu", "*.ncu-rep"))):
    d = raw(path)
    name = os.path.basename(path)[:-8]
    if d is None:
        print(f"{name} | (unreadable report)")
        continue
    dur = f(d, "gpu__time_duration.sum")                 # us
    rd = f(d, "dram__bytes_read.sum")                    # MB
    wr = f(d, "dram__bytes_write.sum")
    gbs = (rd + wr) / dur * 1e3 if dur else 0.0          # MB / us = TB/s -> GB/s
    tens = f(d, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")
    regs = f(d, "launch__registers_per_thread")
    occ = f(d, "sm__warps_active.avg.pct_of_peak_sustained_active")
    stalls = {k.split("smsp__average_warps_issue_stalled_")[-1].replace("_per_issue_active.ratio", ""): v
              for k, v in d.items() if "smsp__average_warps_issue_stalled_" in k and k.endswith("per_issue_active.ratio")}
    top = max(stalls, key=stalls.get) if stalls else "-"
    print(f"{name} | {dur:.1f} | {rd:.1f} | {wr:.1f} | {gbs:.0f} ({gbs / peaks['hbm_gbs'] * 100:.0f} %) | {tens:.1f} | {regs:.0f} | {occ:.0f} | {top}")
