#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box): key raw metrics + the most-stalled SASS lines.
usage: tools/ncu_summary.py gpurun_out/prof.ncu-rep [n_top] [kernel source file ...]  > profiles/<name>.txt
Source files named after n_top are recorded as `source_blob: <path> <git blob hash>` lines: bench.py only takes
roofline.traffic from a summary whose blob hash equals the kernel source it is running."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 30
for path in sys.argv[3:]:
    blob = subprocess.run(["git", "hash-object", path], capture_output=True, text=True).stdout.strip()
    print(f"source_blob: {path} {blob}")
KEYS = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum ", "dram__bytes_write.sum ",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum ", "sm__warps_active.avg.per_cycle_active", "sm__inst_executed.sum.per_cycle_elapsed",
        "sm__cycles_elapsed.max ", "launch__registers_per_thread ", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ", "smsp__inst_executed.sum ", "sm__inst_executed_pipe_alu.sum ",
        "sm__inst_executed_pipe_fma.sum ", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_active.avg "]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, u = rows[0], rows[1]
for v in rows[2:]:
    print("== kernel:", v[h.index("Kernel Name")][:100])
    for a, b, c in zip(h, u, v):
        if any((a + " ").startswith(k) or a == k.strip() for k in KEYS):
            print(f"  {a:75s} {c:>18s} {b}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if "Address" in r and "Source" in r)
h = rows[hi]
isrc, isamp, iex = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
data = [r for r in rows[hi + 1:] if len(r) > isamp and r[isamp].isdigit()]
tot = sum(int(r[isamp]) for r in data)
stall_cols = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
print(f"== source page: {len(data)} SASS instructions, {tot} warp samples; stall totals:")
agg = sorted(((sum(int(r[i]) for r in data if r[i].isdigit()), h[i]) for i in stall_cols), reverse=True)
print("  " + ", ".join(f"{n}={100.0 * v / max(tot, 1):.1f}%" for v, n in agg[:10]))
print(f"== top {ntop} instructions by samples")
for r in sorted(data, key=lambda r: -int(r[isamp]))[:ntop]:
    st = sorted(((int(r[i]) if r[i].isdigit() else 0, h[i]) for i in stall_cols), reverse=True)[:2]
    print(f"  {r[isamp]:>7s} {r[iex]:>10s}  {r[isrc][:80]:80s} {st[0][1]}:{st[0][0]} {st[1][1]}:{st[1][0]}")
