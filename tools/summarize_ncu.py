#!/usr/bin/env python
"""Turn gpurun_out/ ncu artefacts into the small text summaries committed under
profiles/ (the .ncu-rep itself stays in gpurun_out/, which is scratch).

    python tools/summarize_ncu.py launches gpurun_out/launches_r1.csv > profiles/launches_r1.txt
    python tools/summarize_ncu.py kernel   gpurun_out/score_r1.ncu-rep > profiles/score_items_r1.txt
"""
import collections
import csv
import io
import subprocess
import sys

KEY = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__cycles_active.avg",
    "sm__inst_executed.avg.per_cycle_active", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum",
    "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum",
]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    h = rows[0]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        a = agg.setdefault(r[ki].split("(")[0], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"# {path}: per-kernel device time (ncu gpu__time_duration.sum, cold-cache, serialised)")
    print(f"# {'kernel':58s} {'launches':>8s} {'total ms':>10s} {'avg us':>10s} {'share':>7s}")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k[:60]:60s} {v[0]:8d} {v[1] / 1e6:10.3f} {v[1] / v[0] / 1e3:10.1f} {v[1] / tot:7.3f}")


def kernel(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, units, vals = rows[0], rows[1], rows[2]
    print(f"# {path}: ncu --set full, kernel {vals[h.index('Kernel Name')]}")
    for i, name in enumerate(h):
        if name in KEY:
            print(f"{name:75s} {vals[i]:>18s} {units[i]}")
    st = []
    for i, name in enumerate(h):
        if "issue_stalled" in name and name.endswith("per_issue_active.ratio") and "not_issued" not in name:
            try:
                st.append((float(vals[i]), name))
            except ValueError:
                pass
    print("# warp stall reasons (warps per issue-active cycle), top 8")
    for v, n in sorted(st, reverse=True)[:8]:
        print(f"{n:75s} {v:18.3f}")
    src = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    if len(rows) > 2:
        h = rows[1]
        ix = {n: i for i, n in enumerate(h)}
        agg = collections.defaultdict(lambda: [0, 0])
        for r in rows[2:]:
            s = r[ix["Source"]].split()
            if not s:
                continue
            op = (s[1] if s[0].startswith("@") else s[0]).split(".")[0].rstrip(";")
            agg[op][0] += int(r[ix["Instructions Executed"]] or 0)
            agg[op][1] += int(r[ix["# Samples"]] or 0)
        tot = sum(v[0] for v in agg.values())
        print("# executed warp instructions by opcode (source page), top 10")
        for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:10]:
            print(f"{k:12s} {v[0]:14d} {v[0] / tot:7.3f}  stall samples {v[1]}")


def multi(path):
    """One line per captured launch: duration, DRAM bytes and throughput, occupancy."""
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h = rows[0]
    ix = {n: i for i, n in enumerate(h)}
    cols = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "dram__bytes.sum.per_second", "FBSP.TriageCompute.dram__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
            "launch__grid_size", "launch__block_size"]
    cols = [c for c in cols if c in ix]
    print(f"# {path}: ncu --set full, one row per captured launch (units: {', '.join(rows[1][ix[c]] or '-' for c in cols)})")
    print("# kernel | " + " | ".join(cols))
    for r in rows[2:]:
        if len(r) < len(h):
            continue
        print(r[ix["Kernel Name"]].split("(")[0][:40] + " | " + " | ".join(r[ix[c]] for c in cols))


if __name__ == "__main__":
    {"launches": launches, "kernel": kernel, "multi": multi}[sys.argv[1]](sys.argv[2])
