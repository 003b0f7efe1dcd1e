#!/usr/bin/env python
"""Per-source-line instruction counts and stall samples from an ncu report captured
with --import-source on (binary built with -lineinfo).

    python tools/ncu_lines.py gpurun_out/x.ncu-rep [file-filter] [top-n]
"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    filt = sys.argv[2] if len(sys.argv) > 2 else ""
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    cur, h, agg = None, None, []
    for r in csv.reader(io.StringIO(out)):
        if r and r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if r and r[0] == "Line No":
            h = r
            ix_s, ix_i = h.index("# Samples"), h.index("Instructions Executed")
            continue
        if h is None or len(r) < len(h) or r[0] in ("-", "", "Function Name"):
            continue
        try:
            agg.append((int(r[ix_s] or 0), int(r[ix_i] or 0), cur, r[0], r[1][:100]))
        except ValueError:
            continue
    tot_s, tot_i = sum(a[0] for a in agg), sum(a[1] for a in agg)
    print(f"# {rep}: {tot_i} warp instructions, {tot_s} stall samples")
    byfile = {}
    for a in agg:
        f = byfile.setdefault(a[2], [0, 0])
        f[0] += a[0]
        f[1] += a[1]
    for k, v in sorted(byfile.items(), key=lambda kv: -kv[1][1]):
        print(f"# {k:32s} instr {v[1]:14d} ({v[1] / tot_i:6.3f})  samples {v[0]:9d} ({v[0] / tot_s:6.3f})")
    print("# samples  share        instr  file:line  source")
    for a in sorted([x for x in agg if filt in x[2]], key=lambda x: -x[0])[:top]:
        print(f"{a[0]:8d} {a[0] / tot_s:6.3f} {a[1]:12d}  {a[2]}:{a[3]}  {a[4]}")


if __name__ == "__main__":
    main()
