"""Per-source-line instruction counts and stall samples of one kernel from an .ncu-rep captured with --import-source on.
Usage: python tools/ncu_lines.py report.ncu-rep kernel_regex [top]"""
import csv
import io
import os
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
must = sys.argv[4] if len(sys.argv) > 4 else ""          # substring the function name must contain (template arguments)
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                      "regex:" + kern], capture_output=True, text=True).stdout
rows, fname, hdr, on = [], "", None, True
for r in csv.reader(io.StringIO(out)):
    if not r:
        continue
    if r[0] == "File Path":
        fname = os.path.basename(r[1]); continue
    if r[0] == "Function Name":
        on = must in r[1]
        continue
    if r[0] == "Line No":
        hdr = r; continue
    if hdr is None or not on or not r[0].isdigit():
        continue
    try:
        smp = int(r[hdr.index("# Samples")]); inst = int(r[hdr.index("Instructions Executed")])
    except (ValueError, IndexError):
        continue
    rows.append((inst, smp, fname, r[0], r[1].strip()[:110]))
ti, ts = sum(x[0] for x in rows) or 1, sum(x[1] for x in rows) or 1
print("warp instructions %d, samples %d" % (ti, ts))
print("-- by instructions executed")
for inst, smp, fn, ln, src in sorted(rows, reverse=True)[:top]:
    print("  %5.1f%% inst  %5.1f%% smp  %s:%s  %s" % (100.0 * inst / ti, 100.0 * smp / ts, fn, ln, src))
print("-- by stall samples")
for inst, smp, fn, ln, src in sorted(rows, key=lambda x: -x[1])[:top]:
    print("  %5.1f%% smp  %5.1f%% inst  %s:%s  %s" % (100.0 * smp / ts, 100.0 * inst / ti, fn, ln, src))
