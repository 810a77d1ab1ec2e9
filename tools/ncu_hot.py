"""Summarise `ncu --page source --csv --print-source cuda,sass` of one kernel: stall samples per CUDA source line and
per stall reason.  Usage: python tools/ncu_hot.py report.ncu-rep kernel_regex [top]"""
import csv
import io
import os
import subprocess
import sys
from collections import Counter


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                          "regex:" + kern], capture_output=True, text=True).stdout
    lines, reasons = [], Counter()
    hdr, fname = None, ""
    for r in csv.reader(io.StringIO(out)):
        if not r:
            continue
        if r[0] == "File Path":
            fname = os.path.basename(r[1])
            continue
        if r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or r[0] == "" or len(r) < 8:
            continue                                  # SASS rows: already summed into their source line
        try:
            smp = int(r[hdr.index("# Samples")])
        except ValueError:
            continue
        st = {}
        for i, k in enumerate(hdr):
            if k.startswith("stall_") and "Not Issued" not in k and i < len(r):
                try:
                    v = int(r[i])
                except ValueError:
                    v = 0
                if v:
                    st[k[6:]] = v
                    reasons[k[6:]] += v
        inst = r[hdr.index("Instructions Executed")]
        lines.append((smp, fname, r[0], r[1].strip()[:90], st, inst))
    tot = sum(x[0] for x in lines) or 1
    print("total samples", tot)
    print("-- stall reasons")
    for k, v in reasons.most_common(10):
        print("  %-20s %7d %5.1f%%" % (k, v, 100.0 * v / tot))
    print("-- top source lines")
    for smp, fn, ln, src, st, inst in sorted(lines, reverse=True)[:top]:
        top2 = ",".join("%s=%d" % kv for kv in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print("  %6d %5.1f%% inst=%-9s %s:%s  %s   [%s]" % (smp, 100.0 * smp / tot, inst, fn, ln, src, top2))


if __name__ == "__main__":
    main()
