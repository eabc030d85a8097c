#!/usr/bin/env python
"""Per-CUDA-source-line summary of an ncu report (needs -lineinfo + --import-source on).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top_n]

Prints, per source line, the share of warp-stall samples and of executed warp instructions.
"""
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = None
    lines = {}
    cur = None
    fname = ""
    for r in rows:
        if len(r) == 2 and r[0] in ("File Path", "File Name"):
            fname = r[1].split("/")[-1]
            continue
        if len(r) > 8 and r[0] == "Line No":
            h = r
            si, ie, te = h.index("# Samples"), h.index("Instructions Executed"), h.index("Thread Instructions Executed")
            continue
        if h is None or len(r) <= te:
            continue
        if r[0] != "":
            cur = (fname, int(r[0]))
            lines.setdefault(cur, [0, 0, 0, r[1].strip()[:110]])
        else:
            try:
                lines[cur][0] += int(r[si]); lines[cur][1] += int(r[ie]); lines[cur][2] += int(r[te])
            except (ValueError, TypeError, KeyError):
                pass
    ts = sum(v[0] for v in lines.values()) or 1
    ti = sum(v[1] for v in lines.values()) or 1
    tt = sum(v[2] for v in lines.values()) or 1
    print("total stall samples %d, warp instructions %d, thread instructions %d (%.1f threads/inst)" % (ts, ti, tt, tt / ti))
    print("%7s %7s  %s" % ("stall%", "inst%", "file:line source"))
    for (f, ln), v in sorted(lines.items(), key=lambda kv: -kv[1][0])[:top]:
        print("%6.2f%% %6.2f%%  %s:%d %s" % (100.0 * v[0] / ts, 100.0 * v[1] / ti, f, ln, v[3]))


if __name__ == "__main__":
    main()
