#!/usr/bin/env python
"""Per-CUDA-source-line summary of an ncu report (needs -lineinfo + --import-source on), one table per captured launch.

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top_n] [launch_index ...]

Prints, per source line, the share of warp-stall samples and of executed warp instructions, and the mean number of
active threads of the instructions attributed to the line.
"""
import csv
import subprocess
import sys


def launches(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    cur = None
    res = []
    h = None
    line = None
    fname = ""
    pending_file = ""
    for r in csv.reader(out.splitlines()):
        if len(r) == 2 and r[0] in ("File Path", "File Name"):
            fname = r[1].split("/")[-1]
            pending_file = r[1]
            continue
        if len(r) == 2 and r[0] == "Function Name":
            # a launch lists each of its source files once: a repeated file (or another function) starts the next launch
            if cur is None or cur["name"] != r[1] or pending_file in cur["files"]:
                cur = {"name": r[1], "lines": {}, "files": set()}
                res.append(cur)
            cur["files"].add(pending_file)
            continue
        if len(r) > 8 and r[0] == "Line No":
            h = r
            si, ie, te = h.index("# Samples"), h.index("Instructions Executed"), h.index("Thread Instructions Executed")
            continue
        if h is None or cur is None or len(r) <= te:
            continue
        if r[0] != "":
            line = (fname, int(r[0]))
            cur["lines"].setdefault(line, [0, 0, 0, r[1].strip()[:120]])
        elif line is not None:
            try:
                v = cur["lines"][line]
                v[0] += int(r[si]); v[1] += int(r[ie]); v[2] += int(r[te])
            except (ValueError, TypeError, KeyError):
                pass
    return res


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    pick = [int(x) for x in sys.argv[3:]]
    for i, L in enumerate(launches(rep)):
        if pick and i not in pick:
            continue
        lines = L["lines"]
        ts = sum(v[0] for v in lines.values()) or 1
        ti = sum(v[1] for v in lines.values()) or 1
        tt = sum(v[2] for v in lines.values()) or 1
        print("== launch %d: %s" % (i, L["name"][:100]))
        print("total stall samples %d, warp instructions %d, thread instructions %d (%.1f threads/inst)" % (ts, ti, tt, tt / ti))
        print("%7s %7s %5s  %s" % ("stall%", "inst%", "thr", "file:line source"))
        for (f, ln), v in sorted(lines.items(), key=lambda kv: -(kv[1][0] / ts + kv[1][1] / ti))[:top]:
            print("%6.2f%% %6.2f%% %5.1f  %s:%d %s" % (100.0 * v[0] / ts, 100.0 * v[1] / ti, v[2] / max(v[1], 1), f, ln, v[3]))


if __name__ == "__main__":
    main()
