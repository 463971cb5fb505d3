#!/usr/bin/env python
"""Per-source-line instruction / stall-sample totals of one kernel launch in an .ncu-rep (needs --import-source on,
-lineinfo).   usage: source_hot.py <file.ncu-rep> <launch index in file> [top N]"""
import csv, subprocess, sys, io
rep, idx = sys.argv[1], int(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                      "--launch-skip", str(idx), "--launch-count", "1"], capture_output=True, text=True).stdout
lines = raw.splitlines()
src, fname, h = [], "?", None
for l in lines:
    if l.startswith('"File Path"'):
        fname = l.split(",", 1)[1].strip('"').split("/")[-1]
        continue
    if l.startswith('"Function Name"'):
        print(l[:150]); continue
    try:
        r = next(csv.reader([l]))
    except Exception:
        continue
    if l.startswith('"Line No"'):
        h = r; ci, cs = h.index("Instructions Executed"), h.index("Warp Stall Sampling (All Samples)"); continue
    if h is None or not r or not r[0].strip().isdigit():
        continue
    try:
        src.append((fname, int(r[0]), r[1].strip(), int(r[ci]), int(r[cs])))
    except (ValueError, IndexError):
        pass
ti, ts = sum(x[3] for x in src), sum(x[4] for x in src)
print(f"total warp instructions {ti}, stall samples {ts}")
print("  file:line   inst%  stall%  source")
for fn, ln, text, ins, st in sorted(src, key=lambda x: -x[3 if "--by-inst" in sys.argv else 4])[:top]:
    print(f"{fn[:14]:>14}:{ln:<5d} {100 * ins / max(ti, 1):5.1f}  {100 * st / max(ts, 1):5.1f}   {text[:105]}")
