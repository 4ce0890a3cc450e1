"""Warp-instruction and stall-sample totals per source line, dumped as 'file:line inst samples' (development tool).
usage: python tools/ncu_ranges.py file.ncu-rep launch_index > lines.txt"""
import csv, io, subprocess, sys
cmd = ["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", sys.argv[2], "--launch-count", "1"]
out = subprocess.run(cmd, capture_output=True, text=True).stdout
fname, hdr = "", None
for r in csv.reader(io.StringIO(out)):
    if len(r) >= 2 and r[0] in ("File Path", "File Name"):
        fname = r[1].split("/")[-1]; continue
    if len(r) > 5 and r[0] == "Line No":
        hdr = r; continue
    if hdr is None or len(r) < len(hdr) or not r[0].isdigit():
        continue
    try:
        e, s = int(r[hdr.index("Instructions Executed")]), int(r[hdr.index("# Samples")])
    except ValueError:
        continue
    if e or s:
        print(f"{fname}:{r[0]} {e} {s} {r[1].strip()[:80]}")
