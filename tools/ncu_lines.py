"""Per-source-line totals (warp instructions executed, stall samples) from a .ncu-rep captured with --import-source on.
usage: python tools/ncu_lines.py file.ncu-rep [top_n] [launch_index]"""
import csv
import io
import subprocess
import sys


def main():
    cmd = ["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass"]
    if len(sys.argv) > 3:
        cmd += ["--launch-skip", sys.argv[3], "--launch-count", "1"]
    out = subprocess.run(cmd, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    fname = ""
    lines = []
    hdr = None
    for r in rows:
        if len(r) >= 2 and r[0] in ("File Path", "File Name"):
            fname = r[1].split("/")[-1]
            continue
        if len(r) > 5 and r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or len(r) < len(hdr) or not r[0].isdigit():
            continue
        i_s, i_e = hdr.index("# Samples"), hdr.index("Instructions Executed")
        try:
            lines.append((int(r[i_s]), int(r[i_e]), f"{fname}:{r[0]}", r[1].strip()))
        except ValueError:
            pass
    ts, te = sum(x[0] for x in lines), sum(x[1] for x in lines)
    print(f"total samples {ts}, warp instructions {te}")
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    for s, e, loc, src in sorted(lines, reverse=True)[:top]:
        print(f"{s / max(ts, 1):6.1%} smp {e / max(te, 1):6.1%} inst  {loc:28s} {src[:100]}")


if __name__ == "__main__":
    main()
