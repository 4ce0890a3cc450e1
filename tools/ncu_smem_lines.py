"""Per-source-line shared-memory wavefronts, instructions and stall samples from a .ncu-rep captured with --import-source on.
usage: python tools/ncu_smem_lines.py file.ncu-rep launch_index [top_n]"""
import csv
import subprocess
import sys


def main():
    rep, idx = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", idx,
                          "--launch-count", "1"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, fname, lines = None, "", []

    def f(x):
        try:
            return float(x.replace(",", ""))
        except ValueError:
            return 0.0
    for r in rows:
        if len(r) >= 2 and r[0] in ("File Path", "File Name"):
            fname = r[1].split("/")[-1]
            continue
        if len(r) > 5 and r[0] == "Line No":
            hdr = r
            continue
        if hdr is None or len(r) < len(hdr) or not r[0].isdigit():
            continue
        iw, ii, ie, isamp = (hdr.index(k) for k in ("L1 Wavefronts Shared", "L1 Wavefronts Shared Ideal", "Instructions Executed", "# Samples"))
        lines.append((f(r[iw]), f(r[ii]), f(r[ie]), f(r[isamp]), f"{fname}:{r[0]}", r[1].strip()[:100]))
    tw, ti, ts = sum(l[0] for l in lines), sum(l[2] for l in lines), sum(l[3] for l in lines)
    print(f"shared wavefronts {tw / 1e6:.1f}M (ideal {sum(l[1] for l in lines) / 1e6:.1f}M), warp instructions {ti / 1e6:.1f}M, samples {ts:.0f}")
    key = 0 if tw > 0 else 2
    for l in sorted(lines, key=lambda l: -l[key])[:top]:
        print(f"{100 * l[0] / max(tw, 1):5.1f}% wf ({l[0] / 1e6:5.1f}M, ideal {l[1] / 1e6:5.1f}M) {100 * l[2] / max(ti, 1):5.1f}% inst {100 * l[3] / max(ts, 1):5.1f}% smp  {l[4]:24s} {l[5]}")


main()
