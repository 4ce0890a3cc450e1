"""Executed-instruction mix and top stall sites from `ncu --page source` of a .ncu-rep (no GPU needed).
usage: python tools/ncu_source.py file.ncu-rep [top_n]"""
import collections
import csv
import io
import subprocess
import sys


def main():
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hi = next(i for i, r in enumerate(rows) if "Source" in r)
    hdr = rows[hi]
    i_src, i_ex, i_thr, i_smp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    ops, tot, thr, samples = collections.Counter(), 0, 0, []
    for r in rows[hi + 1:]:
        try:
            n = int(r[i_ex])
        except (ValueError, IndexError):
            continue
        toks = r[i_src].split()
        op = toks[1] if toks[0].startswith("@") else toks[0]
        ops[op.split(".")[0]] += n
        tot += n
        thr += int(r[i_thr])
        samples.append((int(r[i_smp]), n, r[i_src].strip()))
    print(f"warp instructions {tot}, thread instructions {thr}, avg active lanes {thr / max(tot, 1):.1f}")
    for k, v in ops.most_common(22):
        print(f"  {k:10s} {v:12d} {v / tot:6.1%}")
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 15
    ssum = sum(s for s, _, _ in samples)
    print("top stall-sample sites:")
    for s, n, src in sorted(samples, reverse=True)[:top]:
        print(f"  {s / ssum:6.1%} execs={n:9d}  {src[:90]}")


if __name__ == "__main__":
    main()
