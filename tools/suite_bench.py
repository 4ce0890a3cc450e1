"""Per-pass device timing of the CIFAR-10-C-style suite (development tool): CUDA events, rotating output buffers.
usage: python tools/suite_bench.py [corruption ...]   ->  gpurun_out/suite_bench.json"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from imagetransformations_b200 import cifar_c as C


def main():
    names = sys.argv[1:] or C.CORRUPTIONS
    n = int(os.environ.get("N", 50000))
    side = int(os.environ.get("SIDE", 32))
    x = torch.from_numpy(np.random.default_rng(1001).integers(0, 256, (n, side, side, 3), dtype=np.uint8)).cuda()
    outs = [torch.empty_like(x) for _ in range(3)]
    peak = 6551.7
    mt = torch.zeros((n, 2), dtype=torch.int32, device="cuda") if os.environ.get("METRICS") else None
    res = {}
    for name in names:
        for sev in range(1, 6):
            for r in range(3):
                C.corrupt(x, name, sev, seed=r, out=outs[r % 3], metrics=mt)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = int(os.environ.get("REPS", 10))
            a.record()
            for r in range(reps):
                C.corrupt(x, name, sev, seed=r, out=outs[r % 3], metrics=mt)
            b.record()
            torch.cuda.synchronize()
            us = a.elapsed_time(b) / reps * 1e3
            gbs = 2 * x.numel() / (us * 1e-6) / 1e9
            res[f"{name}/{sev}"] = {"us": round(us, 1), "GBs": round(gbs, 1), "frac": round(gbs / peak, 3)}
            print(f"{name:18s} s{sev}  {us:9.1f} us  {gbs:8.1f} GB/s  {gbs / peak:6.1%}", flush=True)
    tot = sum(v["us"] for v in res.values())
    print(f"total {tot / 1e3:.2f} ms for {len(res)} passes; floor at {peak} GB/s = {len(res) * 2 * x.numel() / peak / 1e3:.1f} us")
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/suite_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
