"""Per-op device timing of the C1 grid (development tool): 10000 x 32x32x3, rotating outputs > L2.
usage: python tools/c1_bench.py [op ...]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from imagetransformations_b200 import batch as B, _lib as L


def main():
    ops = sys.argv[1:] or ["rotate", "scale", "brightness", "contrast", "blur", "noise"]
    n = int(os.environ.get("N", 10000))
    x = torch.from_numpy(np.random.default_rng(5).integers(0, 256, (n, 32, 32, 3), dtype=np.uint8)).cuda()
    outs = [torch.empty_like(x) for _ in range(8)]
    peak = 6557.8
    grid = {"rotate": [-22.5, -12.5, 7.5, 22.5], "scale": [0.9, 1.1, 1.2000000000000002, 1.3], "brightness": [-0.05, 0.03],
            "contrast": [0.3, 0.9], "blur": [0.5, 1.0, 2.0, 3.0, 4.0, 5.0], "noise": [0.01, 0.05, 0.1]}

    def run(op, v, out, seed=0):
        if op == "rotate":
            return B.rotate(x, -v, out=out)
        if op == "scale":
            return B.apply_scale(x, v, out=out)
        if op == "brightness":
            return B.brightness(x, 1.0 + v, out=out)
        if op == "contrast":
            return B.convert_scale_abs(x, v, out=out)
        if op == "blur":
            return B.apply_blur(x, v, out=out)
        return B.noise_philox(x, L.NOISE_GAUSS_F32, v * 255.0, seed, out=out)

    for op in ops:
        for v in grid[op]:
            for r in range(8):
                run(op, v, outs[r])
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 40
            a.record()
            for r in range(reps):
                run(op, v, outs[r % 8], r)
            b.record()
            torch.cuda.synchronize()
            us = a.elapsed_time(b) / reps * 1e3
            print(f"{op:10s} {v:8.3f}  {us:8.1f} us  {2 * x.numel() / (us * 1e-6) / 1e9 / peak:6.1%}", flush=True)


main()
