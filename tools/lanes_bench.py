"""The 75-pass suite step as a CUDA graph with 1, 2, 3 or 4 independent chains of passes (SuiteGraph lanes): CUDA-event time
per replay, device-resident CIFAR batch, metric epilogue on.  usage: python tools/lanes_bench.py  ->  gpurun_out/lanes_bench.json"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from imagetransformations_b200 import cifar_c as C


def main():
    n = int(os.environ.get("N", 50000))
    x = torch.from_numpy(np.random.default_rng(1001).integers(0, 256, (n, 32, 32, 3), dtype=np.uint8)).cuda()
    passes = C.suite_passes(None, (1, 2, 3, 4, 5))
    res = {}
    ref = None
    for lanes in (1, 2, 3, 4, 2, 1):
        outs = [torch.empty_like(x) for _ in range(lanes * 2 if lanes != 3 else 3)]
        m = torch.zeros((len(passes), n, 2), dtype=torch.int32, device="cuda")
        g = C.generate_suite(x, seed=100, out=outs, metrics=m, graph=True, lanes=lanes)
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10
        a.record()
        for _ in range(reps):
            g.replay()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / reps
        same = None
        if ref is None:
            ref = m.clone()
        else:
            same = bool(torch.equal(ref, m))     # the metrics of every pass do not depend on how the passes are scheduled
        res.setdefault(str(lanes), []).append(round(ms, 4))
        print(f"lanes {lanes}: {ms:.3f} ms per step, metrics equal to lanes=1: {same}", flush=True)
        del g, outs, m
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/lanes_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
