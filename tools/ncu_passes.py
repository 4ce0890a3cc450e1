"""One launch of selected suite passes, for `ncu --set full` captures (development tool).
usage: python tools/ncu_passes.py name:sev [name:sev ...]   (N images from env N, default 50000)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from imagetransformations_b200 import cifar_c as C


def main():
    n = int(os.environ.get("N", 50000))
    x = torch.from_numpy(np.random.default_rng(1001).integers(0, 256, (n, 32, 32, 3), dtype=np.uint8)).cuda()
    out = torch.empty_like(x)
    for spec in sys.argv[1:]:
        name, sev = spec.split(":")
        C.corrupt(x, name, int(sev), seed=3, out=out)
    torch.cuda.synchronize()


if __name__ == "__main__":
    main()
