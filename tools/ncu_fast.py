"""Run one fast-mode pass of the named (corruption, severity) pairs for an ncu capture (development tool).
usage: ncu_fast.py elastic_transform:3 zoom_blur:5 ..."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from imagetransformations_b200 import cifar_c as C

n = int(os.environ.get("N", 50000))
x = torch.from_numpy(np.random.default_rng(1001).integers(0, 256, (n, 32, 32, 3), dtype=np.uint8)).cuda()
out = torch.empty_like(x)
for a in sys.argv[1:]:
    name, sev = a.split(":")
    for r in range(2):
        C.corrupt(x, name, int(sev), seed=r, out=out)
torch.cuda.synchronize()
print("ok")
