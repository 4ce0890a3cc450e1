"""Run one batch op a few times for an ncu capture (development tool).  usage: ncu_one.py blur|scale"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from imagetransformations_b200 import batch as B

x = torch.from_numpy(np.random.default_rng(1000).integers(0, 256, (10000, 32, 32, 3), dtype=np.uint8)).cuda()
op = sys.argv[1] if len(sys.argv) > 1 else "blur"
for _ in range(3):
    if op == "blur":
        B.apply_blur(x, 2.0)
    elif op == "scale":
        B.apply_scale(x, 1.1)
torch.cuda.synchronize()
