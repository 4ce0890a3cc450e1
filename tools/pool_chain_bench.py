"""Per-op cost inside the fused TransformationPool chain (development tool): 10000 individuals, one op each."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from imagetransformations_b200 import batch as B, cifar_image_transformations as M

n = 10000
x = B.to_tensor(torch.from_numpy(np.random.default_rng(1).integers(0, 256, (n, 32, 32, 3), dtype=np.uint8)).cuda(), M.CIFAR10_MEAN, M.CIFAR10_STD)
cases = [[], [("gaussian_noise", {"severity": 3})], [("impulse_noise", {"severity": 3})], [("shot_noise", {"severity": 3})],
         [("shot_noise", {"severity": 1})], [("defocus_blur", {"severity": 3})], [("enhance_contrast", {"factor": 1.3})],
         [("enhance_brightness", {"factor": 1.3})], [("enhance_sharpness", {"factor": 1.3})], [("enhance_color", {"factor": 1.3})],
         [("motion_blur", {"size": 9})], [("histogram_equalization", {})]]
for c in cases:
    ops = np.zeros((n, max(1, len(c))), M.POOL_OP_DTYPE)
    ops["kind"] = -1
    if c:
        ops[:, :len(c)] = M.encode_chain(c)
    for _ in range(3):
        M.pool_chain_batch(x, ops, seed=1)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        M.pool_chain_batch(x, ops, seed=1)
    b.record()
    torch.cuda.synchronize()
    print(f"{(c[0][0] + str(c[0][1])) if c else 'empty chain':45s} {a.elapsed_time(b) / 10 * 1e3:8.1f} us", flush=True)
