"""One launch of each kernel added late in round 2, for an `ncu --set full` capture (development tool)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from imagetransformations_b200 import batch as B, cifar_c as C, cifar_image_transformations as M

rng = np.random.default_rng(1)
x = torch.from_numpy(rng.integers(0, 256, (50000, 32, 32, 3), dtype=np.uint8)).cuda()
o = torch.empty_like(x)
m = torch.zeros((50000, 2), dtype=torch.int32, device="cuda")
for r in range(2):
    C.corrupt(x, "pixelate", 3, out=o, metrics=m)            # OpBoxPixelateMma
    C.corrupt(x, "elastic_transform", 3, seed=r, out=o, metrics=m)
    C.corrupt(x, "jpeg_compression", 3, out=o, metrics=m)
    C.corrupt(x, "frost", 3, seed=r, out=o, metrics=m)
x1 = x[:10000].contiguous()
o1 = torch.empty_like(x1)
for r in range(2):
    B.apply_blur(x1, 2.0, out=o1)                            # OpCvGaussMma
    B.apply_scale(x1, 1.1, out=o1)                           # OpResampleMma
xg = B.to_tensor(x1, M.CIFAR10_MEAN, M.CIFAR10_STD)
base = [("gaussian_noise", {"severity": 3}), ("impulse_noise", {"severity": 3}), ("shot_noise", {"severity": 3}),
        ("defocus_blur", {"severity": 3}), ("enhance_contrast", {"factor": 1.2})]
ops = np.zeros((10000, 5), M.POOL_OP_DTYPE)
ops[:] = M.encode_chain(base)
for r in range(2):
    M.pool_chain_batch(xg, ops, seed=r)                      # pool_chain_kernel
xl = torch.from_numpy(rng.integers(0, 256, (512, 224, 224, 3), dtype=np.uint8)).cuda()
ol = torch.empty_like(xl)
for name, sev in (("elastic_transform", 2), ("zoom_blur", 5), ("glass_blur", 4), ("pixelate", 3), ("fog", 3)):
    for r in range(2):
        C.corrupt(xl, name, sev, seed=r, out=ol)
torch.cuda.synchronize()
print("ok")
