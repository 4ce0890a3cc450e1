"""One launch of the fused 16-view TTA expansion (C3) and of AugMix (C4, N4 images) for ncu captures (development tool)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import random

import numpy as np
import torch

from imagetransformations_b200 import AugMix, batch as B, tta_initial, views as V

n = 10000
x = torch.from_numpy(np.random.default_rng(1000).integers(0, 256, (n, 32, 32, 3), dtype=np.uint8)).cuda()
tn = B.to_tensor(x, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD)
np.random.seed(3)
crop, br, ct = tta_initial.sample_view_draws(n)
out = torch.empty((n, 16, 3, 32, 32), dtype=torch.float32, device="cuda")
V.tta16(tn, crop, br, ct, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD, out=out)
n4 = int(os.environ.get("N4", 592))
x4 = torch.rand((n4, 3, 224, 224), dtype=torch.float32, device="cuda")
np.random.seed(4)
random.seed(4)
d = [AugMix.sample_draws(3, 3) for _ in range(n4)]
AugMix.augmix_batch(x4, 3, np.stack([a[0] for a in d]), np.array([a[1] for a in d]), np.stack([a[2] for a in d]), np.stack([a[3] for a in d]))
torch.cuda.synchronize()
