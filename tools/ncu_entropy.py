"""One launch of the Shannon-entropy gate on 10000 CIFAR-shaped tensors for ncu captures (development tool)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from imagetransformations_b200 import batch as B, tta_initial

x = torch.from_numpy(np.random.default_rng(1000).integers(0, 256, (10000, 32, 32, 3), dtype=np.uint8)).cuda()
tn = B.to_tensor(x, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD)
for _ in range(3):
    e = B.shannon_entropy(tn, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD)
torch.cuda.synchronize()
