"""One launch each of the kernels added or rewritten late in round 1 (TTA views, AugMix, entropy gate, tensor TTA chain,
dp4a Gaussian blur) for an `ncu --set full` capture (development tool)."""
import os
import random
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from imagetransformations_b200 import AugMix, batch as B, post_hoc_aaryaa as PH, tta_initial, views as V

n = 10000
x = torch.from_numpy(np.random.default_rng(1000).integers(0, 256, (n, 32, 32, 3), dtype=np.uint8)).cuda()
tn = B.to_tensor(x, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD)
np.random.seed(3)
crop, br, ct = tta_initial.sample_view_draws(n)
draws = V.tta16_draws(n, crop, br, ct, "cuda")
out = torch.empty((n, 16, 3, 32, 32), dtype=torch.float32, device="cuda")
V.tta16(tn, None, None, None, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD, out=out, draws=draws)
B.shannon_entropy(tn, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD)
torch.manual_seed(5)
base = [PH.chain_record(PH.draw_chain(32, 32)) for _ in range(256)]
PH.tta_chain_batch(tn, np.stack([base[i % 256] for i in range(n * 10)]), views=10)
B.apply_blur(x, 2.0)
n4 = 592
x4 = torch.rand((n4, 3, 224, 224), dtype=torch.float32, device="cuda")
np.random.seed(4)
random.seed(4)
d = [AugMix.sample_draws(3, 3) for _ in range(n4)]
AugMix.augmix_batch(x4, 3, np.stack([a[0] for a in d]), np.array([a[1] for a in d]), np.stack([a[2] for a in d]), np.stack([a[3] for a in d]))
torch.cuda.synchronize()
