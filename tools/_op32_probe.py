import torch
from imagetransformations_b200 import cifar_c as C
x = torch.randint(0, 256, (50000, 32, 32, 3), dtype=torch.uint8, device="cuda")
m = torch.zeros((50000, 2), dtype=torch.int32, device="cuda")
for s in (1, 5):
    C.corrupt(x, "zoom_blur", s, seed=1, metrics=m)
torch.cuda.synchronize()
