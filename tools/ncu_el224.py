import sys; sys.path.insert(0,'/root/repo')
import numpy as np, torch
from imagetransformations_b200 import cifar_c as C
x = torch.from_numpy(np.random.default_rng(1).integers(0,256,(1024,224,224,3),dtype=np.uint8)).cuda()
o = torch.empty_like(x)
for sev in (1,2,5):
    for r in range(2): C.corrupt(x,"elastic_transform",sev,seed=r,out=o)
torch.cuda.synchronize(); print("ok")
