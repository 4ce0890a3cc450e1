import sys, time
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from imagetransformations_b200 import cifar_c as C
from oracle import cifar_c as CC
x = np.random.default_rng(3).integers(0, 256, (2, 224, 224, 3), dtype=np.uint8)
xd = torch.from_numpy(x).cuda()
for name in C.CORRUPTIONS:
    for sev in (1, 4):
        msg = []
        try:
            y = C.corrupt(xd, name, sev, seed=1); torch.cuda.synchronize(); msg.append("onchip ok")
        except Exception as e:
            msg.append("onchip ERR " + str(e)[:80])
        try:
            t0 = time.time()
            d = CC.draw(name, sev, x, np.random.RandomState(sev))
            want = CC.corrupt(name, sev, x, d)
            t1 = time.time() - t0
            t0 = time.time()
            got = C.corrupt(xd, name, sev, draws=d).cpu().numpy()
            t2 = time.time() - t0
            msg.append("exact %s (oracle %.2fs, cuda %.3fs)" % ("EQUAL" if np.array_equal(got, want) else "DIFF %d max %d" % ((got != want).sum(), np.abs(got.astype(int) - want).max()), t1, t2))
        except Exception as e:
            msg.append("exact ERR " + type(e).__name__ + " " + str(e)[:80])
        print(name, sev, " | ".join(msg), flush=True)
