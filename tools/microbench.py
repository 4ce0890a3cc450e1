"""Quick device-side bandwidth probe of individual kernels (development tool, not the bench)."""
import sys, os, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from imagetransformations_b200 import batch as B, _lib as L

def timeit(fn, iters=20, warm=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3

def main():
    n = int(os.environ.get("N", 50000))
    x = torch.randint(0, 256, (n, 32, 32, 3), dtype=torch.uint8, device="cuda")
    outs = [torch.empty_like(x) for _ in range(4)]  # rotate outputs so L2 does not hold them
    k = [0]
    def nxt():
        k[0] = (k[0] + 1) % len(outs); return outs[k[0]]
    res = {}
    bytes_pass = 2 * x.numel()
    def rec(name, fn, nbytes=bytes_pass):
        t = timeit(fn); res[name] = dict(us=t * 1e6, GBs=nbytes / t / 1e9); print(name, res[name], flush=True)
    rec("copy_", lambda: nxt().copy_(x))
    rec("brightness_uniform", lambda: B.brightness(x, 1.03, out=nxt()))
    f = torch.rand(n, device="cuda") + 0.5
    rec("brightness_per_image", lambda: B.brightness(x, f, out=nxt()))
    rec("rotate_uniform", lambda: B.rotate(x, 12.5, out=nxt()))
    rec("hflip", lambda: B.hflip(x, out=nxt()))
    rec("translate", lambda: B.translate(x, 5, -3, out=nxt()))
    t32 = torch.empty((n, 3, 32, 32), dtype=torch.float32, device="cuda")
    rec("to_tensor_norm", lambda: B.to_tensor(x, (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010), out=t32), x.numel() * 5)
    u8 = torch.empty_like(x)
    rec("to_uint8", lambda: B.to_uint8(t32, (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010), clamp=True, out=u8), x.numel() * 5)
    g = torch.randn_like(t32)
    rec("sign_step_E3", lambda: B.sign_step(t32, g, [4/355, 8/255, 12/255]), t32.numel() * 4 * 5)
    for (smem, st, cps) in [(40000, 3, 5), (56000, 3, 4), (72000, 3, 3), (72000, 4, 3), (110000, 3, 2), (110000, 4, 2), (200000, 4, 1), (28000, 3, 8), (28000, 2, 8)]:
        L.load().imgtx_set_tuning(smem, st, cps)
        rec(f"rotate smem={smem} st={st} cps={cps}", lambda: B.rotate(x, 12.5, out=nxt()))
    json.dump(res, open("gpurun_out/microbench.json", "w"), indent=1)

if __name__ == "__main__":
    os.makedirs("gpurun_out", exist_ok=True)
    main()
