"""Device timing of the other BASELINE.json configs (C1, C3, C4, C5) at their stated shapes (development tool; the headline
bench is bench.py = C2).  CUDA events, outputs rotated so the working set exceeds L2 where it can.
usage: python tools/config_bench.py  ->  gpurun_out/config_bench.json"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from imagetransformations_b200 import AugMix, batch as B, _lib as L, tta_initial, views as V

PEAK = 6542.7


def timeit(fn, iters=20, warm=5):
    if os.environ.get("ONCE"):      # one call (for ncu captures)
        fn()
        torch.cuda.synchronize()
        return 1.0
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e-3


def main():
    res = {}

    only = os.environ.get("ONLY", "")

    def rec(name, secs, nbytes, units, unit_name):
        res[name] = {"us": round(secs * 1e6, 1), "GBs": round(nbytes / secs / 1e9, 1), "frac_of_hbm_peak": round(nbytes / secs / 1e9 / PEAK, 3),
                     unit_name + "_per_s": round(units / secs, 1)}
        print(f"{name:44s} {secs * 1e6:9.1f} us {nbytes / secs / 1e9:8.1f} GB/s {nbytes / secs / 1e9 / PEAK:6.1%}  {units / secs:.3e} {unit_name}/s", flush=True)

    rng = np.random.default_rng(1000)
    # ---- C1: transformation.py suite, 10000 x 32x32x3 u8 (6144 B/image/pass) ----
    n1 = 10000
    x = torch.from_numpy(rng.integers(0, 256, (n1, 32, 32, 3), dtype=np.uint8)).cuda()
    outs = [torch.empty_like(x) for _ in range(8)]
    k = [0]

    def nxt():
        k[0] = (k[0] + 1) % len(outs)
        return outs[k[0]]
    b1 = 2 * x.numel()
    rec("C1 rotate 12.5 deg (NEAREST 16.16)", timeit(lambda: B.rotate(x, 12.5, out=nxt())), b1, n1, "img")
    rec("C1 scale 1.1 (LANCZOS Q22 + crop)", timeit(lambda: B.apply_scale(x, 1.1, out=nxt())), b1, n1, "img")
    rec("C1 brightness 1.03 (LUT)", timeit(lambda: B.brightness(x, 1.03, out=nxt())), b1, n1, "img")
    rec("C1 contrast 0.5 (convertScaleAbs LUT)", timeit(lambda: B.convert_scale_abs(x, 0.5, out=nxt())), b1, n1, "img")
    rec("C1 blur r=2 (cv Q8 11x11 separable)", timeit(lambda: B.apply_blur(x, 2.0, out=nxt())), b1, n1, "img")
    rec("C1 gaussian noise 0.05 (Philox)", timeit(lambda: B.noise_philox(x, L.NOISE_GAUSS_F32, 0.05, 7, out=nxt())), b1, n1, "img")
    rec("C1 translate (+5,-3)", timeit(lambda: B.translate(x, 5, -3, out=nxt())), b1, n1, "img")
    # ---- C3: 16-view TTA expansion, 10000 images, f32 normalised in/out (208 896 B/image) ----
    tn = B.to_tensor(x, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD)
    np.random.seed(3)
    crop, br, ct = tta_initial.sample_view_draws(n1)
    vout = [torch.empty((n1, 16, 3, 32, 32), dtype=torch.float32, device="cuda") for _ in range(2)]
    draws = V.tta16_draws(n1, crop, br, ct, "cuda")
    rec("C3 TTA 16 views / image (fused)", timeit(lambda: V.tta16(tn, None, None, None, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD, out=vout[k[0] & 1], draws=draws), iters=10),
        n1 * (12288 + 16 * 12288), n1 * 16, "view")
    # ---- 8f-3: Shannon-entropy gate over the same normalised batch (12 288 B read per image) ----
    rec("8f-3 Shannon-entropy gate (256-bin, f64)", timeit(lambda: B.shannon_entropy(tn, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD)),
        n1 * 12288, n1, "img")
    # ---- 8f-4: tensor-backend TTA chain, 10 views per image (12 288 B read + 10 * 12 288 B written per image) ----
    from imagetransformations_b200 import post_hoc_aaryaa as PH
    torch.manual_seed(5)
    base = [PH.chain_record(PH.draw_chain(32, 32)) for _ in range(512)]
    recs = np.stack([base[i % 512] for i in range(n1 * 10)])
    rdev = torch.from_numpy(recs.view(np.uint8).reshape(-1).copy()).cuda()
    tout = torch.empty((n1 * 10, 3, 32, 32), dtype=torch.float32, device="cuda")
    lib0 = L.load()
    rec("8f-4 tensor TTA chain, 10 views / image (fused)",
        timeit(lambda: L.check(lib0.imgtx_tensor_tta_f32(tn.data_ptr(), tout.data_ptr(), n1, 10, 3, 32, 32, rdev.data_ptr(), B._stream())), iters=10),
        n1 * 11 * 12288, n1 * 10, "view")
    # ---- C5: FGSM eps sweep, 10000 x 3x32x32 f32 (24 576 B read + E * 12 288 B written per image) ----
    xf = B.to_tensor(x)
    g = torch.randn_like(xf)
    eps = [4 / 355, 8 / 255, 12 / 255]
    rec("C5 FGSM sign-step, 3 eps in one pass", timeit(lambda: B.sign_step(xf, g, eps, lo=0.0, hi=1.0)), n1 * (24576 + 3 * 12288), n1 * 3, "img")
    rec("C5 FGSM sign-step, 1 eps", timeit(lambda: B.sign_step(xf, g, eps[0], lo=0.0, hi=1.0)), n1 * (24576 + 12288), n1, "img")
    # ---- C4: AugMix width 3 depth 3 on 4096 x 3x224x224 f32 (1 204 224 B/image) ----
    n4 = int(os.environ.get("N4", 4096))
    x4 = torch.rand((n4, 3, 224, 224), dtype=torch.float32, device="cuda")
    o4 = torch.empty_like(x4)
    np.random.seed(4)
    import random
    random.seed(4)
    draws = [AugMix.sample_draws(3, 3) for _ in range(n4)]
    ws = np.stack([d[0] for d in draws]); m = np.array([d[1] for d in draws])
    ops = np.stack([d[2] for d in draws]); signs = np.stack([d[3] for d in draws])
    # build the per-image op records once (host work of augmix_batch), then time the launch alone
    recs = np.zeros((n4, 3, 3), AugMix.AUG_OP_DTYPE)
    cache = {}
    for i in range(n4):
        for a in range(3):
            for b in range(3):
                key = (int(ops[i, a, b]), 3.0, int(signs[i, a, b]))
                if key not in cache:
                    cache[key] = AugMix.compile_op(key[0], 3, key[2], 224, 224)
                recs[i, a, b] = cache[key]
    wts = np.concatenate([ws.astype(np.float32), m.astype(np.float32)[:, None], (1.0 - m).astype(np.float32)[:, None]], axis=1)
    rd = torch.from_numpy(recs.view(np.uint8).reshape(-1).copy()).cuda()
    wd = torch.from_numpy(np.ascontiguousarray(wts)).cuda()
    lib = L.load()
    rec("C4 AugMix w3 d3, 224x224 (fused, one pass)",
        timeit(lambda: L.check(lib.imgtx_augmix_f32(x4.data_ptr(), o4.data_ptr(), n4, 224, 224, 3, 3, rd.data_ptr(), wd.data_ptr(), B._stream())), iters=5, warm=2),
        n4 * 1204224, n4, "img")
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/config_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
