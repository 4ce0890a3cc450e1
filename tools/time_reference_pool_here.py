"""Time the REFERENCE's own four CIFAR-C-style generators and the GA base chain on the CPU cores of the build container
(the only place /root/reference exists): a second, kind "reference" CPU figure next to the oracle
port that bench.py times on the GPU box.  usage: python tools/time_reference_pool_here.py > profiles/<name>.json"""
import json
import multiprocessing as mp
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def worker(args):
    seed, n = args
    import random
    import numpy as np
    from PIL import Image
    from oracle import ref_loader
    P = ref_loader.pool().TransformationPool
    rs = np.random.RandomState(seed)
    np.random.seed(seed)
    random.seed(seed)
    imgs = [Image.fromarray(rs.randint(0, 256, (32, 32, 3)).astype(np.uint8)) for _ in range(n)]
    res = {}
    for name in ("gaussian_noise", "impulse_noise", "shot_noise", "defocus_blur"):
        t0 = time.perf_counter()
        for im in imgs:
            getattr(P, name)(im, severity=3)
        res[name] = time.perf_counter() - t0
    t0 = time.perf_counter()
    for im in imgs:
        u = P.gaussian_noise(im, severity=3)
        u = P.impulse_noise(u, severity=3)
        u = P.shot_noise(u, severity=3)
        P.defocus_blur(u, severity=3)
    res["base_chain"] = time.perf_counter() - t0
    return res


def main():
    cores = os.cpu_count() or 1
    n = 2000
    with mp.get_context("fork").Pool(cores) as pool:
        out = pool.map(worker, [(10 + i, n) for i in range(cores)])
    keys = out[0].keys()
    rates = {k: round(n * cores / max(o[k] for o in out), 1) for k in keys}
    json.dump({"what": "the reference's own TransformationPool generators (pipenline/cifar_image_transformations.py:39-77) "
                       "on PIL images, one process per core, in the build container", "kind": "reference", "cores": cores,
               "images_per_core": n, "unit": "images/s", "rates": rates}, sys.stdout, indent=1)
    print()


main()
