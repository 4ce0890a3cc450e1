"""Device->host bandwidth matrix of one box (development tool): what bounds the end-to-end number.
One process drives every visible GPU: pinned cudaMemcpyAsync D2H of 153.6 MB (one output batch) and 768 MB (a 5-severity
group) per GPU alone, then 2 / 4 / 8 GPUs at once, write-combined host buffers, and the host's own memcpy rate.
usage: python tools/d2h_matrix.py > profiles/<name>.json"""
import json
import sys
import time

import numpy as np
import torch


def timed_copies(pairs, reps):
    """pairs: list of (device tensor, pinned host tensor, stream). Returns seconds for reps rounds of all copies at once."""
    for d, h, s in pairs:
        with torch.cuda.stream(s):
            h.copy_(d, non_blocking=True)
    for _, _, s in pairs:
        s.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        for d, h, s in pairs:
            with torch.cuda.stream(s):
                h.copy_(d, non_blocking=True)
    for _, _, s in pairs:
        s.synchronize()
    return time.perf_counter() - t0


def main():
    ng = torch.cuda.device_count()
    res = {"gpus": ng, "unit": "GB/s", "single": {}, "concurrent": {}, "notes": []}
    sizes = {"153.6MB": 153_600_000, "768MB": 768_000_000}
    dev, host, streams = {}, {}, {}
    for g in range(ng):
        torch.cuda.set_device(g)
        dev[g] = torch.empty(sizes["768MB"], dtype=torch.uint8, device=f"cuda:{g}")
        host[g] = torch.empty(sizes["768MB"], dtype=torch.uint8).pin_memory()
        streams[g] = torch.cuda.Stream(device=g)
    for name, nb in sizes.items():
        reps = 8 if nb < 2e8 else 3
        res["single"][name] = {}
        for g in range(ng):
            t = timed_copies([(dev[g][:nb], host[g][:nb], streams[g])], reps)
            res["single"][name][f"gpu{g}"] = round(nb * reps / t / 1e9, 2)
        res["concurrent"][name] = {}
        for k in (1, 2, 4, 8):
            if k > ng:
                break
            t = timed_copies([(dev[g][:nb], host[g][:nb], streams[g]) for g in range(k)], reps)
            res["concurrent"][name][f"{k}_gpus_aggregate"] = round(k * nb * reps / t / 1e9, 2)
    # pairs (two GPUs at once): do they share a host link?
    if ng >= 2:
        nb = sizes["153.6MB"]
        res["pairs_153.6MB"] = {}
        for a, b in [(0, 1), (0, 2), (0, 4), (0, ng - 1), (ng // 2, ng - 1)]:
            if a < ng and b < ng and a != b:
                t = timed_copies([(dev[a][:nb], host[a][:nb], streams[a]), (dev[b][:nb], host[b][:nb], streams[b])], 8)
                res["pairs_153.6MB"][f"gpu{a}+gpu{b}"] = round(2 * nb * 8 / t / 1e9, 2)
    # staggered: ranks start 1/k of a copy apart (does de-phasing the ranks help?)
    if ng >= 4:
        nb = sizes["153.6MB"]
        k = min(8, ng)
        one = nb / (res["single"]["153.6MB"]["gpu0"] * 1e9)
        for g in range(k):
            streams[g].synchronize()
        t0 = time.perf_counter()
        for r in range(8):
            for g in range(k):
                with torch.cuda.stream(streams[g]):
                    host[g][:nb].copy_(dev[g][:nb], non_blocking=True)
                time.sleep(one / k)
        for g in range(k):
            streams[g].synchronize()
        res["concurrent"]["153.6MB"][f"{k}_gpus_staggered"] = round(k * nb * 8 / (time.perf_counter() - t0) / 1e9, 2)
    # write-combined host memory (cudaHostAllocWriteCombined) through the runtime API
    try:
        from cuda import cudart
        nb = sizes["153.6MB"]
        k = min(8, ng)
        bufs = []
        for g in range(k):
            cudart.cudaSetDevice(g)
            err, p = cudart.cudaHostAlloc(nb, cudart.cudaHostAllocWriteCombined | cudart.cudaHostAllocPortable)
            assert int(err) == 0, err
            bufs.append(p)

        def wc(gs, reps=8):
            for rep in range(reps + 1):
                if rep == 1:
                    for g in gs:
                        cudart.cudaSetDevice(g)
                        cudart.cudaStreamSynchronize(streams[g].cuda_stream)
                    t0 = time.perf_counter()
                for g in gs:
                    cudart.cudaSetDevice(g)
                    cudart.cudaMemcpyAsync(bufs[g], dev[g].data_ptr(), nb, cudart.cudaMemcpyKind.cudaMemcpyDeviceToHost,
                                           streams[g].cuda_stream)
            for g in gs:
                cudart.cudaSetDevice(g)
                cudart.cudaStreamSynchronize(streams[g].cuda_stream)
            return len(gs) * nb * reps / (time.perf_counter() - t0) / 1e9
        res["write_combined_153.6MB"] = {"1_gpu": round(wc([0]), 2), f"{k}_gpus_aggregate": round(wc(list(range(k))), 2)}
        for p in bufs:
            cudart.cudaFreeHost(p)
    except Exception as e:   # noqa: BLE001
        res["notes"].append(f"write-combined probe failed: {type(e).__name__}: {e}")
    # host memory: one-thread and all-thread copy rates (what a consumer of the pinned buffers could sustain)
    a = np.ones(1 << 30, np.uint8)
    b = np.empty_like(a)
    np.copyto(b, a)
    t0 = time.perf_counter()
    for _ in range(3):
        np.copyto(b, a)
    res["host_memcpy_1_thread"] = round(3 * a.nbytes / (time.perf_counter() - t0) / 1e9, 2)
    ta, tb = torch.from_numpy(a), torch.from_numpy(b)
    tb.copy_(ta)
    t0 = time.perf_counter()
    for _ in range(3):
        tb.copy_(ta)
    res["host_memcpy_torch_threads"] = {"threads": torch.get_num_threads(), "GBs": round(3 * a.nbytes / (time.perf_counter() - t0) / 1e9, 2)}
    json.dump(res, sys.stdout, indent=1)
    print()


main()
