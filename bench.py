#!/usr/bin/env python
"""bench.py — headline benchmark of the image-transformation hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--no-cpu] [--no-configs]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Headline workload (BASELINE.json configs[1], "C2"): the CIFAR-10-C-style suite, 15 corruptions x 5 severities = 75 passes over
a synthetic 50000 x 32 x 32 x 3 uint8 batch resident in HBM.  One STEP = the 75 passes over one batch.  Random corruptions
draw on chip (Philox) and run in the throughput mode (every byte within 1 LSB of the bit-exact mode on the same draws;
tests/test_gpu_fast_mode.py, tests/test_gpu_draw_laws.py).  Every pass also writes its per-image metric from the kernel
epilogue; with N > 1 ranks (one per GPU, each with its own 50000-image shard: weak scaling) the metrics of every pass are
all-gathered over NCCL on a side stream, overlapped with the next pass — the only collective the path has.

`value`        transformed images/s, device-resident input and output, CUDA-event timed, max over ranks.
`e2e`          the same through the public API with HOST buffers: pinned host batch -> device, 75 passes, every output batch
               copied back to pinned host memory, all inside the timed region; `e2e.bound` names the measured ceiling.
`roofline`     the corruption type with the largest share of the step, per-launch CUDA-event time: algorithmic bytes
               (6144 B/image: 3072 read + 3072 written) / duration vs the measured HBM copy peak (MEASURED_PEAKS.json).
`configs`      the other BASELINE.json configs (C1 transformation.py grid, C3 16 TTA views, C4 AugMix 224x224, C5 FGSM
               eps sweep fused with the post-attack leg; plus H19: the TransformationPool population step as one fused
               launch, and C2_224: the headline suite on Imagenette-shaped 224x224 images), each with device time, fraction of the HBM peak, e2e with host
               buffers and a CPU baseline (oracle/np_ops.py / np_chains.py: restatements pinned by goldens frozen from the
               live reference, tests/golden/).
`cpu_baseline` / `--impl reference`: the CPU restatement of the suite (oracle/cifar_c.py — the reference has no code for 11
               of the 15 types and loads pre-made files; its Python cannot travel to the GPU box) on all host cores.
"""
from __future__ import annotations

import argparse
import json
import multiprocessing as mp
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_IMAGES = 50000
H = W = 32
ALGO_BYTES_PER_IMAGE = 2 * H * W * 3
METRIC = "transformed images/sec (CIFAR-10-C-style suite, 15 corruptions x 5 severities)"
CONFIG = {"workload": "cifar10c_suite_15x5", "images_per_gpu": N_IMAGES, "image": "32x32x3 uint8", "passes": 75}
CIFAR10_MEAN, CIFAR10_STD = (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 50 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower() == "active"})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arms (oracle ports; the only place bench.py executes oracle/)
# ------------------------------------------------------------------------------------------------
def _cpu_suite_worker(args):
    seed, n = args
    import numpy as np
    from oracle import cifar_c as CC
    x = np.random.default_rng(seed).integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    rng = np.random.RandomState(seed)
    for s in range(1, 6):
        CC.zoom_blur(x[:1], s)      # the merged zoom operator is a cached table on both arms
    t0 = time.perf_counter()
    for name in CC.CORRUPTIONS:
        for sev in range(1, 6):
            CC.corrupt(name, sev, x, CC.draw(name, sev, x, rng))
    return time.perf_counter() - t0


def _cpu_h19_worker(args):
    seed, n = args
    import random
    import numpy as np
    from oracle import np_chains as CH, np_ops as O
    rs = np.random.RandomState(seed)
    random.seed(seed)
    tn = O.to_tensor_normalize(rs.randint(0, 256, (n, H, W, 3)).astype(np.uint8), CH.CIFAR10_MEAN, CH.CIFAR10_STD)
    enh = [("enhance_contrast", "factor"), ("enhance_brightness", "factor"), ("enhance_sharpness", "factor"),
           ("enhance_color", "factor"), ("motion_blur", "size")]
    t0 = time.perf_counter()
    for i in range(n):
        u = O.tensor_to_pil(tn[i:i + 1], CH.CIFAR10_MEAN, CH.CIFAR10_STD)
        u = CH.pool_op(u, "gaussian_noise", {"severity": 3}, {"noise": rs.normal(0, CH.POOL_GAUSS[2] * 255, u.shape)})
        u = CH.pool_op(u, "impulse_noise", {"severity": 3}, {"mask": rs.random_sample(u.shape[:3])})
        u = CH.pool_op(u, "shot_noise", {"severity": 3}, {"counts": rs.poisson(O.shot_noise_scaled(u, CH.POOL_SHOT[2]))})
        u = CH.pool_op(u, "defocus_blur", {"severity": 3})
        for _ in range(random.randint(0, 3)):
            name, key = random.choice(enh)
            u = CH.pool_op(u, name, {key: random.choice([5, 7, 9, 11]) if key == "size" else random.uniform(0.5, 2.0)})
        O.to_tensor_normalize(u, CH.CIFAR10_MEAN, CH.CIFAR10_STD)
    return time.perf_counter() - t0


def _cpu_suite224_worker(args):
    seed, n = args
    import numpy as np
    from oracle import cifar_c as CC
    x = np.random.default_rng(seed).integers(0, 256, (n, 224, 224, 3), dtype=np.uint8)
    rng = np.random.RandomState(seed)
    for s in range(1, 6):
        CC.zoom_blur(x[:1], s)      # the merged zoom operator is a cached table on both arms
    t0 = time.perf_counter()
    for name in CC.CORRUPTIONS:
        for sev in range(1, 6):
            CC.corrupt(name, sev, x, CC.draw(name, sev, x, rng))
    return time.perf_counter() - t0


C1_GRID = ([("rotate", a) for a in [-22.5 + 2.5 * i for i in range(19)]] + [("scale", s) for s in (0.9, 1.0, 1.1, 1.2000000000000002, 1.3)] +
           [("brightness", -0.05 + 0.01 * i) for i in range(11)] + [("contrast", 0.1 * i) for i in range(11)] +
           [("blur", 0.5 * i) for i in range(11)] + [("noise", 0.01 * i) for i in range(11)])   # transformation.py:92-170


def _cpu_c1_worker(args):
    seed, n = args
    import numpy as np
    from oracle import np_ops as O
    rng = np.random.default_rng(seed)
    x = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    t0 = time.perf_counter()
    for op, v in C1_GRID:
        if op == "rotate":
            O.pil_rotate_nearest(x, -v)
        elif op == "scale":
            O.apply_scale(x, v)
        elif op == "brightness":
            O.enhance_brightness(x, 1.0 + v)
        elif op == "contrast":
            O.convert_scale_abs(x, v)
        elif op == "blur":
            O.apply_blur(x, v)
        else:
            O.gaussian_noise_f32(x, rng.normal(0, v * 255, x.shape).astype(np.float32))
    return time.perf_counter() - t0


def _cpu_c3_worker(args):
    seed, n = args
    import numpy as np
    from oracle import np_chains as CH, np_ops as O
    rng = np.random.default_rng(seed)
    tn = O.to_tensor_normalize(rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8), CH.CIFAR100_MEAN, CH.CIFAR100_STD)
    crop = rng.integers(0, 9, (n, 5, 2))
    br, ct = rng.uniform(0.5, 1.5, (n, 5)), rng.uniform(0.8, 1.2, (n, 5))
    t0 = time.perf_counter()
    for i in range(n):
        CH.tta_views16(tn[i], crop[i], br[i], ct[i])
    return time.perf_counter() - t0


def _cpu_c4_worker(args):
    seed, n = args
    import numpy as np
    from oracle import np_chains as CH
    rng = np.random.default_rng(seed)
    t = rng.random((n, 3, 224, 224), dtype=np.float32)
    ops, signs = rng.integers(0, 8, (n, 3, 3)), rng.choice([-1, 1], (n, 3, 3))
    ws, m = rng.dirichlet([1, 1, 1], n), rng.beta(1, 1, n)
    t0 = time.perf_counter()
    for i in range(n):
        CH.augmix(t[i], 3, ws[i], m[i], ops[i], signs[i])
    return time.perf_counter() - t0


C5_EPS = [4 / 355, 8 / 255, 12 / 255]     # attacks/fgsm.py:297 ("4/355" is the reference's literal value)


def _cpu_c5_worker(args):
    seed, n = args
    import numpy as np
    from oracle import np_ops as O
    rng = np.random.default_rng(seed)
    x = O.to_tensor_normalize(rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8), CIFAR10_MEAN, CIFAR10_STD)
    g = rng.standard_normal(x.shape).astype(np.float32)
    lo = [(0 - m) / s for m, s in zip(CIFAR10_MEAN, CIFAR10_STD)]
    hi = [(1 - m) / s for m, s in zip(CIFAR10_MEAN, CIFAR10_STD)]
    lut = O.pil_blend_lut(0, 1.3)
    t0 = time.perf_counter()
    for e in C5_EPS:
        u = O.tensor_to_pil(O.fgsm_step(x, g, e, -1, lo, hi), CIFAR10_MEAN, CIFAR10_STD)
        O.to_tensor_normalize(lut[u], CIFAR10_MEAN, CIFAR10_STD)
    return time.perf_counter() - t0


def cpu_rate(worker, units_per_item, items_per_core, cores, what):
    """units/s of a CPU restatement, one process per core over a bounded sample."""
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        times = pool.map(worker, [(100 + i, items_per_core) for i in range(cores)])
    v = units_per_item * items_per_core * cores / max(times)
    return {"value": v, "unit": "images/s", "cores": cores, "kind": "port",
            "sample": f"{items_per_core} images/core x {cores} cores, {what} ({max(times):.1f} s)"}, max(times)


def run_reference(args, rank):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_core = 256   # about 10-15 s of CPU work per core and step
    warm = max(0, args.warmup)
    for _ in range(warm):
        cpu_rate(_cpu_suite_worker, 75, 2, cores, "warm-up")          # a real (tiny) pass through every code path per warm-up
    vals = []
    for _ in range(args.steps):
        cb, tmax = cpu_rate(_cpu_suite_worker, 75, per_core, cores, "75 passes, random draws included, numpy/scipy restatement")
        vals.append((cb, tmax))
    vals.sort(key=lambda t: t[0]["value"])
    cb, tmax = vals[len(vals) // 2]
    v = cb["value"]
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": warm, "ms_per_step": tmax * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic", "config": CONFIG,
        "cpu_baseline": cb,
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    from imagetransformations_b200 import cifar_c as C

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    passes = C.suite_passes()
    no_epilogue = set()                     # every pass writes its per-image metric from its own epilogue
    launches_per_step = len(passes) + sum(1 for p in passes if p[0] in no_epilogue)
    peak, peak_src = peaks()

    def ev():
        return torch.cuda.Event(enable_timing=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    host = torch.from_numpy(np.random.default_rng(1001 + rank).integers(0, 256, (N_IMAGES, H, W, 3), dtype=np.uint8)).pin_memory()
    x = host.to(dev)
    # The 75 passes read the same batch and write their own outputs and metric slices, so nothing orders them: the step is
    # captured as ONE CUDA graph of `lanes` independent chains of passes (pass k on chain k % lanes, writing out[k % 2]), and
    # a kernel's tail -- its last, partly idle wave of CTAs and the launch gap behind it -- is filled by the other chain.
    lanes = max(1, int(os.environ.get("IMGTX_BENCH_LANES", "4")))
    outs = [torch.empty_like(x) for _ in range(2 if lanes <= 2 else lanes)]  # rotate: lanes x 153.6 MB, with x 768 MB > 126 MB of L2
    # Per-pass metrics: this rank's kernels write their slice of a (rank, pass, image, 2) buffer from their epilogue, and ONE
    # NCCL all-gather per step fills in every other rank's slices in place.  Two buffers at N > 1: the gather of
    # step i runs on a side stream under the passes of step i + 1, which write the other buffer.
    n_buf = 2 if world > 1 else 1
    gbufs = [torch.zeros((world, len(passes), N_IMAGES, 2), dtype=torch.int32, device=dev) for _ in range(n_buf)]
    mviews_b = [[gb[rank, k] for k in range(len(passes))] for gb in gbufs]
    mviews = mviews_b[0]
    side = torch.cuda.Stream()
    graphs = None
    gdone = [None] * n_buf     # the gather that read buffer b has finished

    def gather(b):
        done = torch.cuda.Event()
        done.record()
        with torch.cuda.stream(side):
            side.wait_event(done)
            dist.all_gather_into_tensor(gbufs[b].view(-1), gbufs[b][rank].view(-1))
            gdone[b] = torch.cuda.Event()
            gdone[b].record()

    def step(i):
        b = i % n_buf
        if world > 1 and gdone[b] is not None:
            torch.cuda.current_stream().wait_event(gdone[b])   # the passes overwrite the slices that gather was reading
        if graphs is not None:
            graphs[b].replay()
        else:
            for k, (name, sev) in enumerate(passes):
                C.corrupt(x, name, sev, seed=i * 1000 + k, out=outs[k % len(outs)], metrics=mviews_b[b][k])
        if world > 1:
            gather(b)

    def drain():
        if world > 1:
            torch.cuda.current_stream().wait_stream(side)

    warm = max(args.warmup, 3)
    for i in range(warm):
        step(i)
    host_us_per_launch = None
    barrier()
    # host cost of the eager step (enqueue only): what a CUDA graph removes
    t0 = time.perf_counter()
    step(50)
    host_us_per_launch = (time.perf_counter() - t0) * 1e6 / launches_per_step
    barrier()

    if not args.no_graph:
        try:
            gg = [C.generate_suite(x, seed=100 + b, out=outs, metrics=mviews_b[b], graph=True, lanes=lanes) for b in range(n_buf)]
            torch.cuda.synchronize()
            graphs = gg
            for i in range(n_buf):
                step(i)
            barrier()
        except Exception as e:   # capture is an optimisation of the launch path, never a requirement
            graphs = None
            sys.stderr.write(f"[bench] CUDA graph capture unavailable ({type(e).__name__}: {e}); timing eager launches\n")

    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = ev(), ev()
    barrier()
    e0.record()
    for i in range(args.steps):
        step(100 + i)
    drain()            # every gather of the K steps is inside the timed region
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    timed_graph = graphs is not None
    n_outs = len(outs)
    clocks = sampler.stop()
    ms_per_step = ms / args.steps
    value = world * len(passes) * N_IMAGES / (ms_per_step * 1e-3)

    # the gathered metrics are real: every rank now holds every rank's words for every pass
    metrics_ok = all(bool((gb[r].abs().sum() > 0).item()) for gb in gbufs for r in range(world))
    if world > 1:
        for gb in gbufs:
            ref = gb[:, 15:20].clone()
            dist.broadcast(ref, 0)
            metrics_ok = metrics_ok and bool(torch.equal(ref, gb[:, 15:20]))

    # ---- per-pass timing (CUDA events on the launching stream) -> dominant kernel + roofline ----
    per_pass = {}
    reps = 5
    for k, (name, sev) in enumerate(passes):
        C.corrupt(x, name, sev, seed=k, out=outs[0], metrics=mviews[k])
        a, b = ev(), ev()
        a.record()
        for r in range(reps):
            C.corrupt(x, name, sev, seed=k + r, out=outs[r & 1], metrics=mviews[k])
        b.record()
        torch.cuda.synchronize()
        per_pass[(name, sev)] = a.elapsed_time(b) / reps
    by_type = {}
    for (name, sev), t in per_pass.items():
        by_type[name] = by_type.get(name, 0.0) + t
    top = max(by_type, key=by_type.get)
    top_ms_per_launch = by_type[top] / 5
    algo_bytes = ALGO_BYTES_PER_IMAGE * N_IMAGES
    achieved = algo_bytes / (top_ms_per_launch * 1e-3) / 1e9
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        tj = json.load(open(tp))
        traffic, traffic_src = tj.get(top), tj.get("_source")
    suite_gbs = len(passes) * algo_bytes / (ms_per_step * 1e-3) / 1e9
    worst = min(5 * algo_bytes / (v * 1e-3) / 1e9 / peak for v in by_type.values())

    # ---- output all-gather (optional in the API), timed on its own: 7/8 of a pass's bytes into every GPU ----
    out_gather = None
    if world > 1:
        big = torch.empty((world,) + tuple(x.shape), dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(big.view(-1), outs[0].view(-1))
        barrier()
        a, b = ev(), ev()
        a.record()
        for _ in range(3):
            dist.all_gather_into_tensor(big.view(-1), outs[0].view(-1))
        b.record()
        barrier()
        gms = max_over_ranks(a.elapsed_time(b) / 3)
        out_gather = {"ms_per_pass": gms, "recv_gbs_per_rank": (world - 1) * x.numel() / (gms * 1e-3) / 1e9,
                      "note": "timed separately; never part of the transform step"}
        del big

    # ---- end to end through the public API with host buffers ----
    # D2H ceiling of this box for this process: plain pinned copies of one output batch
    pin_probe = torch.empty_like(host).pin_memory()
    for _ in range(2):
        pin_probe.copy_(outs[0], non_blocking=True)
    barrier()
    a, b = ev(), ev()
    a.record()
    n_probe = 16                                             # 2.5 GB per rank: long enough to be a steady-state figure
    for _ in range(n_probe):
        pin_probe.copy_(outs[0], non_blocking=True)
    b.record()
    barrier()
    d2h_ceiling = max_over_ranks(a.elapsed_time(b))          # slowest rank, all ranks copying at once
    d2h_ceiling_gbs = world * n_probe * x.numel() / (d2h_ceiling * 1e-3) / 1e9
    del pin_probe

    e2e_steps = max(1, min(args.steps, 2))
    host_out = [torch.empty((N_IMAGES, H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
    copy_stream = torch.cuda.Stream()
    dev_out = [torch.empty_like(x) for _ in range(2)]
    done = [torch.cuda.Event() for _ in range(2)]
    barrier()
    s0, s1 = ev(), ev()
    s0.record()
    for i in range(e2e_steps):
        xin = host.to(dev, non_blocking=True)
        for k, (name, sev) in enumerate(passes):
            b2 = k & 1
            torch.cuda.current_stream().wait_event(done[b2])      # the copy that last read dev_out[b2] has finished
            C.corrupt(xin, name, sev, seed=i * 1000 + k, out=dev_out[b2])
            ready = torch.cuda.Event()
            ready.record()
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(ready)
                host_out[b2].copy_(dev_out[b2], non_blocking=True)
                done[b2].record()
    torch.cuda.current_stream().wait_stream(copy_stream)
    s1.record()
    barrier()
    e2e_ms = max_over_ranks(s0.elapsed_time(s1))
    e2e_value = world * len(passes) * N_IMAGES * e2e_steps / (e2e_ms * 1e-3)
    e2e_d2h_gbs = world * len(passes) * x.numel() * e2e_steps / (e2e_ms * 1e-3) / 1e9
    del host_out, dev_out

    configs = None
    if not args.no_configs:
        del outs, gbufs, mviews_b, mviews, graphs
        torch.cuda.empty_cache()
        configs = bench_configs(dev, rank, world, peak, barrier, max_over_ranks, want_cpu=(rank == 0 and world == 1 and not args.no_cpu))

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        cpu, _ = cpu_rate(_cpu_suite_worker, 75, 384, cores, "75 passes, oracle/cifar_c.py")   # about 15-25 s of CPU work per core

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": CONFIG,
            "details": {"severity_table": "cifar", "rng": "on-chip Philox4x32-10", "mode": "fast (<= 1 LSB from the exact mode)",
                        "l2": f"working set {153.6 * (1 + n_outs):.0f} MB (input + {n_outs} rotating outputs) > 126 MB L2",
                        "timed_with": (f"CUDA graph replay, {lanes} independent chain(s) of passes per graph" if timed_graph
                                       else "eager launches"),
                        "collective": ("one NCCL all_gather_into_tensor per step of the per-image metrics of its 75 passes "
                                       "(30 MB per rank), on a side stream under the next step's passes (two metric buffers)"
                                       if world > 1 else "none at N=1 (metrics still written by every pass)"),
                        "metrics_gathered_ok": metrics_ok},
            "suite_algorithmic_gbs": suite_gbs, "suite_frac_of_hbm_peak": suite_gbs / peak,
            "worst_type_frac_of_hbm_peak": worst,
            "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                         "share_of_step": by_type[top] / sum(by_type.values()),
                         "algorithmic_bytes_per_launch": algo_bytes, "ms_per_launch": top_ms_per_launch},
            "per_corruption_ms": {k: round(v, 4) for k, v in by_type.items()},
            "per_corruption_frac": {k: round(5 * algo_bytes / (v * 1e-3) / 1e9 / peak, 3) for k, v in by_type.items()},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": int(x.numel()),
                    "d2h_bytes_per_step": int(x.numel()) * len(passes), "steps": e2e_steps,
                    "ms_per_step": e2e_ms / e2e_steps, "d2h_gbs": e2e_d2h_gbs,
                    "bound": {"what": "PCIe device->host copies of the 75 output batches into pinned memory",
                              "ceiling_gbs": d2h_ceiling_gbs, "ceiling_how": "plain pinned cudaMemcpyAsync of one output batch, "
                              "all ranks at once, measured in this run", "achieved_frac_of_ceiling": e2e_d2h_gbs / d2h_ceiling_gbs}},
            "gpu_launches": launches_per_step * args.steps, "launches_per_step": launches_per_step,
            "host_us_per_launch_eager": host_us_per_launch,
            "output_all_gather": out_gather,
            "configs": configs,
            "cpu_baseline": cpu,
        }))
    if world > 1:
        dist.destroy_process_group()


def bench_configs(dev, rank, world, peak, barrier, max_over_ranks, want_cpu):
    """C1, C3, C4, C5 of BASELINE.json on this rank's own synthetic shard (weak scaling, time = max over ranks)."""
    import numpy as np
    import torch
    from imagetransformations_b200 import AugMix, batch as B, _lib as L, tta_initial, views as V
    cores = os.cpu_count() or 1
    res = {}
    rng = np.random.default_rng(2000 + rank)

    def ev():
        return torch.cuda.Event(enable_timing=True)

    def timed(fn, iters, warm=2):
        for _ in range(warm):
            fn()
        barrier()
        a, b = ev(), ev()
        a.record()
        for _ in range(iters):
            fn()
        b.record()
        barrier()
        return max_over_ranks(a.elapsed_time(b) / iters)

    def entry(ms, units, nbytes, e2e_ms, h2d, d2h, cpu, note):
        d = {"ms": ms, "images_per_s": world * units / (ms * 1e-3), "algorithmic_bytes": nbytes,
             "gbs": nbytes / (ms * 1e-3) / 1e9, "frac_of_hbm_peak": nbytes / (ms * 1e-3) / 1e9 / peak,
             "e2e": {"value": world * units / (e2e_ms * 1e-3), "unit": "images/s", "ms": e2e_ms, "h2d_bytes": h2d, "d2h_bytes": d2h},
             "cpu_baseline": cpu, "workload": note}
        return d

    # ---- C1: transformation.py grid (:92-170) on 10000 x 32x32x3, one pass per (op, value); 8 rotating outputs ----
    n1 = 10000
    h1 = torch.from_numpy(rng.integers(0, 256, (n1, H, W, 3), dtype=np.uint8)).pin_memory()
    x1 = h1.to(dev)
    rot = [torch.empty_like(x1) for _ in range(8)]          # 30.7 MB x 9 = 276 MB working set > 126 MB of L2
    k = [0]

    def nxt():
        k[0] = (k[0] + 1) % len(rot)
        return rot[k[0]]

    def c1_pass(op, v, out):
        if op == "rotate":
            return B.rotate(x1, -v, out=out)
        if op == "scale":
            return B.apply_scale(x1, v, out=out)
        if op == "brightness":
            return B.brightness(x1, 1.0 + v, out=out)
        if op == "contrast":
            return B.convert_scale_abs(x1, v, out=out)
        if op == "blur":
            return B.apply_blur(x1, v, out=out) if v > 0 else B.pointwise(x1, L.PW_IDENTITY, 0.0, out=out)
        return B.noise_philox(x1, L.NOISE_GAUSS_F32, v * 255.0, 7, out=out)

    def c1_all():
        for op, v in C1_GRID:
            c1_pass(op, v, nxt())
    ms = timed(c1_all, 3)
    per_op = {}
    for op in ("rotate", "scale", "brightness", "contrast", "blur", "noise"):
        sub = [(o, v) for o, v in C1_GRID if o == op]
        t = timed(lambda: [c1_pass(o, v, nxt()) for o, v in sub], 3, warm=1)
        per_op[op] = {"ms_per_pass": t / len(sub), "frac_of_hbm_peak": 2 * x1.numel() * len(sub) / (t * 1e-3) / 1e9 / peak}
    pin = [torch.empty_like(h1) for _ in range(2)]
    pin = [p.pin_memory() for p in pin]

    def c1_e2e():
        xin = h1.to(dev, non_blocking=True)
        for j, (op, v) in enumerate(C1_GRID):
            o = c1_pass(op, v, rot[j & 1])
            pin[j & 1].copy_(o, non_blocking=True)
        del xin
    e2e = timed(c1_e2e, 1, warm=1)
    cpu = cpu_rate(_cpu_c1_worker, len(C1_GRID), 1024, cores, f"{len(C1_GRID)} passes, oracle/np_ops.py (golden-pinned)")[0] if want_cpu else None
    res["C1"] = entry(ms, n1 * len(C1_GRID), 2 * x1.numel() * len(C1_GRID), e2e, int(x1.numel()), int(x1.numel()) * len(C1_GRID), cpu,
                      f"transformation.py grid: {len(C1_GRID)} (op, value) passes over 10000x32x32x3 u8, 8 rotating outputs (276 MB > L2)")
    res["C1"]["per_op"] = per_op
    del rot, pin, x1

    # ---- C3: 16 TTA views per image (tta_initial_claudeFixed.py:78,218-230), f32 normalised in and out, one fused launch ----
    x8 = torch.from_numpy(rng.integers(0, 256, (n1, H, W, 3), dtype=np.uint8)).to(dev)
    tn = B.to_tensor(x8, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD)
    np.random.seed(3)
    crop, br, ct = tta_initial.sample_view_draws(n1)
    draws = V.tta16_draws(n1, crop, br, ct, dev)
    vout = [torch.empty((n1, 16, 3, 32, 32), dtype=torch.float32, device=dev) for _ in range(2)]
    j = [0]

    def c3():
        j[0] ^= 1
        V.tta16(tn, None, None, None, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD, out=vout[j[0]], draws=draws)
    ms = timed(c3, 5)
    htn = tn.cpu().pin_memory()
    hv = torch.empty((n1, 16, 3, 32, 32), dtype=torch.float32).pin_memory()

    def c3_e2e():
        t = htn.to(dev, non_blocking=True)
        V.tta16(t, None, None, None, tta_initial.CIFAR100_MEAN, tta_initial.CIFAR100_STD, out=vout[0], draws=draws)
        hv.copy_(vout[0], non_blocking=True)
    e2e = timed(c3_e2e, 1, warm=1)
    cpu = cpu_rate(_cpu_c3_worker, 1, 2048, cores, "16 views per image, oracle/np_chains.py:tta_views16 (golden-pinned)")[0] if want_cpu else None
    res["C3"] = entry(ms, n1, n1 * (12288 + 16 * 12288), e2e, n1 * 12288, n1 * 16 * 12288, cpu,
                      "16 views per image (1 flip, 5 crop+BICUBIC, 5 brightness, 5 contrast), 10000 images, fused; 208896 B/image")
    res["C3"]["views_per_s"] = res["C3"]["images_per_s"] * 16
    del vout, hv, htn, tn, x8

    # ---- C4: AugMix width 3 depth 3 on 4096 x 3x224x224 f32 (AugMix.py:45-62), one fused launch ----
    n4 = 4096
    x4 = torch.rand((n4, 3, 224, 224), dtype=torch.float32, device=dev)
    o4 = torch.empty_like(x4)
    np.random.seed(4)
    import random
    random.seed(4)
    dr = [AugMix.sample_draws(3, 3) for _ in range(n4)]
    ws = np.stack([d[0] for d in dr]); m = np.array([d[1] for d in dr])
    ops = np.stack([d[2] for d in dr]); signs = np.stack([d[3] for d in dr])
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
    rd = torch.from_numpy(recs.view(np.uint8).reshape(-1).copy()).to(dev)
    wd = torch.from_numpy(np.ascontiguousarray(wts)).to(dev)
    lib = L.load()

    def c4(xi=x4, oi=o4, cnt=n4, roff=0, woff=0):
        L.check(lib.imgtx_augmix_f32(xi.data_ptr(), oi.data_ptr(), cnt, 224, 224, 3, 3, rd.data_ptr() + roff, wd.data_ptr() + woff, B._stream()))
    ms = timed(c4, 3, warm=1)
    chunk = 512                                           # e2e in 512-image chunks through two pinned buffers (308 MB each)
    hin = torch.empty((chunk, 3, 224, 224), dtype=torch.float32).pin_memory()
    hout = torch.empty((chunk, 3, 224, 224), dtype=torch.float32).pin_memory()
    rec_b, wt_b = recs[0].nbytes, wts[0].nbytes

    def c4_e2e():
        for c0 in range(0, n4, chunk):
            xi = hin.to(dev, non_blocking=True)
            c4(xi, o4[:chunk], chunk, c0 * rec_b, c0 * wt_b)
            hout.copy_(o4[:chunk], non_blocking=True)
    e2e = timed(c4_e2e, 1, warm=0)
    cpu = cpu_rate(_cpu_c4_worker, 1, 384, cores, "oracle/np_chains.py:augmix at 224x224 (golden-pinned)")[0] if want_cpu else None
    res["C4"] = entry(ms, n4, n4 * 1204224, e2e, n4 * 602112, n4 * 602112, cpu,
                      "AugMix width 3 depth 3 severity 3, 4096 x 3x224x224 f32, fused; 1204224 B/image")
    del x4, o4, hin, hout

    # ---- C5: FGSM eps sweep fused with clamp and the post-attack leg (attacks/fgsm.py:190-192,297; pool chain :141-152) ----
    x8 = torch.from_numpy(rng.integers(0, 256, (n1, H, W, 3), dtype=np.uint8)).to(dev)
    xf = B.to_tensor(x8, CIFAR10_MEAN, CIFAR10_STD)
    g = torch.randn(xf.shape, device=dev, generator=torch.Generator(dev).manual_seed(5))
    lo = [(0 - mm) / s for mm, s in zip(CIFAR10_MEAN, CIFAR10_STD)]
    hi = [(1 - mm) / s for mm, s in zip(CIFAR10_MEAN, CIFAR10_STD)]
    from imagetransformations_b200 import batch as Bm
    lut = _brightness_lut(1.3)
    ms = timed(lambda: Bm.sign_step_post(xf, g, C5_EPS, lut, CIFAR10_MEAN, CIFAR10_STD, -1.0, lo, hi), 10)
    ms_step_only = timed(lambda: Bm.sign_step(xf, g, C5_EPS, -1.0, lo, hi), 10)
    hx, hg = xf.cpu().pin_memory(), g.cpu().pin_memory()
    ho = torch.empty((3,) + tuple(xf.shape), dtype=torch.float32).pin_memory()

    def c5_e2e():
        a, b = hx.to(dev, non_blocking=True), hg.to(dev, non_blocking=True)
        ho.copy_(Bm.sign_step_post(a, b, C5_EPS, lut, CIFAR10_MEAN, CIFAR10_STD, -1.0, lo, hi), non_blocking=True)
    e2e = timed(c5_e2e, 1, warm=1)
    cpu = cpu_rate(_cpu_c5_worker, 3, 8192, cores, "3 eps x (sign step, clamp, tensor_to_pil, brightness map, pil_to_tensor), oracle/np_ops.py")[0] if want_cpu else None
    res["C5"] = entry(ms, n1 * 3, n1 * (24576 + 3 * 12288), e2e, n1 * 24576, n1 * 3 * 12288, cpu,
                      "reverse-FGSM 3-eps sweep + per-channel clamp + tensor_to_pil -> Brightness(1.3) -> pil_to_tensor, ONE kernel, "
                      "10000 x 3x32x32 f32 image + gradient; 61440 B per input image (images_per_s counts output images)")
    res["C5"]["sign_step_only_ms"] = ms_step_only
    del x8, xf, g, hx, hg, ho

    # ---- H19: the GA's population step (pipenline/cifar_image_transformations.py:141-152, 203-226): base chain + 0..3 enhancers,
    #      10000 individuals (500 images x 20 individuals), ONE fused launch, noise drawn on chip ----
    from imagetransformations_b200 import cifar_image_transformations as Mp
    import random as _random
    _random.seed(1234 + rank)
    base = [("gaussian_noise", {"severity": 3}), ("impulse_noise", {"severity": 3}), ("shot_noise", {"severity": 3}),
            ("defocus_blur", {"severity": 3})]
    enh = ["enhance_contrast", "enhance_brightness", "enhance_sharpness", "enhance_color", "motion_blur"]
    chains = [base + [(_random.choice(enh), {}) for _ in range(_random.randint(0, 3))] for _ in range(n1)]
    recs = [Mp.encode_chain(c) for c in chains]
    ops_np = np.zeros((n1, 7), Mp.POOL_OP_DTYPE)
    ops_np["kind"] = -1
    for i, r in enumerate(recs):
        ops_np[i, :len(r)] = r
    xg = B.to_tensor(torch.from_numpy(rng.integers(0, 256, (n1, H, W, 3), dtype=np.uint8)).to(dev), CIFAR10_MEAN, CIFAR10_STD)
    ms = timed(lambda: Mp.pool_chain_batch(xg, ops_np, seed=3), 5)

    def unfused():   # the same ops as separate launches over the whole batch (base chain only), on-chip noise
        u = B.to_uint8(xg, CIFAR10_MEAN, CIFAR10_STD, clamp=True)
        u = B.noise_philox(u, L.NOISE_GAUSS_F64, Mp.GAUSSIAN_LEVELS[2] * 255.0, 3)
        u = B.noise_philox(u, L.NOISE_IMPULSE, Mp.IMPULSE_LEVELS[2], 4)
        u = B.noise_philox(u, L.NOISE_SHOT, Mp.SHOT_LEVELS[2], 5)
        u = B.pil_gaussian_blur(u, Mp.DEFOCUS_LEVELS[2])
        return B.to_tensor(u, CIFAR10_MEAN, CIFAR10_STD)
    ms_unfused = timed(unfused, 5)
    hxg = xg.cpu().pin_memory()
    hog = torch.empty_like(hxg).pin_memory()

    def h19_e2e():
        hog.copy_(Mp.pool_chain_batch(hxg.to(dev, non_blocking=True), ops_np, seed=3), non_blocking=True)
    e2e = timed(h19_e2e, 1, warm=1)
    cpu = cpu_rate(_cpu_h19_worker, 1, 256, cores, "base chain + 0..3 enhancers per individual, oracle/np_chains.py:individual_apply (golden-pinned)")[0] if want_cpu else None
    res["H19"] = entry(ms, n1, n1 * 24576, e2e, n1 * 12288, n1 * 12288, cpu,
                       "TransformationPool population step: tensor_to_pil -> gaussian/impulse/shot noise, defocus_blur (severity 3) "
                       "-> 0..3 enhancers -> pil_to_tensor, 10000 individuals, ONE launch; 24576 B/individual")
    res["H19"]["base_chain_as_separate_launches_ms"] = ms_unfused
    del xg, hxg, hog

    # ---- C2 on Imagenette-shaped images (north star: "CIFAR- and Imagenette-shaped batches"): the same 75 passes over
    #      1024 x 224x224x3 (154 MB per batch: the bytes of the 50000-image CIFAR batch), on-chip draws, two rotating outputs ----
    from imagetransformations_b200 import cifar_c as Cc
    n2 = 1024
    hL = torch.from_numpy(rng.integers(0, 256, (n2, 224, 224, 3), dtype=np.uint8)).pin_memory()
    xL = hL.to(dev)
    oL = [torch.empty_like(xL) for _ in range(2)]
    passes = Cc.suite_passes()

    def suite224():
        for j, (name, sev) in enumerate(passes):
            Cc.corrupt(xL, name, sev, seed=3, out=oL[j & 1])
    ms = timed(suite224, 2, warm=1)
    per_type = {}
    for name in Cc.CORRUPTIONS:
        t = timed(lambda: [Cc.corrupt(xL, name, sev, seed=3, out=oL[sev & 1]) for sev in range(1, 6)], 2, warm=1)
        per_type[name] = round(t, 3)
    pinL = [torch.empty_like(hL).pin_memory() for _ in range(2)]

    def suite224_e2e():
        xin = hL.to(dev, non_blocking=True)
        for j, (name, sev) in enumerate(passes):
            pinL[j & 1].copy_(Cc.corrupt(xin, name, sev, seed=3, out=oL[j & 1]), non_blocking=True)
    e2e = timed(suite224_e2e, 1, warm=1)
    cpu = cpu_rate(_cpu_suite224_worker, 75, 2, cores, "75 passes at 224x224, random draws included, oracle/cifar_c.py")[0] if want_cpu else None
    res["C2_224"] = entry(ms, n2 * 75, 2 * int(xL.numel()) * 75, e2e, int(xL.numel()), int(xL.numel()) * 75, cpu,
                          "the 15 x 5 suite on 1024 x 224x224x3 uint8 (Imagenette-shaped), on-chip draws; general-size kernels "
                          "(global workspaces), not the CIFAR throughput forms")
    res["C2_224"]["per_corruption_ms"] = per_type
    return res


def _brightness_lut(factor):
    """ImageEnhance.Brightness(factor) as the 256-entry map Pillow's blend produces (f32 arithmetic, libImaging/Blend.c) — host
    table for the C5 post-attack leg (the product's own LUT builder lives in csrc/lut.cuh; this is its 6-line host twin)."""
    import numpy as np
    p = np.arange(256, dtype=np.float32)
    t = (np.float32(factor) * p).astype(np.float32)
    return np.where(t <= 0, 0, np.where(t >= 255, 255, np.trunc(t))).astype(np.uint8)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--no-configs", action="store_true", help="headline workload only")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of a CUDA graph replay")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world == 1 and args.gpus > 1:
        # convenience: re-launch under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29511", os.path.abspath(__file__), "--gpus", str(args.gpus),
               "--steps", str(args.steps), "--warmup", str(args.warmup)] + (["--no-cpu"] if args.no_cpu else []) + \
              (["--no-configs"] if args.no_configs else []) + (["--no-graph"] if args.no_graph else [])
        raise SystemExit(subprocess.call(cmd))
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
