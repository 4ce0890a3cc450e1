#!/usr/bin/env python
"""bench.py — headline benchmark of the image-transformation hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1]): the CIFAR-10-C-style suite, 15 corruptions x 5 severities = 75 passes over a
synthetic 50000 x 32 x 32 x 3 uint8 batch resident in HBM.  One STEP = the 75 passes over one batch (75 kernel
launches).  Random corruptions draw on chip (Philox).  Multi-GPU: every rank transforms its own
50000-image shard (weak scaling, no data-path collective); the step time is the max over ranks.

`value`   transformed images/s, device-resident input and output, CUDA-event timed.
`e2e`     the same through the public API with HOST buffers: pinned host batch -> device, 75 passes, every output batch
          copied back to pinned host memory, all inside the timed region.
`roofline` the pass that takes the largest share of the step, timed per launch with CUDA events: algorithmic bytes
          (6144 B/image: 3072 read + 3072 written) / duration vs the measured HBM copy peak.
`cpu_baseline` / `--impl reference`: the CPU restatement of the same suite (oracle/cifar_c.py — the reference has no
          code for 11 of the 15 types and loads pre-made files; its Python cannot travel to the GPU box) on the host cores.
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


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
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
# CPU arm (oracle port)
# ------------------------------------------------------------------------------------------------
def _cpu_worker(args):
    seed, n = args
    import numpy as np
    from oracle import cifar_c as CC
    x = np.random.default_rng(seed).integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    rng = np.random.RandomState(seed)
    CC.zoom_blur(x[:1], 1)  # build nothing heavy inside the timed loop that the GPU arm also caches
    for s in range(2, 6):
        CC.zoom_blur(x[:1], s)
    t0 = time.perf_counter()
    for name in CC.CORRUPTIONS:
        for sev in range(1, 6):
            CC.corrupt(name, sev, x, CC.draw(name, sev, x, rng))
    return time.perf_counter() - t0


def cpu_suite_rate(images_per_core: int, cores: int):
    """images/s of the CPU restatement over the full 75-pass suite, one process per core."""
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        t0 = time.perf_counter()
        times = pool.map(_cpu_worker, [(100 + i, images_per_core) for i in range(cores)])
        wall = time.perf_counter() - t0
    total = 75 * images_per_core * cores
    return total / max(times), wall, max(times)


def run_reference(args, rank):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_core = 256   # about 10-15 s of CPU work per core and step
    vals = []
    for _ in range(args.warmup if args.warmup < 1 else 1):
        cpu_suite_rate(4, cores)
    for _ in range(args.steps):
        v, wall, tmax = cpu_suite_rate(per_core, cores)
        vals.append((v, tmax))
    v = statistics.median([a for a, _ in vals])
    ms = statistics.median([b for _, b in vals]) * 1e3
    sample = f"{per_core} images/core x {cores} cores x 75 passes per step (random draws included, numpy/scipy restatement)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "images/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": "cifar10c_suite_15x5", "images_per_gpu": N_IMAGES, "image": "32x32x3 uint8", "passes": 75},
        "cpu_baseline": {"value": v, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
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
    numa_cores = 0
    if world > 1 and not os.environ.get("IMGTX_NO_BIND"):
        # one process per GPU: keep it (and the pinned buffers it allocates below) on the GPU's own NUMA node
        from imagetransformations_b200 import shard as S
        numa_cores = S.bind_to_gpu_cpus(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    passes = C.suite_passes()
    launches_per_step = len(passes)  # every pass of the 32x32 suite is one kernel launch (pixelate is a fused resize chain)

    host = torch.from_numpy(np.random.default_rng(1001 + rank).integers(0, 256, (N_IMAGES, H, W, 3), dtype=np.uint8)).pin_memory()
    x = host.to(dev)
    outs = [torch.empty_like(x) for _ in range(2)]  # alternate: 2 x 153.6 MB, with x 460 MB > 126 MB of L2

    def step(seed):
        for k, (name, sev) in enumerate(passes):
            C.corrupt(x, name, sev, seed=seed * 1000 + k, out=outs[k & 1])

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(max(args.warmup, 1)):
        step(i)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(100 + i)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop()
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    value = world * len(passes) * N_IMAGES / (ms_per_step * 1e-3)

    # ---- per-pass timing (CUDA events on the launching stream) -> dominant kernel + roofline ----
    per_pass = {}
    reps = 5
    for k, (name, sev) in enumerate(passes):
        C.corrupt(x, name, sev, seed=k, out=outs[0])
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for r in range(reps):
            C.corrupt(x, name, sev, seed=k + r, out=outs[r & 1])
        b.record()
        torch.cuda.synchronize()
        per_pass[(name, sev)] = a.elapsed_time(b) / reps
    by_type = {}
    for (name, sev), t in per_pass.items():
        by_type[name] = by_type.get(name, 0.0) + t
    top = max(by_type, key=by_type.get)
    top_launches = 5
    top_ms_per_launch = by_type[top] / top_launches
    peak, peak_src = peaks()
    algo_bytes = ALGO_BYTES_PER_IMAGE * N_IMAGES
    achieved = algo_bytes / (top_ms_per_launch * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get(top)
    suite_gbs = len(passes) * algo_bytes / (ms_per_step * 1e-3) / 1e9

    # ---- end to end through the public API with host buffers ----
    e2e_steps = max(1, min(args.steps, 2))
    host_out = [torch.empty((N_IMAGES, H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
    copy_stream = torch.cuda.Stream()
    dev_out = [torch.empty_like(x) for _ in range(2)]
    done = [torch.cuda.Event() for _ in range(2)]
    barrier()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    for i in range(e2e_steps):
        xin = host.to(dev, non_blocking=True)
        for k, (name, sev) in enumerate(passes):
            b = k & 1
            torch.cuda.current_stream().wait_event(done[b])      # the copy that last read dev_out[b] has finished
            C.corrupt(xin, name, sev, seed=i * 1000 + k, out=dev_out[b])
            ready = torch.cuda.Event()
            ready.record()
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(ready)
                host_out[b].copy_(dev_out[b], non_blocking=True)
                done[b].record()
    torch.cuda.current_stream().wait_stream(copy_stream)
    s1.record()
    barrier()
    e2e_ms = s0.elapsed_time(s1)
    if world > 1:
        t = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    e2e_value = world * len(passes) * N_IMAGES * e2e_steps / (e2e_ms * 1e-3)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        per_core = 384   # about 15-25 s of CPU work per core
        v, wall, tmax = cpu_suite_rate(per_core, cores)
        cpu = {"value": v, "unit": "images/s", "cores": cores, "kind": "port",
               "sample": f"{per_core} images/core x {cores} cores x 75 passes (oracle/cifar_c.py, {tmax:.1f} s)"}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 1), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "cifar10c_suite_15x5", "images_per_gpu": N_IMAGES, "image": "32x32x3 uint8",
                       "passes": len(passes), "severity_table": "cifar", "rng": "on-chip Philox4x32-10",
                       "l2": "working set 461 MB (input + 2 rotating outputs) > 126 MB L2",
                       "host_cores_bound_per_rank": numa_cores},
            "suite_algorithmic_gbs": suite_gbs, "suite_frac_of_hbm_peak": suite_gbs / peak,
            "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "share_of_step": by_type[top] / sum(by_type.values()),
                         "algorithmic_bytes_per_launch": algo_bytes, "ms_per_launch": top_ms_per_launch},
            "per_corruption_ms": {k: round(v, 4) for k, v in by_type.items()},
            "per_corruption_gbs": {k: round(5 * algo_bytes / (v * 1e-3) / 1e9, 1) for k, v in by_type.items()},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "images/s", "h2d_bytes_per_step": int(x.numel()),
                    "d2h_bytes_per_step": int(x.numel()) * len(passes), "steps": e2e_steps,
                    "ms_per_step": e2e_ms / e2e_steps},
            "gpu_launches": launches_per_step * args.steps,
            "cpu_baseline": cpu,
        }))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
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
               "--steps", str(args.steps), "--warmup", str(args.warmup)] + (["--no-cpu"] if args.no_cpu else [])
        raise SystemExit(subprocess.call(cmd))
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
