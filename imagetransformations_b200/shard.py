"""Image-sharded data parallelism.

Every operator on the path is per image, so a batch is cut into contiguous index ranges, one per rank; per-image
parameters and injected draws are cut the same way and the sharded result is bit-identical to the single-GPU result.
There is NO collective on the data path.  The only exchanges are the ones BASELINE.json names: an all-gather of the small
per-image metrics (always) and, on request, of the transformed outputs.  One process per GPU (torchrun); the backend is
whatever the process group was created with (nccl on the B200 box, gloo in the CPU tests of the host logic).

The reference has no distributed code at all; its per-image loops (transformation.py:113-164,
fall_2025/TTA_transforms.py:244-254) are the unit that is sharded here.
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def world() -> Tuple[int, int]:
    """(rank, world_size); (0, 1) when torch.distributed is not initialised."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n: int, rank: Optional[int] = None, world_size: Optional[int] = None) -> Tuple[int, int]:
    """Contiguous slice [lo, hi) of n images owned by `rank`: rank r of R gets [r*n//R, (r+1)*n//R) (ragged when R does not
    divide n; empty when n < R for the trailing ranks)."""
    r, w = world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    if not 0 <= rank < world_size:
        raise ValueError("rank outside the world")
    return rank * n // world_size, (rank + 1) * n // world_size


def shard_sizes(n: int, world_size: int) -> list:
    return [shard_range(n, r, world_size)[1] - shard_range(n, r, world_size)[0] for r in range(world_size)]


def shard(x, rank: Optional[int] = None, world_size: Optional[int] = None):
    """This rank's rows of a per-image array/tensor (images, parameters or draws: anything indexed by image first)."""
    lo, hi = shard_range(len(x), rank, world_size)
    return x[lo:hi]


def shard_draws(draws: Optional[Dict[str, object]], n: int, rank: Optional[int] = None,
                world_size: Optional[int] = None) -> Optional[Dict[str, object]]:
    """Slice every per-image entry (leading dimension n) of an injected-draw dict; other entries pass through."""
    if draws is None:
        return None
    lo, hi = shard_range(n, rank, world_size)
    return {k: (v[lo:hi] if hasattr(v, "__len__") and len(v) == n else v) for k, v in draws.items()}


def all_gather_rows(local: torch.Tensor, n_total: int, group=None) -> torch.Tensor:
    """Concatenate the ranks' row blocks (ragged allowed) in rank order -> (n_total, ...).  Used for per-image metrics
    (a few bytes per image) and, optionally, for outputs; the data path itself never calls this."""
    rank, ws = world()
    if ws == 1:
        return local
    sizes = shard_sizes(n_total, ws)
    if local.shape[0] != sizes[rank]:
        raise ValueError(f"rank {rank} holds {local.shape[0]} rows, expected {sizes[rank]}")
    if len(set(sizes)) == 1:
        out = torch.empty((n_total,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    # ragged shards: pad every block to the largest one, gather once, drop the padding
    m = max(sizes)
    padded = torch.zeros((m,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    padded[:local.shape[0]] = local
    out = torch.empty((ws * m,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded, group=group)
    return torch.cat([out[r * m:r * m + sizes[r]] for r in range(ws)], dim=0)


def image_checksums(x: torch.Tensor) -> torch.Tensor:
    """Per-image metric that is all-gathered after a pass: (n, 2) int32 holding the uint32 words
    [sum of bytes, sum of (byte index + 1) * byte mod 2^32].  Order-sensitive inside an image, so sharded and single-GPU
    runs can be compared without moving the images.  CUDA tensors: one libimgtx kernel (imgtx_image_checksums_u8; the suite's
    throughput kernels produce the same words in their epilogue, see cifar_c.corrupt(metrics=...)).  CPU tensors (the gloo
    tests of the host logic): the same formula in torch integer arithmetic."""
    n = x.shape[0]
    nbytes = int(x[0].numel()) if n else 0
    if x.is_cuda:
        from . import _lib as L, batch as B
        x = x.contiguous()
        out = torch.empty((n, 2), dtype=torch.int32, device=x.device)
        if n:
            with torch.cuda.device_of(x):
                L.check(L.load().imgtx_image_checksums_u8(x.data_ptr(), n, nbytes, out.data_ptr(), B._stream()))
        return out
    flat = x.reshape(n, max(nbytes, 1)).to(torch.int64) if n else x.new_zeros((0, 1), dtype=torch.int64)
    wts = torch.arange(1, flat.shape[1] + 1, dtype=torch.int64)
    both = torch.stack([flat.sum(1), (flat * wts).sum(1)], dim=1) & 0xFFFFFFFF
    return torch.where(both >= 2 ** 31, both - 2 ** 32, both).to(torch.int32)   # the same 32 bits as the device words


def corrupt_sharded(x_all, name: str, severity: int, draws: Optional[dict] = None, seed: int = 0,
                    gather_outputs: bool = False, corrupt_fn=None):
    """One (corruption, severity) pass over this rank's shard of x_all (host or device (n,h,w,3) uint8).
    Returns (local_output, gathered per-image checksums (n,2), gathered outputs or None).  Injected draws are sliced with
    the images, so the sharded result equals the single-GPU result bit for bit; on-chip draws are keyed by (seed, index in
    the launch), so each rank derives its own stream from (seed, first global index)."""
    if corrupt_fn is None:
        from . import cifar_c
        corrupt_fn = cifar_c.corrupt
    n = len(x_all)
    lo, hi = shard_range(n)
    xl = x_all[lo:hi]
    if not isinstance(xl, torch.Tensor):
        xl = torch.from_numpy(xl)
    if torch.cuda.is_available() and not xl.is_cuda:
        xl = xl.cuda(non_blocking=True)
    rank_seed = seed if draws else (seed * 0x9E3779B97F4A7C15 + lo) & 0xFFFFFFFFFFFFFFFF
    if xl.is_cuda and corrupt_fn.__module__.endswith("cifar_c"):
        # the pass kernel's epilogue writes the metric (no second read of the output)
        local = torch.empty((xl.shape[0], 2), dtype=torch.int32, device=xl.device)
        out = corrupt_fn(xl, name, severity, draws=shard_draws(draws, n), seed=rank_seed, metrics=local)
    else:
        out = corrupt_fn(xl, name, severity, draws=shard_draws(draws, n), seed=rank_seed)
        local = image_checksums(out)
    sums = all_gather_rows(local, n)
    gathered = all_gather_rows(out, n) if gather_outputs else None
    return out, sums, gathered
