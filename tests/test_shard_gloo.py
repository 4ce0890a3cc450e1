"""Multi-process host logic of the image-sharded path (world_size 2, gloo, CPU): the index ranges tile the batch, injected
draws are cut with the images, and the metric / output all-gathers reassemble rank order, including ragged shards.  The
transform itself is stubbed with a per-image CPU function here (the CUDA path needs a GPU; its sharded == single-GPU
identity is a -m gpu test)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from imagetransformations_b200 import shard as S


def test_shard_ranges_tile_the_batch():
    for n in (0, 1, 2, 7, 50000, 50001):
        for ws in (1, 2, 3, 4, 8):
            rs = [S.shard_range(n, r, ws) for r in range(ws)]
            assert rs[0][0] == 0 and rs[-1][1] == n
            assert all(rs[i][1] == rs[i + 1][0] for i in range(ws - 1))
            assert max(S.shard_sizes(n, ws)) - min(S.shard_sizes(n, ws)) <= 1
    assert S.shard_range(50000, 3, 8) == (18750, 25000)
    with pytest.raises(ValueError):
        S.shard_range(10, 2, 2)


def _stub_corrupt(x, name, severity, draws=None, seed=0):
    """Per-image, order-sensitive stand-in: out = (x * (severity + 1) + per-image draw) mod 256."""
    d = draws["k"] if draws else torch.zeros(len(x), dtype=torch.int64)
    d = torch.as_tensor(np.asarray(d)).reshape(-1, 1, 1, 1).to(x.device)   # the shard may live on a GPU when one is present
    return ((x.to(torch.int64) * (severity + 1) + d) % 256).to(torch.uint8)


def _worker(rank, ws, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        x = np.random.default_rng(3).integers(0, 256, (n, 4, 4, 3), dtype=np.uint8)
        draws = {"k": np.arange(n) * 7 % 251, "shared": 5}
        lo, hi = S.shard_range(n)
        sd = S.shard_draws(draws, n)
        assert len(sd["k"]) == hi - lo and sd["shared"] == 5
        out, sums, gathered = S.corrupt_sharded(x, "stub", 2, draws=draws, gather_outputs=True, corrupt_fn=_stub_corrupt)
        full = _stub_corrupt(torch.from_numpy(x), "stub", 2, draws=draws)
        # on a GPU box the shard lives on cuda:<rank>: compare on the host
        assert torch.equal(out.cpu(), full[lo:hi])
        assert torch.equal(gathered.cpu(), full)
        assert torch.equal(sums.cpu(), S.image_checksums(full).cpu())
        q.put((rank, True))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [10, 11, 1])
def test_sharded_pass_equals_single_process(n):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert sorted(q.get(timeout=5)[0] for _ in range(2)) == [0, 1]


def test_checksum_is_order_sensitive():
    x = torch.arange(2 * 4 * 4 * 3, dtype=torch.int64).reshape(2, 4, 4, 3).to(torch.uint8)
    a = S.image_checksums(x)
    b = S.image_checksums(x.flip(2))
    assert torch.equal(a[:, 0], b[:, 0]) and not torch.equal(a[:, 1], b[:, 1])
