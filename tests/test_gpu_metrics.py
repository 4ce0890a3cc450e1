"""The per-image pass metric (north star: "NCCL ... only to all-gather per-image outputs and metrics"):
{byte sum, position-weighted byte sum mod 2^32} written by the suite kernels' epilogue, by the standalone kernel and by the
host formula must be the same words; over NCCL on two ranks the gathered metrics equal the single-GPU ones."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import cifar_c as CC

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def host_metric(y):
    y = y.reshape(len(y), -1).astype(np.int64)
    w = np.arange(1, y.shape[1] + 1, dtype=np.int64)
    return np.stack([y.sum(1), (y * w).sum(1)], 1) & 0xFFFFFFFF


@pytest.mark.parametrize("mode", ["fast", "exact"])
def test_epilogue_metric_equals_host_formula(cuda, mode):
    from imagetransformations_b200 import cifar_c as C, shard as S
    x = torch.from_numpy(np.random.default_rng(2).integers(0, 256, (777, 32, 32, 3), dtype=np.uint8)).cuda()
    for name in CC.CORRUPTIONS:
        m = torch.full((len(x), 2), -1, dtype=torch.int32, device="cuda")
        y = C.corrupt(x, name, 3, seed=5, mode=mode, metrics=m)
        want = host_metric(y.cpu().numpy())
        got = m.cpu().numpy().astype(np.int64) & 0xFFFFFFFF
        assert np.array_equal(got, want), (name, mode)
        assert np.array_equal(S.image_checksums(y).cpu().numpy().astype(np.int64) & 0xFFFFFFFF, want), name
    # CPU tensors (host logic under gloo) give the same words
    yc = y.cpu()
    assert torch.equal(S.image_checksums(yc), S.image_checksums(y).cpu())
    # odd sizes / unaligned
    z = torch.from_numpy(np.random.default_rng(3).integers(0, 256, (5, 7, 9, 3), dtype=np.uint8)).cuda()
    assert np.array_equal(S.image_checksums(z).cpu().numpy().astype(np.int64) & 0xFFFFFFFF, host_metric(z.cpu().numpy()))


WORKER = r'''
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from imagetransformations_b200 import cifar_c as C, shard as S
from oracle import cifar_c as CC
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
x = np.random.default_rng(7).integers(0, 256, (1001, 32, 32, 3), dtype=np.uint8)
ok = True
for name, sev in (("contrast", 2), ("glass_blur", 4), ("jpeg_compression", 3), ("elastic_transform", 3)):
    d = CC.draw(name, sev, x, np.random.RandomState(3))
    out, sums, gathered = S.corrupt_sharded(x, name, sev, draws=d, gather_outputs=True)
    full = C.corrupt(torch.from_numpy(x).cuda(), name, sev, draws=d)
    ok &= torch.equal(gathered, full) and torch.equal(sums, S.image_checksums(full))
dist.barrier()
print("RANK", rank, "OK" if ok else "MISMATCH", flush=True)
dist.destroy_process_group()
sys.exit(0 if ok else 1)
'''


def test_nccl_allgather_of_metrics_and_outputs_two_ranks(cuda, tmp_path):
    """Two NCCL ranks (one per GPU): every rank corrupts its shard, the epilogue metrics and (optionally) the outputs are
    all-gathered, and both equal the single-GPU result.  Skipped on boxes with fewer than two GPUs."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", str(script), ROOT]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("OK") == 2
