"""Allocation-free entry point (imgtx_cv_gaussian_blur_noalloc_u8): called through the C ABI, its output equals the
existing imgtx_cv_gaussian_blur_u8 byte for byte (integer arithmetic: tolerance 0, as in tests/test_gpu_ops2.py) on calls
in a row that share the static scratch array, sizes beyond the stated maximum are rejected with a clear error, and the
device's free memory does not move across calls.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from imagetransformations_b200 import _lib as L
    return L, L.load()


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _existing(L, lib, x, ks, sigma):
    n, h, w, c = x.shape
    out = torch.empty_like(x)
    L.check(lib.imgtx_cv_gaussian_blur_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, ks, float(sigma), 0, _stream()))
    return out


def _noalloc(L, lib, x, ks, sigma, out=None):
    n, h, w, c = x.shape
    out = torch.empty_like(x) if out is None else out
    L.check(lib.imgtx_cv_gaussian_blur_noalloc_u8(x.data_ptr(), out.data_ptr(), n, h, w, c, ks, float(sigma), _stream()))
    return out


def _imgs(seed, n, h, w, c=3):
    return torch.from_numpy(np.random.default_rng(seed).integers(0, 256, (n, h, w, c), dtype=np.uint8)).cuda()


def test_noalloc_blur_equals_existing_entry_point_on_calls_in_a_row(cuda):
    L, lib = _lib()
    assert lib.imgtx_noalloc_scratch_bytes() == 32 << 20
    # 224x224: the existing entry point spills its 16-bit intermediate to a cudaMallocAsync workspace; 64x48 and 32x32: it
    # stays in shared memory / on the tensor cores.  The second and third calls find the first call's data in the scratch.
    cases = [(_imgs(1, 5, 224, 224), 13, 2.0), (_imgs(2, 7, 64, 48), 31, 5.0), (_imgs(3, 4, 224, 224), 5, 0.8),
             (_imgs(4, 300, 224, 224), 7, 1.0),      # more images than scratch slots (111 CTAs walk the batch)
             (_imgs(5, 9, 32, 32), 13, 2.0), (_imgs(6, 3, 37, 53, 1), 9, 1.5)]
    for x, ks, sigma in cases:
        want = _existing(L, lib, x, ks, sigma)
        got = _noalloc(L, lib, x, ks, sigma)
        torch.cuda.synchronize()
        assert torch.equal(got, want), (tuple(x.shape), ks, int((got != want).sum()))
    # the same input again, after other shapes have used the scratch
    x, ks, sigma = cases[0]
    assert torch.equal(_noalloc(L, lib, x, ks, sigma), _existing(L, lib, x, ks, sigma))
    # n == 0 is a successful no-op
    assert lib.imgtx_cv_gaussian_blur_noalloc_u8(None, None, 0, 224, 224, 3, 5, 1.0, _stream()) == 0


def test_noalloc_blur_rejects_images_beyond_the_scratch(cuda):
    L, lib = _lib()
    x = torch.zeros(16, dtype=torch.uint8, device="cuda")
    # 2 * 4096 * 2048 * 3 bytes = 48 MiB > 32 MiB: refused before anything is launched or touched
    rc = lib.imgtx_cv_gaussian_blur_noalloc_u8(x.data_ptr(), x.data_ptr(), 1, 4096, 2048, 3, 5, 1.0, _stream())
    assert rc < 0
    msg = lib.imgtx_last_error().decode()
    assert "scratch" in msg and "4096x2048x3" in msg, msg
    with pytest.raises(L.ImgtxError):
        L.check(rc)
    # the largest image that fits (2 * h * w * c == 32 MiB exactly) is accepted
    big = _imgs(7, 1, 4096, 4096, 1)
    got = _noalloc(L, lib, big, 5, 1.0)
    assert torch.equal(got, _existing(L, lib, big, 5, 1.0))


def test_noalloc_blur_does_not_move_free_memory(cuda):
    L, lib = _lib()
    x, out = _imgs(8, 64, 224, 224), None
    out = torch.empty_like(x)
    _noalloc(L, lib, x, 13, 2.0, out)          # first call: module / kernel loading may take device memory
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        _noalloc(L, lib, x, 13, 2.0, out)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    # a workspace for this batch would be 64 x 301 056 B = 19 MB; allow 2 MiB of unrelated driver noise
    assert free0 - free1 <= 2 << 20, (free0, free1)
