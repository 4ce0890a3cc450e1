"""PIL <-> device helpers for the drop-in (single image) signatures.  The drop-ins exist for call-site
compatibility with the reference; throughput work goes through the batched API (`batch.py`, `cifar_c.py`,
`chains.py`) and is what bench.py times."""
from __future__ import annotations

import numpy as np
import torch
from PIL import Image


def device() -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("imagetransformations_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def to_dev(img) -> torch.Tensor:
    """PIL image / HWC uint8 array -> (1, h, w, c) uint8 CUDA tensor (RGB; RGBA is converted like the reference's
    cv2.COLOR_RGBA2RGB, i.e. alpha dropped)."""
    if isinstance(img, Image.Image):
        if img.mode not in ("RGB", "L"):
            img = img.convert("RGB")
        a = np.array(img)
    else:
        a = np.asarray(img)
    if a.ndim == 2:
        a = a[:, :, None]
    if a.shape[2] == 4:
        a = a[:, :, :3]
    return torch.from_numpy(np.ascontiguousarray(a)).unsqueeze(0).to(device(), non_blocking=False)


def batch_to_dev(imgs) -> torch.Tensor:
    a = np.stack([np.array(i.convert("RGB")) if isinstance(i, Image.Image) else np.asarray(i) for i in imgs])
    return torch.from_numpy(np.ascontiguousarray(a)).to(device())


def to_pil(t: torch.Tensor) -> Image.Image:
    a = t[0].cpu().numpy()
    if a.shape[2] == 1:
        a = a[:, :, 0]
    return Image.fromarray(a)


def batch_to_pil(t: torch.Tensor):
    a = t.cpu().numpy()
    return [Image.fromarray(x) for x in a]
