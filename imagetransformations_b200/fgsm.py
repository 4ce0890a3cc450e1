"""Drop-ins for the sign-step family (attacks/fgsm.py:167-196, attacks/fgsm/cifar-10/fgsm_cifar10.py:25-28,
reverse_fgsm_cifar10.py:74-80, pgd_cifar10_corrupted.py:45-85 and their copies).  The model forward/backward stays
PyTorch; the elementwise step (sign, eps-scale, add, projection, clamp) is one fused libimgtx kernel.
"""
from __future__ import annotations

import torch

from . import batch as B

CIFAR10_MEAN = (0.4914, 0.4822, 0.4465)
CIFAR10_STD = (0.2023, 0.1994, 0.2010)


def _chan_bounds(t, c):
    """(1,C,1,1) tensor / sequence / scalar -> python list of floats (length 1 or C)."""
    if isinstance(t, torch.Tensor):
        v = t.detach().reshape(-1).cpu().tolist()
    elif isinstance(t, (tuple, list)):
        v = [float(a) for a in t]
    else:
        v = [float(t)]
    if len(v) not in (1, c):
        raise ValueError("bounds must be scalar or one per channel")
    return v


def normalized_bounds(mean=CIFAR10_MEAN, std=CIFAR10_STD, device=None):
    """min_vals / max_vals of reverse_fgsm_cifar10.py:71-72."""
    lo = torch.tensor([(0 - m) / s for m, s in zip(mean, std)], device=device).view(1, 3, 1, 1)
    hi = torch.tensor([(1 - m) / s for m, s in zip(mean, std)], device=device).view(1, 3, 1, 1)
    return lo, hi


def fgsm_attack(image, epsilon, data_grad):
    """clamp(image + epsilon*sign(data_grad), 0, 1)."""
    return B.sign_step(image, data_grad, float(epsilon), +1.0, 0.0, 1.0)


def reverse_fgsm_attack(image, epsilon, data_grad, min_pixel_val, max_pixel_val):
    """min(max(image - epsilon*sign(data_grad), min_vals), max_vals) with per-channel (1,3,1,1) bounds."""
    c = image.shape[-3]
    return B.sign_step(image, data_grad, float(epsilon), -1.0, _chan_bounds(min_pixel_val, c),
                       _chan_bounds(max_pixel_val, c))


def fgsm_eps_sweep(image, data_grad, eps_list, direction=+1.0, lo=None, hi=None):
    """E epsilons from ONE pass over image and gradient: returns (E, *image.shape)."""
    return B.sign_step(image, data_grad, [float(e) for e in eps_list], direction, lo, hi)


def FGSM(net, x, y, eps):
    """attacks/fgsm.py:167-196: returns (x_adv, h_adv, y_adv, pert)."""
    x_ = x.detach().clone().requires_grad_(True)
    h_ = net(x_)
    cost = torch.nn.CrossEntropyLoss()(h_, y)
    net.zero_grad()
    cost.backward()
    x_adv, pert = B.sign_step(x_.detach(), x_.grad.detach(), float(eps), +1.0, return_pert=True)
    h_adv = net(x_adv)
    _, y_adv = torch.max(h_adv.data, 1)
    return x_adv, h_adv, y_adv, pert


def pgd_step(perturbed_image, original_image, data_grad, epsilon, alpha, min_pixel_val, max_pixel_val):
    """One PGD iteration (pgd_cifar10_corrupted.py:72-83) as a single kernel."""
    c = perturbed_image.shape[-3]
    return B.pgd_step(perturbed_image, original_image, data_grad, float(alpha), float(epsilon),
                      _chan_bounds(min_pixel_val, c), _chan_bounds(max_pixel_val, c))


def pgd_attack(model, image, label, epsilon, alpha, num_iter, min_pixel_val, max_pixel_val, criterion=None):
    """pgd_cifar10_corrupted.py:45-85: same torch RNG draw (uniform_ init), same iteration structure."""
    criterion = torch.nn.CrossEntropyLoss() if criterion is None else criterion
    original_image = image.clone().detach()
    random_noise = torch.empty_like(original_image).uniform_(-epsilon, epsilon)
    c = image.shape[-3]
    lo, hi = _chan_bounds(min_pixel_val, c), _chan_bounds(max_pixel_val, c)
    # x0 + noise, clamped: a zero-gradient PGD step with an infinite ball is not the same arithmetic, so the init
    # uses the float-corrupt kernel (x + r*1 + 0) followed by the per-channel clamp of a zero-eps sign step.
    perturbed = B.float_corrupt(original_image, B.L.FC_ADD_NOISE, 1.0, 0.0, noise=random_noise, clamp=False)
    perturbed = B.sign_step(perturbed, perturbed, 0.0, +1.0, lo, hi)
    for _ in range(num_iter):
        perturbed.requires_grad = True
        output = model(perturbed)
        loss = criterion(output, label)
        model.zero_grad()
        loss.backward()
        data_grad = perturbed.grad.data
        perturbed = B.pgd_step(perturbed.detach(), original_image, data_grad, float(alpha), float(epsilon), lo, hi)
    return perturbed


def patch_step_(patch: torch.Tensor, grad: torch.Tensor, epsilon: float) -> torch.Tensor:
    """fall_2025/angellic_patches_fgsm.py:67-69, in place like the reference:  patch -= eps*grad.sign(); patch.clamp_(0, 1).
    The same step without the in-place update is the entropy / margin descent of pipenline/cifar100_shannon.py:72-74."""
    B.sign_step(patch, grad, float(epsilon), -1.0, 0.0, 1.0, out=patch)
    return patch


def odin_perturb(imgs: torch.Tensor, grads: torch.Tensor, epsilon: float) -> torch.Tensor:
    """fall_2025/generalizedOdin2.py:47,76:  imgs - eps*torch.sign(-grads).  a - (-b) == a + b bit for bit in IEEE arithmetic
    (sign(-g) == -sign(g) including zeros and NaN), so this is the un-clamped forward sign step."""
    return B.sign_step(imgs, grads, float(epsilon), +1.0)

