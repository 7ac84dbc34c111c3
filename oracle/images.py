"""Deterministic synthetic images (oracle / test infrastructure).

Structured rather than i.i.d. noise — a per-channel mix of sinusoids plus N(0, 0.05) noise, clamped —
because pure-noise images make every random-init classifier predict one class."""
from __future__ import annotations

import numpy as np
import torch


def synthetic_image(seed: int, h: int, w: int) -> np.ndarray:
    """uint8 HWC [h, w, 3], C-contiguous, fully determined by (seed, h, w)."""
    g = torch.Generator().manual_seed(int(seed))
    yy = torch.linspace(0, 1, h).view(h, 1)
    xx = torch.linspace(0, 1, w).view(1, w)
    chans = []
    for _ in range(3):
        f = torch.rand(6, generator=g) * 6 + 0.5          # spatial frequencies
        ph = torch.rand(6, generator=g) * 6.2831853
        a = torch.rand(3, generator=g) * 0.5 + 0.1
        v = (a[0] * torch.sin(6.2831853 * f[0] * xx + ph[0]) * torch.cos(6.2831853 * f[1] * yy + ph[1])
             + a[1] * torch.sin(6.2831853 * (f[2] * xx + f[3] * yy) + ph[2])
             + a[2] * torch.cos(6.2831853 * (f[4] * xx - f[5] * yy) + ph[3]))
        v = 0.5 + 0.5 * v / float(a.sum())
        chans.append(v)
    img = torch.stack(chans, -1) + 0.05 * torch.randn(h, w, 3, generator=g)
    return (img.clamp(0, 1) * 255).round().to(torch.uint8).numpy().copy()


def synthetic_batch(seed: int, n: int, h: int, w: int):
    return [synthetic_image(seed * 100003 + i, h, w) for i in range(n)]
