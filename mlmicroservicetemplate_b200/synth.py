"""Deterministic synthetic uploads for the benchmarks (product side: bench.py and benchmarks/ must not lean on the
test oracle for their inputs).  Structured rather than i.i.d. noise — per channel a mix of three sinusoid patterns plus
N(0, 0.05) noise, clamped — because pure-noise images make every random-init classifier predict one class
(SURVEY.md §8d).  Fully determined by (seed, h, w)."""
from __future__ import annotations

from typing import List

import numpy as np
import torch

_TWO_PI = 6.2831853


def synthetic_image(seed: int, h: int, w: int) -> np.ndarray:
    """uint8 HWC [h, w, 3], C-contiguous."""
    gen = torch.Generator().manual_seed(int(seed))
    rows = torch.linspace(0, 1, h).view(h, 1)
    cols = torch.linspace(0, 1, w).view(1, w)
    planes = []
    for _ in range(3):
        freq = torch.rand(6, generator=gen) * 6 + 0.5
        phase = torch.rand(6, generator=gen) * _TWO_PI
        amp = torch.rand(3, generator=gen) * 0.5 + 0.1
        wave = (amp[0] * torch.sin(_TWO_PI * freq[0] * cols + phase[0]) * torch.cos(_TWO_PI * freq[1] * rows + phase[1])
                + amp[1] * torch.sin(_TWO_PI * (freq[2] * cols + freq[3] * rows) + phase[2])
                + amp[2] * torch.cos(_TWO_PI * (freq[4] * cols - freq[5] * rows) + phase[3]))
        planes.append(0.5 + 0.5 * wave / float(amp.sum()))
    img = torch.stack(planes, -1) + 0.05 * torch.randn(h, w, 3, generator=gen)
    return (img.clamp(0, 1) * 255).round().to(torch.uint8).numpy().copy()


def synthetic_batch(seed: int, n: int, h: int, w: int) -> List[np.ndarray]:
    return [synthetic_image(seed * 100003 + i, h, w) for i in range(n)]
