"""Wavelets with the deepwave call shape the reference uses.

Reference call site: ``deepwave.wavelets.ricker(freq, nt, dt, 1/freq)`` at
models/networks.py:5357 (also :2745, :3092, ...; numpy round trips at :4541, :4949).
"""
from __future__ import annotations

import math
from typing import Optional

import torch


def ricker(freq: float, length: int, dt: float, peak_time: float,
           dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """Ricker wavelet ``(1 - 2 pi^2 f^2 t^2) exp(-pi^2 f^2 t^2)``, ``t = arange(length)*dt - peak_time``.

    Returns a 1-D tensor of ``length`` samples (float32 unless ``dtype`` is given), on the CPU
    like deepwave's; callers move it with ``.to(device)`` (models/networks.py:5360).
    Evaluated in float64 and rounded once.
    """
    if length < 0:
        raise RuntimeError("ricker: length must be non-negative")
    t = torch.arange(int(length), dtype=torch.float64) * float(dt) - float(peak_time)
    arg = (math.pi * float(freq)) ** 2 * t * t
    out = (1.0 - 2.0 * arg) * torch.exp(-arg)
    return out.to(dtype if dtype is not None else torch.float32)
