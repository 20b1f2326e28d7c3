"""Host-side preparation for the propagators (runs as a handful of torch ops on the model's device).

Mirrors what deepwave's Python layer does before it enters its native time loop for the call
``Propagator({'vp': v}, dx)(src, x_s, x_r, dt)`` (reference models/networks.py:5449,5464):
pick the inner time step from the CFL limit, resample the wavelet, pad the model, build the
PML profiles and turn metre coordinates into cells.  Conventions: SURVEY.md section 7.0.
"""
from __future__ import annotations

import math
from typing import Sequence, Tuple

import torch


def spacing(dx) -> Tuple[float, float]:
    """dx as float, 1-sequence or (dz, dy) sequence/tensor -> (dz, dy)."""
    if isinstance(dx, torch.Tensor):
        dx = dx.detach().cpu().reshape(-1).tolist()
    if isinstance(dx, (int, float)):
        return float(dx), float(dx)
    vals = [float(x) for x in dx]
    if len(vals) == 1:
        return vals[0], vals[0]
    if len(vals) != 2:
        raise RuntimeError(f"dx must have 1 or 2 entries for a 2-D model, got {len(vals)}")
    return vals[0], vals[1]


def inner_step(dt: float, dz: float, dy: float, vmax: float) -> Tuple[int, float]:
    """Number of inner steps per user step and the inner step: dt_max = 0.6/(|1/d| vmax)."""
    if not (vmax > 0.0) or not math.isfinite(vmax):
        raise RuntimeError(f"maximum velocity must be positive and finite, got {vmax}")
    limit = 0.6 / (math.sqrt(1.0 / (dz * dz) + 1.0 / (dy * dy)) * vmax)
    ratio = int(math.ceil(abs(dt) / limit - 1e-12))
    ratio = max(ratio, 1)
    return ratio, dt / ratio


def sigma_profile(n: int, halo: int, pml_width: int, d: float, vmax: float) -> torch.Tensor:
    """Quadratic PML ramp over a padded axis of n cells (float64, CPU)."""
    sig = torch.zeros(n, dtype=torch.float64)
    if pml_width <= 0:
        return sig
    length = pml_width * d
    top = 3.0 * vmax * math.log(1000.0) / (2.0 * length)
    idx = torch.arange(n, dtype=torch.float64)
    below = (halo + pml_width) - idx
    above = idx - (n - halo - pml_width - 1)
    depth = torch.maximum(below, above).clamp_(0.0, float(pml_width))
    return top * (depth / pml_width) ** 2


def resample_up(x: torch.Tensor, ratio: int) -> torch.Tensor:
    """Band-limited upsampling of dim 0 by an integer factor (Fourier zero padding, the
    scipy.signal.resample convention).  Differentiable; linear in x."""
    if ratio == 1:
        return x
    n = x.shape[0]
    m = n * ratio
    spec = torch.fft.rfft(x, dim=0)
    if n % 2 == 0:
        # the Nyquist bin of the short signal is split between +-n/2 of the long one
        # (no host scalar written into a device tensor: this runs inside CUDA-graph captures)
        spec = torch.cat([spec[:-1], spec[-1:] * 0.5], dim=0)
    pad = torch.zeros((m // 2 + 1 - spec.shape[0],) + tuple(x.shape[1:]), dtype=spec.dtype,
                      device=x.device)
    return torch.fft.irfft(torch.cat([spec, pad], dim=0), n=m, dim=0) * ratio


def resample_up_adjoint(g: torch.Tensor, ratio: int, like: torch.Tensor) -> torch.Tensor:
    """Transpose of resample_up applied to g [n*ratio, ...] -> [n, ...]."""
    if ratio == 1:
        return g
    with torch.enable_grad():
        x = torch.zeros_like(like, dtype=g.dtype).requires_grad_(True)
        y = resample_up(x, ratio)
        (gx,) = torch.autograd.grad(y, x, g)
    return gx


def to_cells(x: torch.Tensor, dz: float, dy: float, mode: str = "trunc") -> torch.Tensor:
    """Metres [..., 2] as (z, y) -> integer model cells [..., 2] (int64), on x's device."""
    s = torch.stack([x[..., 0].double() / dz, x[..., 1].double() / dy], dim=-1)
    if mode == "trunc":
        # arange(n)*dx in fp32 can land one ulp under the integer; nudge before flooring
        c = torch.floor(s + 1e-6)
    elif mode == "round":
        c = torch.round(s)
    else:
        raise RuntimeError(f"unknown loc_mode {mode!r}")
    return c.long()


def replicate_pad_adjoint(gpad: torch.Tensor, pad: int) -> torch.Tensor:
    """Transpose of replicate padding by `pad` cells on the four sides of a 2-D array."""
    if pad == 0:
        return gpad
    g = gpad.clone()
    g[pad, :] += g[:pad, :].sum(0)
    g[-pad - 1, :] += g[-pad:, :].sum(0)
    g = g[pad:-pad, :]
    g[:, pad] += g[:, :pad].sum(1)
    g[:, -pad - 1] += g[:, -pad:].sum(1)
    return g[:, pad:-pad].contiguous()


def injection_lists(cells_dev: torch.Tensor, nzp: int, pitch: int, keep=None):
    """Sorts the per-shot point list by device cell so a work unit finds its points by row range.

    cells_dev: [S, n] device cell (z*pitch + col).  Returns (row_start [S, nzp+1] int32 absolute
    offsets into the flat lists, cell [S*n] int32 sorted by (shot, cell), col [S*n] int32 = index
    of the entry within its shot before sorting).  The sort is stable, so duplicate points stay
    adjacent and in their original order: the kernel applies them in that fixed order.
    `keep` [S, n] bool: points to leave out of the lists altogether (they sort behind the last
    row of the last shot, where no row range reaches them).
    """
    S, n = cells_dev.shape
    dev = cells_dev.device
    if n == 0:
        return (torch.zeros(S * (nzp + 1), dtype=torch.int32, device=dev),
                torch.zeros(1, dtype=torch.int32, device=dev), torch.zeros(1, dtype=torch.int32, device=dev))
    per_shot = nzp * pitch
    key = (torch.arange(S, device=dev)[:, None] * per_shot + cells_dev).reshape(-1)
    if keep is not None:
        key = torch.where(keep.reshape(-1).to(dev), key, torch.full_like(key, S * per_shot))
    skey, order = torch.sort(key, stable=True)
    bounds = (torch.arange(S, device=dev)[:, None] * per_shot
              + torch.arange(nzp + 1, device=dev)[None, :] * pitch).reshape(-1)
    start = torch.searchsorted(skey, bounds, right=False).to(torch.int32)
    cell = cells_dev.reshape(-1)[order].to(torch.int32).contiguous()
    col = (order % n).to(torch.int32).contiguous()
    return start.contiguous(), cell, col


def duplicate_links(cells: torch.Tensor) -> torch.Tensor:
    """Chains the points of each shot that share a cell, for deterministic parallel injection.

    cells [S, n] -> link [S, n] int32: bits 0..29 = index (within the shot) of the next point in the
    same cell + 1, 0 if none; bit 30 set if the point is not the first of its chain.  Chains follow
    the original list order (stable sort)."""
    S, n = cells.shape
    dev = cells.device
    if n == 0:
        return torch.zeros(S, 0, dtype=torch.int32, device=dev)
    c = cells.long()
    order = torch.argsort(c, dim=1, stable=True)
    cs = torch.gather(c, 1, order)
    same_next = torch.zeros(S, n, dtype=torch.bool, device=dev)
    same_next[:, :-1] = cs[:, 1:] == cs[:, :-1]
    same_prev = torch.zeros(S, n, dtype=torch.bool, device=dev)
    same_prev[:, 1:] = same_next[:, :-1]
    nxt = torch.zeros(S, n, dtype=torch.long, device=dev)
    nxt[:, :-1] = order[:, 1:] + 1
    link_sorted = torch.where(same_next, nxt, torch.zeros_like(nxt)) + same_prev.long() * 0x40000000
    link = torch.zeros(S, n, dtype=torch.long, device=dev)
    link.scatter_(1, order, link_sorted)
    return link.to(torch.int32).contiguous()
