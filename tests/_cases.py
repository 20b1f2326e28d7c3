"""Seeded synthetic inputs shared by the parity tests (CPU tensors; tests move them)."""
from __future__ import annotations

import math

import torch


def layered_model(nz, ny, seed=0, vmin=1500.0, vmax=3500.0, blob=True, noise=20.0):
    """Three horizontal layers + a Gaussian blob + small random perturbation."""
    g = torch.Generator().manual_seed(seed)
    v = torch.empty(nz, ny, dtype=torch.float64)
    z = torch.arange(nz, dtype=torch.float64)[:, None].expand(nz, ny)
    y = torch.arange(ny, dtype=torch.float64)[None, :].expand(nz, ny)
    v[:] = vmin
    v[z >= nz // 3] = 0.5 * (vmin + vmax)
    v[z >= 2 * nz // 3] = vmax
    if blob:
        v = v + 400.0 * torch.exp(-((z - 0.6 * nz) ** 2 + (y - 0.5 * ny) ** 2) / (2 * 8.0 ** 2))
    if noise:
        v = v + noise * torch.randn(nz, ny, generator=g, dtype=torch.float64)
    return v.float()


def ricker(freq, nt, dt, peak):
    t = torch.arange(nt, dtype=torch.float64) * dt - peak
    a = (math.pi * freq) ** 2 * t * t
    return ((1 - 2 * a) * torch.exp(-a)).float()


def survey(nz, ny, dx, S, R, Ns=1, z_src=0, z_rec=0, seed=0, scattered=False):
    """x_s [S, Ns, 2], x_r [S, R, 2] in metres (z, y), on cell centres."""
    g = torch.Generator().manual_seed(seed + 7)
    x_s = torch.zeros(S, Ns, 2)
    x_r = torch.zeros(S, R, 2)
    if scattered:
        x_s[..., 0] = torch.randint(0, nz, (S, Ns), generator=g).float() * dx
        x_s[..., 1] = torch.randint(0, ny, (S, Ns), generator=g).float() * dx
        x_r[..., 0] = torch.randint(0, nz, (S, R), generator=g).float() * dx
        x_r[..., 1] = torch.randint(0, ny, (S, R), generator=g).float() * dx
    else:
        for s in range(S):
            for j in range(Ns):
                x_s[s, j, 0] = z_src * dx
                x_s[s, j, 1] = float(int((s + 0.5) * ny / S + 3 * j) % ny) * dx
        yr = torch.clamp(torch.round(torch.linspace(0, ny, R)), max=ny - 1)
        x_r[:, :, 0] = z_rec * dx
        x_r[:, :, 1] = yr[None, :] * dx
    return x_s, x_r


def rel_l2(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-300))
