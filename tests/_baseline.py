"""The BASELINE.json configurations as SURVEY.md 8(d) defines them, for the parity tests.

C1: v[100, 300] three layers 1500/2500/3500 m/s + Gaussian blob (+400 m/s, sigma 8 cells) at (60, 150) +
    N(0, 20 m/s) perturbation (seed 0); dx 10 m; Ricker 15 Hz, peak 1/15 s; dt 1e-3, nt 1000; 10 shots at
    z = 0, y = cells 15, 45, ..., 285; 101 receivers at z = 0, y = cells 0, 3, ..., 300 clipped to 299;
    observed data = forward on the unperturbed layered model; loss = MSE; order 4, pml 10.
C2: v[751, 2301] Marmousi-like synthetic, dx 4 m, order 8, pml 20, dt 4e-4, shots at z = 2 cells,
    576 receivers at z = 2 every 4th cell, 15 Hz Ricker (bench.py's generator: same grid, same survey).
C3: vp as C2's generator at [376, 1151], vs = vp / sqrt(3) with a 24-row water layer, rho = 310 vp^0.25,
    dx 10 m, dt 1e-3, order 4, C-PML 10, explosive sources, two-component receivers.
"""
from __future__ import annotations

import math

import torch

from tests import _cases as K


def c1_model(perturbed=True):
    nz, ny = 100, 300
    z = torch.arange(nz, dtype=torch.float64)[:, None].expand(nz, ny)
    y = torch.arange(ny, dtype=torch.float64)[None, :].expand(nz, ny)
    v = torch.full((nz, ny), 1500.0, dtype=torch.float64)
    v[z >= nz // 3] = 2500.0
    v[z >= 2 * nz // 3] = 3500.0
    if perturbed:
        v = v + 400.0 * torch.exp(-((z - 60.0) ** 2 + (y - 150.0) ** 2) / (2 * 8.0 ** 2))
        v = v + 20.0 * torch.randn(nz, ny, generator=torch.Generator().manual_seed(0), dtype=torch.float64)
    return v.float().contiguous()


def c1(nt=1000, shots=10):
    nz, ny, dx, dt, f = 100, 300, 10.0, 1e-3, 15.0
    R = 101
    x_s = torch.zeros(shots, 1, 2)
    x_s[:, 0, 1] = (15.0 + 30.0 * torch.arange(shots)) * dx
    x_r = torch.zeros(shots, R, 2)
    x_r[:, :, 1] = ((3.0 * torch.arange(R)).clamp(max=ny - 1) * dx)[None, :]
    src = K.ricker(f, nt, dt, 1.0 / f).reshape(-1, 1, 1).repeat(1, shots, 1).contiguous()
    return dict(nz=nz, ny=ny, dx=dx, dt=dt, nt=nt, order=4, pml=10, S=shots, R=R, x_s=x_s, x_r=x_r, src=src,
                v=c1_model(True), v_true=c1_model(False))


def c2(nt=600, shots=(100, 180)):
    """Two shots of the 256-shot survey on the full C2 grid, `nt` of the 3000 steps."""
    import bench
    w = dict(bench.WORKLOADS["c2"])
    v = bench.synth_model(w)
    xs, xr = [], []
    for s in shots:
        a, b = bench.synth_survey(w, 1, first_shot=s)
        xs.append(a)
        xr.append(b)
    x_s, x_r = torch.cat(xs, 0), torch.cat(xr, 0)
    src = K.ricker(w["freq"], nt, w["dt"], 1.0 / w["freq"]).reshape(-1, 1, 1).repeat(1, len(shots), 1).contiguous()
    return dict(nz=w["nz"], ny=w["ny"], dx=w["dx"], dt=w["dt"], nt=nt, order=w["order"], pml=w["pml"], S=len(shots),
                R=w["n_rec"], x_s=x_s, x_r=x_r, src=src, v=v, vmax=w["vmax"],
                v_smooth=smooth(v, 12))


def c3(nt=120):
    """One shot of the 128-shot elastic survey on the full C3 grid."""
    import bench
    w = dict(bench.WORKLOADS["c3"])
    vp = bench.synth_model(w)
    vs = vp / math.sqrt(3.0)
    vs[:24] = 0.0
    rho = 310.0 * vp ** 0.25
    x_s, x_r = bench.synth_survey(w, 1, first_shot=w["survey_shots"] // 2)
    x_s[..., 0] = 30 * w["dx"]          # below the water layer so P and S energy reach the receivers in nt steps
    x_r[..., 0] = 26 * w["dx"]
    src = (K.ricker(25.0, nt, w["dt"], 1.0 / 25.0).reshape(-1, 1, 1) * 1e6).contiguous()
    return dict(nz=w["nz"], ny=w["ny"], dx=w["dx"], dt=w["dt"], nt=nt, order=w["order"], pml=w["pml"], S=1,
                R=w["n_rec"], x_s=x_s, x_r=x_r, src=src, vp=vp, vs=vs, rho=rho, vmax=w["vmax"])


def smooth(v, k):
    """Box smoothing by 2k+1 cells (the 'smooth' starting model of the C2 parity run: true - smooth residual)."""
    p = torch.nn.functional.pad(v[None, None], (k, k, k, k), mode="replicate")
    return torch.nn.functional.avg_pool2d(p, 2 * k + 1, stride=1)[0, 0].contiguous()
