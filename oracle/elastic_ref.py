"""CPU oracle for the 2-D P-SV elastic propagator: pure-PyTorch restatement (autograd gradient).

TEST INFRASTRUCTURE ONLY.  Nothing under ``physicsbasedfwi2_b200/`` may import this file.

PARITY UNPINNED: the reference's elastic path is the external C/MPI code DENISE-Black-Edition
driven through ``pyapi_denise`` (``models/networks.py:31``, ``api.Denise(...)`` ``:8986``,
``d.grad(model_init, src, rec)`` ``:9173``, ``d.get_fwi_gradients`` ``:9185``); it is not under
/root/reference, has no pinned version and cannot run here (needs MPI and its own binaries).  The
reference holds no test or golden vector for it.  This file restates the published scheme DENISE
implements (Virieux/Levander velocity-stress staggered grid, C-PML after Komatitsch & Martin) as
fixed in SURVEY.md section 7.0, behind the deepwave-style call the north star asks for:

    Propagator({'vp': vp, 'vs': vs, 'rho': rho}, dx)(source_amplitudes, x_s, x_r, dt)
        -> receiver_amplitudes [2, nt, n_shots, n_receivers]     (0: v_y horizontal, 1: v_z vertical)

anchored on the reference's call-site contract: physics/geometry parameters at
models/networks.py:8994-9048, 9080-9147 (PSV, explosive sources, two-component receivers
``seis_x``/``seis_y`` :9054-9061, gradients for vp, vs, rho :9185-9188).

Scheme (per shot; z = dim 0 depth, y = dim 1 horizontal; staggered nodes):
    sxx, szz at (i, j);  vy at (i, j+1/2);  vz at (i+1/2, j);  syz at (i+1/2, j+1/2)
    D+ f[i] = sum_k c_k (f[i+k] - f[i-k+1]) / d      D- f[i] = sum_k c_k (f[i+k-1] - f[i-k]) / d
  every derivative carries a C-PML memory variable  psi <- b psi + a Df,  Df <- Df + psi
    F1:  vy += dt by (D+y sxx + D-z syz)          vz += dt bz (D-y syz + D+z szz)
    F2:  sxx += dt (l2m D-y vy + lam D-z vz)      szz += dt (lam D-y vy + l2m D-z vz)
         syz += dt mu_yz (D+z vy + D+y vz)
         sxx, szz [source cells] += dt s(t)       (explosive source)
    receivers record vy, vz before the step.
Effective parameters (computed with torch so autograd chains to vp, vs, rho):
    l2m = rho vp^2, lam = rho (vp^2 - 2 vs^2) at (i, j);  mu_yz = harmonic 4-point mean of rho vs^2
    (0 where any corner is 0);  by, bz = 1 / (2-point arithmetic mean of rho).
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn.functional as F

from .torch_ref import _pair, cfl_substep, locations_to_cells, upsample

# staggered first-derivative coefficients c_k, k = 1..order/2
STAG = {
    2: [1.0],
    4: [9.0 / 8.0, -1.0 / 24.0],
    6: [75.0 / 64.0, -25.0 / 384.0, 3.0 / 640.0],
    8: [1225.0 / 1024.0, -245.0 / 3072.0, 49.0 / 5120.0, -5.0 / 7168.0],
}


def cpml_profiles(n_pad: int, h: int, pml_width: int, d: float, dt: float, vmax: float, pml_freq: float,
                  half: bool, dtype=torch.float64):
    """(a, b) of the recursive-convolution PML along an axis of n_pad nodes, at integer nodes
    (half=False) or at the half nodes i + 1/2 (half=True)."""
    pos = torch.arange(n_pad, dtype=torch.float64) + (0.5 if half else 0.0)
    if pml_width <= 0:
        return torch.zeros(n_pad, dtype=dtype), torch.ones(n_pad, dtype=dtype)
    lo = (h + pml_width) - pos
    hi = pos - (n_pad - h - pml_width - 1)
    depth = torch.clamp(torch.maximum(lo, hi), min=0.0, max=float(pml_width)) / pml_width
    d0 = 3.0 * vmax * math.log(1000.0) / (2.0 * pml_width * d)
    dd = d0 * depth ** 2
    alpha = math.pi * pml_freq * (1.0 - depth)
    b = torch.exp(-(dd + alpha) * dt)
    a = torch.where(dd > 0, dd / (dd + alpha).clamp_min(1e-30) * (b - 1.0), torch.zeros_like(b))
    return a.to(dtype), b.to(dtype)


def effective_parameters(vp_pad, vs_pad, rho_pad):
    """(l2m, lam, mu_yz, by, bz) on the padded grid."""
    mu = rho_pad * vs_pad * vs_pad
    l2m = rho_pad * vp_pad * vp_pad
    lam = l2m - 2.0 * mu

    def sh(x, dz, dy):   # x[i + dz, j + dy] with edge clamp
        if dz:
            x = torch.cat([x[1:], x[-1:]], 0)
        if dy:
            x = torch.cat([x[:, 1:], x[:, -1:]], 1)
        return x
    m00, m10, m01, m11 = mu, sh(mu, 1, 0), sh(mu, 0, 1), sh(mu, 1, 1)
    ok = (m00 > 0) & (m10 > 0) & (m01 > 0) & (m11 > 0)
    one = torch.ones_like(mu)
    inv = 1.0 / torch.where(ok, m00, one) + 1.0 / torch.where(ok, m10, one) + 1.0 / torch.where(ok, m01, one) \
        + 1.0 / torch.where(ok, m11, one)
    mu_yz = torch.where(ok, 4.0 / inv, torch.zeros_like(mu))
    by = 2.0 / (rho_pad + sh(rho_pad, 0, 1))
    bz = 2.0 / (rho_pad + sh(rho_pad, 1, 0))
    return l2m, lam, mu_yz, by, bz


def _dp(f, dim, h, c, inv_d):
    """D+ along dim on the computational region (result [.., nz-2h, ny-2h])."""
    nz, ny = f.shape[-2], f.shape[-1]
    out = None
    for k in range(1, h + 1):
        if dim == 0:
            t = f[..., h + k:nz - h + k, h:ny - h] - f[..., h - k + 1:nz - h - k + 1, h:ny - h]
        else:
            t = f[..., h:nz - h, h + k:ny - h + k] - f[..., h:nz - h, h - k + 1:ny - h - k + 1]
        t = (c[k - 1] * inv_d) * t
        out = t if out is None else out + t
    return out


def _dm(f, dim, h, c, inv_d):
    """D- along dim on the computational region."""
    nz, ny = f.shape[-2], f.shape[-1]
    out = None
    for k in range(1, h + 1):
        if dim == 0:
            t = f[..., h + k - 1:nz - h + k - 1, h:ny - h] - f[..., h - k:nz - h - k, h:ny - h]
        else:
            t = f[..., h:nz - h, h + k - 1:ny - h + k - 1] - f[..., h:nz - h, h - k:ny - h - k]
        t = (c[k - 1] * inv_d) * t
        out = t if out is None else out + t
    return out


def elastic_forward(vp, vs, rho, dx, source_amplitudes, x_s, x_r, dt, *, order: int = 4, pml_width: int = 10,
                    vpmax: Optional[float] = None, pml_freq: float = 25.0, loc_mode: str = "trunc",
                    source: str = "explosive", free_surface: bool = False):
    """Differentiable forward model (vp, vs, rho, source_amplitudes) -> [2, nt, S, R] (vy, vz).
    source: 'explosive' (sxx, szz += dt s; DENISE QUELLTYPB 1, the reference's default), 'fz' (vertical force:
    vz += dt bz s; QUELLTYPB 2, models/networks.py:10452) or 'fy' (horizontal force: vy += dt by s)."""
    if source not in ("explosive", "fz", "fy"):
        raise RuntimeError(f"unknown source kind {source!r}")
    # free_surface: stress-imaging free surface at the first model row (DENISE FREE_SURF = 1, models/networks.py:9811;
    # Levander 1988): szz = 0 on that row, szz and syz antisymmetric about it (the derivatives of F1 read mirrored values
    # above it), the sxx update of the surface row uses l2m - lam^2 / l2m (szz = 0 eliminates e2), nothing above the
    # surface is updated (velocities there stay zero).  Replaces the top C-PML.
    dtype = vp.dtype
    dz, dy = _pair(dx)
    nz, ny = vp.shape
    h = order // 2
    P = pml_width + h
    c = STAG[order]
    vmax = float(vpmax) if vpmax is not None else float(vp.detach().abs().max())
    ratio, dti = cfl_substep(dt, dz, dy, vmax)

    def pad(x):
        return F.pad(x[None, None], (P, P, P, P), mode="replicate")[0, 0]
    vp_p, vs_p, rho_p = pad(vp), pad(vs), pad(rho)
    nzp, nyp = vp_p.shape
    l2m, lam, mu_yz, by, bz = effective_parameters(vp_p, vs_p, rho_p)
    I = (slice(h, nzp - h), slice(h, nyp - h))
    a_z, b_z = cpml_profiles(nzp, h, pml_width, dz, dti, vmax, pml_freq, False, dtype)
    a_zh, b_zh = cpml_profiles(nzp, h, pml_width, dz, dti, vmax, pml_freq, True, dtype)
    a_y, b_y = cpml_profiles(nyp, h, pml_width, dy, dti, vmax, pml_freq, False, dtype)
    a_yh, b_yh = cpml_profiles(nyp, h, pml_width, dy, dti, vmax, pml_freq, True, dtype)
    az, bzp, azh, bzh = (t[h:nzp - h, None] for t in (a_z, b_z, a_zh, b_zh))
    ay, byp, ayh, byh = (t[None, h:nyp - h] for t in (a_y, b_y, a_yh, b_yh))

    nt, S, Ns = source_amplitudes.shape
    sc = locations_to_cells(x_s, dz, dy, P, nz, ny, loc_mode)
    rc = locations_to_cells(x_r, dz, dy, P, nz, ny, loc_mode)
    s_lin = sc[..., 0] * nyp + sc[..., 1]
    r_lin = rc[..., 0] * nyp + rc[..., 1]
    amp = upsample(source_amplitudes.to(dtype), ratio)
    nti = nt * ratio

    z = lambda: torch.zeros(S, nzp, nyp, dtype=dtype)   # noqa: E731
    vy, vz, sxx, szz, syz = z(), z(), z(), z(), z()
    zi = lambda: torch.zeros(S, nzp - 2 * h, nyp - 2 * h, dtype=dtype)   # noqa: E731
    psi = [zi() for _ in range(8)]
    pad4 = (h, h, h, h)
    l2m_i, lam_i, mu_i, by_i, bz_i = l2m[I], lam[I], mu_yz[I], by[I], bz[I]
    EI = (Ellipsis,) + I
    rf = P - h                                    # surface row inside the computational region (free_surface)
    if free_surface:
        live = torch.zeros(nzp - 2 * h, 1, dtype=dtype)
        live[rf:] = 1.0                           # rows at and below the surface
        surf = torch.zeros(nzp - 2 * h, 1, dtype=dtype)
        surf[rf] = 1.0
        below = live - surf
        c_surf = l2m_i - lam_i * lam_i / l2m_i

    def mirrored(szz_, syz_):
        """Copies of szz / syz with the h rows above the surface filled antisymmetrically."""
        a, b = szz_.clone(), syz_.clone()
        for m in range(1, h + 1):
            a[:, P - m] = -szz_[:, P + m]
            b[:, P - m] = -syz_[:, P + m - 1]
        return a, b

    rec_y, rec_z = [], []
    for n in range(nti):
        if n % ratio == 0:
            rec_y.append(vy.reshape(S, -1).gather(1, r_lin))
            rec_z.append(vz.reshape(S, -1).gather(1, r_lin))
        # F1: velocities from stresses
        szz_e, syz_e = mirrored(szz, syz) if free_surface else (szz, syz)
        d1 = _dp(sxx, 1, h, c, 1.0 / dy); psi[0] = byh * psi[0] + ayh * d1
        d2 = _dm(syz_e, 0, h, c, 1.0 / dz); psi[1] = bzp * psi[1] + az * d2
        fy = (d1 + psi[0]) + (d2 + psi[1])
        d3 = _dm(syz, 1, h, c, 1.0 / dy); psi[2] = byp * psi[2] + ay * d3
        d4 = _dp(szz_e, 0, h, c, 1.0 / dz); psi[3] = bzh * psi[3] + azh * d4
        fz = (d3 + psi[2]) + (d4 + psi[3])
        if free_surface:
            psi[0], psi[1], psi[2], psi[3] = (live * q for q in psi[:4])
            fy, fz = live * fy, live * fz
        vy = vy + F.pad((dti * by_i) * fy, pad4)
        vz = vz + F.pad((dti * bz_i) * fz, pad4)
        # F2: stresses from the new velocities
        d5 = _dm(vy, 1, h, c, 1.0 / dy); psi[4] = byp * psi[4] + ay * d5
        d6 = _dm(vz, 0, h, c, 1.0 / dz); psi[5] = bzp * psi[5] + az * d6
        e1, e2 = d5 + psi[4], d6 + psi[5]
        d7 = _dp(vy, 0, h, c, 1.0 / dz); psi[6] = bzh * psi[6] + azh * d7
        d8 = _dp(vz, 1, h, c, 1.0 / dy); psi[7] = byh * psi[7] + ayh * d8
        e3 = (d7 + psi[6]) + (d8 + psi[7])
        if free_surface:
            psi[4], psi[5], psi[6], psi[7] = (live * q for q in psi[4:])
            sxx = sxx + F.pad(below * (dti * (l2m_i * e1 + lam_i * e2)) + surf * (dti * (c_surf * e1)), pad4)
            szz = szz + F.pad(below * (dti * (lam_i * e1 + l2m_i * e2)), pad4)
            syz = syz + F.pad(live * ((dti * mu_i) * e3), pad4)
        else:
            sxx = sxx + F.pad(dti * (l2m_i * e1 + lam_i * e2), pad4)
            szz = szz + F.pad(dti * (lam_i * e1 + l2m_i * e2), pad4)
            syz = syz + F.pad((dti * mu_i) * e3, pad4)
        if source == "explosive":
            inj = torch.zeros(S, nzp * nyp, dtype=dtype).scatter_add(1, s_lin, dti * amp[n]).reshape(S, nzp, nyp)
            sxx, szz = sxx + inj, szz + inj
        else:
            b = (bz if source == "fz" else by).reshape(-1)[s_lin]                 # buoyancy at the source cells [S, Ns]
            inj = torch.zeros(S, nzp * nyp, dtype=dtype).scatter_add(1, s_lin, (dti * b) * amp[n]).reshape(S, nzp, nyp)
            if source == "fz":
                vz = vz + inj
            else:
                vy = vy + inj
    return torch.stack([torch.stack(rec_y, 0), torch.stack(rec_z, 0)], 0)


def elastic_gradient(vp, vs, rho, dx, source_amplitudes, x_s, x_r, dt, grad_rec, **kw):
    """(rec, grad_vp, grad_vs, grad_rho, grad_src) for J = <grad_rec, rec>."""
    vp = vp.detach().clone().requires_grad_(True)
    vs = vs.detach().clone().requires_grad_(True)
    rho = rho.detach().clone().requires_grad_(True)
    s = source_amplitudes.detach().clone().requires_grad_(True)
    rec = elastic_forward(vp, vs, rho, dx, s, x_s, x_r, dt, **kw)
    (rec * grad_rec.to(rec.dtype)).sum().backward()
    return rec.detach(), vp.grad, vs.grad, rho.grad, s.grad
