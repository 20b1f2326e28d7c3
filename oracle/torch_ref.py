"""CPU oracle (1/2): pure-PyTorch restatement of the 2-D acoustic propagator.

TEST INFRASTRUCTURE ONLY.  Nothing under ``physicsbasedfwi2_b200/`` may import
this file; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline``
/ ``--impl reference`` legs of ``bench.py`` use it, and only as the checker.

PARITY UNPINNED: the reference repo (/root/reference) holds no propagator
arithmetic.  The path it calls is the third-party package ``deepwave``
(``import deepwave`` at models/networks.py:10; no version pinned anywhere; the
class API ``deepwave.scalar.Propagator({'vp': v}, dx)(src, x_s, x_r, dt)`` used
at models/networks.py:5408-5411,5449,5464 exists only in deepwave v0.0.1-0.0.9).
deepwave is absent from this image and cannot be installed (no network), and
the reference has no test or golden vector for this path.  This file restates
the published scheme of that API generation as recorded in SURVEY.md section 7.0
(4th-order-default centred differences, 2nd-order leapfrog, Grote-Sim style
auxiliary-variable PML on all four sides, replicate padding, CFL sub-stepping
with FFT wavelet upsampling) and anchors parity on the reference's call sites:

* geometry / wavelet / layouts ....... models/networks.py:5339-5358, 5406-5411
* forward -> loss -> backward ........ models/networks.py:5445-5476, 5491-5493
* wavelet gradient ................... models/networks.py:13235-13237
* inner-loop FWI (Adam on v) ......... models/networks.py:15372-15415
* in-repo statement of an acoustic scheme (Devito, unused):
  seisgan/fwi/pde/seismic/acoustic/operators.py:24-51 (update), :54-89
  (inject src*dt^2/m, sample receivers), :127-165 (adjoint + imaging)

The gradient here comes from ``torch.autograd`` applied to the time loop, so it
is the exact discrete adjoint by construction; ``oracle/cpu_ref.c`` holds the
hand-derived adjoint and must agree with this file to round-off in fp64.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

# Central-difference coefficients, index k = offset 0..order/2.
# D2[order][0] is the centre weight of the 2nd derivative.
D2 = {
    2: [-2.0, 1.0],
    4: [-5.0 / 2.0, 4.0 / 3.0, -1.0 / 12.0],
    6: [-49.0 / 18.0, 3.0 / 2.0, -3.0 / 20.0, 1.0 / 90.0],
    8: [-205.0 / 72.0, 8.0 / 5.0, -1.0 / 5.0, 8.0 / 315.0, -1.0 / 560.0],
}
# D1[order][k-1] multiplies (f[+k] - f[-k]).
D1 = {
    2: [1.0 / 2.0],
    4: [2.0 / 3.0, -1.0 / 12.0],
    6: [3.0 / 4.0, -3.0 / 20.0, 1.0 / 60.0],
    8: [4.0 / 5.0, -1.0 / 5.0, 4.0 / 105.0, -1.0 / 280.0],
}


def ricker(freq: float, length: int, dt: float, peak_time: float,
           dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """deepwave.wavelets.ricker as called at models/networks.py:5357."""
    t = torch.arange(length, dtype=torch.float64) * dt - peak_time
    a = (math.pi * freq) ** 2 * t * t
    y = (1.0 - 2.0 * a) * torch.exp(-a)
    return y.to(dtype or torch.float32)


def _pair(dx) -> Tuple[float, float]:
    if isinstance(dx, (int, float)):
        return float(dx), float(dx)
    if isinstance(dx, torch.Tensor):
        dx = dx.tolist()
        if not isinstance(dx, list):
            return float(dx), float(dx)
    dx = list(dx)
    if len(dx) == 1:
        return float(dx[0]), float(dx[0])
    return float(dx[0]), float(dx[1])


def cfl_substep(dt: float, dz: float, dy: float, vmax: float) -> Tuple[int, float]:
    """dt_max = 0.6 / (sqrt(1/dz^2 + 1/dy^2) * vmax)."""
    dt_max = 0.6 / (math.sqrt(1.0 / dz ** 2 + 1.0 / dy ** 2) * vmax)
    ratio = max(1, int(math.ceil(abs(dt) / dt_max - 1e-12)))
    return ratio, dt / ratio


def pml_profile(n_pad: int, h: int, pml_width: int, d: float, vmax: float,
                dtype=torch.float64) -> torch.Tensor:
    """sigma(i) over padded index i; quadratic ramp, sigma_max = 3 vmax ln(1000) / (2 L)."""
    sig = torch.zeros(n_pad, dtype=torch.float64)
    if pml_width > 0:
        L = pml_width * d
        smax = 3.0 * vmax * math.log(1000.0) / (2.0 * L)
        i = torch.arange(n_pad, dtype=torch.float64)
        lo = (h + pml_width) - i                      # depth into the low-side layer (cells)
        hi = i - (n_pad - h - pml_width - 1)          # depth into the high-side layer
        depth = torch.clamp(torch.maximum(lo, hi), min=0.0, max=float(pml_width))
        sig = smax * (depth / pml_width) ** 2
    return sig.to(dtype)


def upsample(x: torch.Tensor, ratio: int) -> torch.Tensor:
    """FFT resampling along dim 0 with scipy.signal.resample semantics (differentiable)."""
    if ratio == 1:
        return x
    n = x.shape[0]
    X = torch.fft.rfft(x, dim=0)
    m = n * ratio
    if n % 2 == 0:
        # the old Nyquist bin is shared between +n/2 and -n/2 of the longer spectrum
        scale = torch.ones(n // 2 + 1, dtype=x.dtype)
        scale[-1] = 0.5
        X = X * scale.reshape((-1,) + (1,) * (x.dim() - 1))
    tail = torch.zeros((m // 2 + 1 - X.shape[0],) + tuple(x.shape[1:]), dtype=X.dtype)
    Y = torch.cat([X, tail], dim=0)
    return torch.fft.irfft(Y, n=m, dim=0) * ratio


def locations_to_cells(x: torch.Tensor, dz: float, dy: float, pad: int,
                       nz: int, ny: int, mode: str = "trunc") -> torch.Tensor:
    """metres [..., 2] (z, y) -> padded integer cell [..., 2]."""
    s = torch.stack([x[..., 0].double() / dz, x[..., 1].double() / dy], dim=-1)
    # fp32 inputs such as 30.0/10.0 are exact; guard 1-ulp-low products of arange()*dx
    c = torch.floor(s + 1e-6) if mode == "trunc" else torch.round(s)
    c = c.long()
    if (c[..., 0].min() < 0 or c[..., 0].max() >= nz or
            c[..., 1].min() < 0 or c[..., 1].max() >= ny):
        raise RuntimeError("source/receiver location outside the model")
    return c + pad


def _d1(f: torch.Tensor, dim: int, h: int, c1: Sequence[float], inv_d: float) -> torch.Tensor:
    """First derivative on the computational region [h:-h, h:-h] of a [S, nzp, nyp] field."""
    nz, ny = f.shape[-2], f.shape[-1]
    out = None
    for k in range(1, h + 1):
        if dim == 0:
            p = f[..., h + k:nz - h + k, h:ny - h]
            m = f[..., h - k:nz - h - k, h:ny - h]
        else:
            p = f[..., h:nz - h, h + k:ny - h + k]
            m = f[..., h:nz - h, h - k:ny - h - k]
        term = (c1[k - 1] * inv_d) * (p - m)
        out = term if out is None else out + term
    return out


def _lap(f: torch.Tensor, h: int, c2: Sequence[float], inv_dz2: float, inv_dy2: float) -> torch.Tensor:
    nz, ny = f.shape[-2], f.shape[-1]
    c = f[..., h:nz - h, h:ny - h]
    out = (c2[0] * inv_dz2 + c2[0] * inv_dy2) * c
    for k in range(1, h + 1):
        out = out + (c2[k] * inv_dy2) * (f[..., h:nz - h, h - k:ny - h - k] + f[..., h:nz - h, h + k:ny - h + k])
    for k in range(1, h + 1):
        out = out + (c2[k] * inv_dz2) * (f[..., h - k:nz - h - k, h:ny - h] + f[..., h + k:nz - h + k, h:ny - h])
    return out


def acoustic_forward(v: torch.Tensor, dx, source_amplitudes: torch.Tensor,
                     x_s: torch.Tensor, x_r: torch.Tensor, dt: float, *,
                     order: int = 4, pml_width: int = 10, vpmax: Optional[float] = None,
                     loc_mode: str = "trunc", return_wavefield: bool = False):
    """Differentiable forward model  (v, source_amplitudes) -> receiver_amplitudes.

    v [nz, ny]; source_amplitudes [nt, S, Ns]; x_s [S, Ns, 2]; x_r [S, R, 2] (metres, (z, y));
    returns [nt, S, R].  Follows SURVEY.md section 7.0:

        q      = Lap(u) + Dz(phi_z) + Dy(phi_y)
        u+     = a [ 2u - b u-  + dt^2 ( v^2 q - sz sy u ) ] + v^2 dt^2 S s(t)
        phi_z+ = (1 - dt sz) phi_z - dt (sz - sy) Dz(u)
        phi_y+ = (1 - dt sy) phi_y - dt (sy - sz) Dy(u)
        r(t)   = u[receivers]   sampled before the step, every `ratio`-th inner step
    with a = 1/(1 + dt (sz+sy)/2), b = 1 - dt (sz+sy)/2.
    """
    dtype = v.dtype
    dz, dy = _pair(dx)
    nz, ny = v.shape
    h = order // 2
    P = pml_width + h
    c2, c1 = D2[order], D1[order]
    vmax = float(vpmax) if vpmax is not None else float(v.detach().abs().max())
    ratio, dti = cfl_substep(dt, dz, dy, vmax)

    v_pad = F.pad(v[None, None], (P, P, P, P), mode="replicate")[0, 0]
    nzp, nyp = v_pad.shape
    w = (v_pad * v_pad) * (dti * dti)                                  # v^2 dt^2
    sz = pml_profile(nzp, h, pml_width, dz, vmax, dtype)[:, None]
    sy = pml_profile(nyp, h, pml_width, dy, vmax, dtype)[None, :]
    ssum = sz + sy
    a = 1.0 / (1.0 + dti * ssum / 2.0)
    c_u = a * (2.0 - dti * dti * sz * sy)                              # multiplies u
    c_um = -a * (1.0 - dti * ssum / 2.0)                               # multiplies u-
    ez, fz = 1.0 - dti * sz, -dti * (sz - sy)
    ey, fy = 1.0 - dti * sy, -dti * (sy - sz)
    I = (slice(h, nzp - h), slice(h, nyp - h))
    aw_i, cu_i, cum_i = (a * w)[I], c_u[I].expand(nzp - 2 * h, nyp - 2 * h), c_um[I]
    ez_i, fz_i, ey_i, fy_i = ez.expand(nzp, nyp)[I], fz[I], ey.expand(nzp, nyp)[I], fy[I]

    nt, S, Ns = source_amplitudes.shape
    R = x_r.shape[1]
    sc = locations_to_cells(x_s, dz, dy, P, nz, ny, loc_mode)          # [S, Ns, 2]
    rc = locations_to_cells(x_r, dz, dy, P, nz, ny, loc_mode)          # [S, R, 2]
    s_lin = (sc[..., 0] * nyp + sc[..., 1])                            # [S, Ns]
    r_lin = (rc[..., 0] * nyp + rc[..., 1])
    w_src = w.reshape(-1)[s_lin]                                       # [S, Ns]
    amp = upsample(source_amplitudes.to(dtype), ratio)                 # [nt*ratio, S, Ns]
    nti = nt * ratio

    u = torch.zeros(S, nzp, nyp, dtype=dtype)
    um = torch.zeros_like(u)
    pz = torch.zeros_like(u)
    py = torch.zeros_like(u)
    pad4 = (h, h, h, h)
    rec = []
    for n in range(nti):
        if n % ratio == 0:
            rec.append(u.reshape(S, -1).gather(1, r_lin))
        q = (_lap(u, h, c2, 1.0 / dz ** 2, 1.0 / dy ** 2)
             + _d1(pz, 0, h, c1, 1.0 / dz) + _d1(py, 1, h, c1, 1.0 / dy))
        un_i = cu_i * u[(Ellipsis,) + I] + cum_i * um[(Ellipsis,) + I] + aw_i * q
        pz_i = ez_i * pz[(Ellipsis,) + I] + fz_i * _d1(u, 0, h, c1, 1.0 / dz)
        py_i = ey_i * py[(Ellipsis,) + I] + fy_i * _d1(u, 1, h, c1, 1.0 / dy)
        un = F.pad(un_i, pad4)
        inj = torch.zeros(S, nzp * nyp, dtype=dtype).scatter_add(1, s_lin, w_src * amp[n])
        un = un + inj.reshape(S, nzp, nyp)
        um, u = u, un
        pz, py = F.pad(pz_i, pad4), F.pad(py_i, pad4)
    out = torch.stack(rec, dim=0)                                      # [nt, S, R]
    if return_wavefield:
        return out, u
    return out


def acoustic_gradient(v, dx, source_amplitudes, x_s, x_r, dt, grad_rec=None, observed=None, **kw):
    """Convenience: returns (rec, grad_v, grad_src) for J = <grad_rec, rec> or 0.5||rec-observed||^2."""
    v = v.detach().clone().requires_grad_(True)
    s = source_amplitudes.detach().clone().requires_grad_(True)
    rec = acoustic_forward(v, dx, s, x_s, x_r, dt, **kw)
    if grad_rec is not None:
        J = (rec * grad_rec.to(rec.dtype)).sum()
    else:
        J = 0.5 * ((rec - observed.to(rec.dtype)) ** 2).sum()
    J.backward()
    return rec.detach(), v.grad.detach(), s.grad.detach()
