"""CPU stand-in for scalar._Plan used ONLY by the gloo tests of the x-slab driver (slab.py).

The time loop is the C oracle's state-carrying run (oracle/cpu_ref.c, fp32 build); layout is the
oracle's pitch-free [nzp][nyp].  It mirrors the attributes and methods slab._SlabFunction uses, so
the decomposition logic (windows, ghost exchange, masks, gradient assembly) is exercised on CPU
with world_size 2 while the product always runs the CUDA plan."""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

from oracle import cpu_ref as O
from physicsbasedfwi2_b200 import _host as H


class CpuPlan:
    def __init__(self, v, dx, dt, order, pml_width, vpmax, n_shots, chunk_rows=0, window=None):
        self.dev = torch.device("cpu")
        self.dz, self.dy = H.spacing(dx)
        self.nz, self.ny = int(v.shape[0]), int(v.shape[1])
        self.order, self.h, self.pw = order, order // 2, int(pml_width)
        self.P = self.pw + self.h
        self.nzp, self.nyp_global = self.nz + 2 * self.P, self.ny + 2 * self.P
        self.c0, self.c1 = (0, self.nyp_global) if window is None else window
        self.nyp = self.c1 - self.c0
        self.vmax = float(vpmax) if vpmax is not None else float(v.detach().abs().max())
        self.ratio, self.dti = H.inner_step(float(dt), self.dz, self.dy, self.vmax)
        self.S = n_shots
        self.pitch, self.col0, self.stride = self.nyp, 0, self.nzp * self.nyp
        self.n_fields, self.n_images = 6, n_shots
        vp = F.pad(v.detach().float()[None, None], (self.P,) * 4, mode="replicate")[0, 0]
        self.v_pad_global = vp
        self.v_pad = vp[:, self.c0:self.c1]
        dti = torch.tensor(self.dti, dtype=torch.float32)
        self.w = ((self.v_pad * self.v_pad) * (dti * dti)).contiguous().numpy()
        self.sigz = H.sigma_profile(self.nzp, self.h, self.pw, self.dz, self.vmax).float().numpy()
        self.sigy = H.sigma_profile(self.nyp_global, self.h, self.pw, self.dy, self.vmax).float()[
            self.c0:self.c1].contiguous().numpy()

    def points(self, x, what, loc_mode):
        c = H.to_cells(x, self.dz, self.dy, loc_mode)
        bad = ((c[..., 0] < 0) | (c[..., 0] >= self.nz) | (c[..., 1] < 0) | (c[..., 1] >= self.ny)).any()
        z = c[..., 0].clamp(0, self.nz - 1) + self.P
        ycol = (c[..., 1].clamp(0, self.ny - 1) + self.P - self.c0).clamp(self.h, self.nyp - self.h - 1)
        return z * self.pitch + ycol, z, ycol, bad

    def padded_columns(self, x, loc_mode):
        return H.to_cells(x, self.dz, self.dy, loc_mode)[..., 1] + self.P

    def lists(self, cell, keep=None):      # dropped points carry zero amplitude; the oracle takes a dense list
        return (None, cell.to(torch.int32).contiguous(), None)

    def new_wave(self):
        return torch.zeros(6 * self.S * self.stride, dtype=torch.float32)

    def settle(self, wave, nsteps):
        return wave          # the oracle leaves its state in canonical order

    def run(self, *, nsteps, adjoint, inj, inj_amp, smp_cell, n_smp, q_write=None, q_read=None, image=None,
            wave=None):
        if wave is None:
            wave = self.new_wave()
        _, icell, _ = inj
        np_ = lambda t: None if t is None else t.numpy()
        amp = None if inj_amp is None else np.ascontiguousarray(inj_amp.numpy())
        smp = None if n_smp == 0 else np.ascontiguousarray(smp_cell.numpy().reshape(self.S, n_smp))
        out = O.run(np.float32, self.nzp, self.nyp, self.order, self.pw, self.S, self.dz, self.dy, self.dti,
                    self.w, self.sigz, self.sigy, nsteps, adjoint,
                    None if amp is None else np.ascontiguousarray(icell.numpy().reshape(self.S, -1)), amp,
                    1 if adjoint else 0, smp, 1 if adjoint else 0, q_write=np_(q_write), q_read=np_(q_read),
                    G=np_(image), state=wave.numpy())
        if out is None:
            return torch.zeros(nsteps, self.S, 0)
        return torch.from_numpy(out)
