"""ctypes front end of the C oracle (oracle/cpu_ref.c).  TEST INFRASTRUCTURE ONLY.

Mirrors the reference call  Propagator({'vp': v}, dx)(src, x_s, x_r, dt)  (models/networks.py:
5449,5464) and its backward (:5491) on the CPU.  PARITY UNPINNED (see cpu_ref.c header).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional

import numpy as np
import torch

from . import torch_ref as T

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "libfwi_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".c", ".h"))]
    stale = (not os.path.exists(so)) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "libfwi_oracle.so"],
                              stdout=subprocess.DEVNULL)
    return so


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        _LIB.fwi_oracle_num_threads.restype = ctypes.c_int
    return _LIB


def num_threads() -> int:
    return int(lib().fwi_oracle_num_threads())


def set_threads(n: int) -> int:
    """OpenMP threads of the C oracle (bench.py's CPU legs: torchrun exports OMP_NUM_THREADS=1)."""
    lib().fwi_oracle_set_threads(ctypes.c_int(int(n)))
    return num_threads()


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def run(np_dtype, nzp, nyp, order, pml_width, S, dz, dy, dt, w, sigz, sigy, nsteps, adjoint,
        inj_cell, inj_amp, inj_scale_a, smp_cell, smp_div_a, q_write=None, q_read=None, G=None,
        u_final=None, state=None, save_every=1):
    """Thin call of fwi_oracle_acoustic_run[_state]_{f32,f64}; returns smp_out [nsteps, S, n_smp].
    `state` [6, S, nzp*nyp] (u, u-, pz, -, py, -) is read and left updated when given."""
    if save_every > 1:
        assert state is None and u_final is None
    f = getattr(lib(), "fwi_oracle_acoustic_run_" + ("state_" if state is not None else "k_" if save_every > 1 else "")
                + ("f32" if np_dtype == np.float32 else "f64"))
    f.restype = ctypes.c_int
    n_inj = 0 if inj_cell is None else inj_cell.shape[1]
    n_smp = 0 if smp_cell is None else smp_cell.shape[1]
    smp_out = np.zeros((nsteps, S, n_smp), dtype=np_dtype) if n_smp else None
    for a in (w, sigz, sigy, inj_amp, q_write, q_read, G, u_final, state):
        assert a is None or (a.dtype == np_dtype and a.flags.c_contiguous)
    for a in (inj_cell, smp_cell):
        assert a is None or (a.dtype == np.int32 and a.flags.c_contiguous)
    rc = f(ctypes.c_int(nzp), ctypes.c_int(nyp), ctypes.c_int(order), ctypes.c_int(pml_width),
           ctypes.c_int(S), ctypes.c_double(dz), ctypes.c_double(dy), ctypes.c_double(dt),
           _ptr(w), _ptr(sigz), _ptr(sigy), ctypes.c_int(nsteps), ctypes.c_int(adjoint),
           ctypes.c_int(n_inj), _ptr(inj_cell), _ptr(inj_amp), ctypes.c_int(inj_scale_a),
           ctypes.c_int(n_smp), _ptr(smp_cell), _ptr(smp_out), ctypes.c_int(smp_div_a),
           _ptr(q_write), _ptr(q_read), _ptr(G),
           ctypes.c_int(save_every) if save_every > 1 else _ptr(state if state is not None else u_final))
    if rc != 0:
        raise RuntimeError(f"fwi_oracle_acoustic_run failed rc={rc}")
    return smp_out


class AcousticSetup:
    """Host-side preparation shared by forward/backward (restates SURVEY.md 7.0)."""

    def __init__(self, v: torch.Tensor, dx, dt: float, order=4, pml_width=10, vpmax=None,
                 dtype=torch.float64):
        self.tdtype = dtype
        self.npdtype = np.float32 if dtype == torch.float32 else np.float64
        self.dz, self.dy = T._pair(dx)
        self.nz, self.ny = v.shape
        self.order, self.h, self.pw = order, order // 2, pml_width
        self.P = pml_width + self.h
        vmax = float(vpmax) if vpmax is not None else float(v.abs().max())
        self.ratio, self.dti = T.cfl_substep(dt, self.dz, self.dy, vmax)
        vp = torch.nn.functional.pad(v.to(dtype)[None, None], (self.P,) * 4, mode="replicate")[0, 0]
        self.v_pad = vp
        self.nzp, self.nyp = vp.shape
        dti = torch.tensor(self.dti, dtype=dtype)
        self.w = ((vp * vp) * (dti * dti)).contiguous().numpy()
        self.sigz = T.pml_profile(self.nzp, self.h, pml_width, self.dz, vmax, dtype).numpy()
        self.sigy = T.pml_profile(self.nyp, self.h, pml_width, self.dy, vmax, dtype).numpy()

    def cells(self, x, mode="trunc"):
        c = T.locations_to_cells(x, self.dz, self.dy, self.P, self.nz, self.ny, mode)
        return (c[..., 0] * self.nyp + c[..., 1]).to(torch.int32).contiguous().numpy()


def _buffer(workspace, name, shape, dtype):
    """Zero-initialised array; with a `workspace` dict the array is kept and handed out again (same shape), so a
    timed repeat neither allocates nor first-touches gigabytes of snapshots (a CPU code would reuse them too).
    Stale contents are harmless: every cell the adjoint reads is rewritten by the forward pass before it."""
    if workspace is None:
        return np.zeros(shape, dtype=dtype)
    a = workspace.get(name)
    if a is None or a.shape != tuple(shape) or a.dtype != dtype:
        a = workspace[name] = np.zeros(shape, dtype=dtype)
    return a


def acoustic_forward(v, dx, source_amplitudes, x_s, x_r, dt, *, order=4, pml_width=10,
                     vpmax=None, loc_mode="trunc", dtype=torch.float64, save=False, workspace=None, save_every=1):
    su = AcousticSetup(v, dx, dt, order, pml_width, vpmax, dtype)
    su.save_every = max(1, int(save_every))
    nt, S, _ = source_amplitudes.shape
    amp = T.upsample(source_amplitudes.detach().to(dtype), su.ratio).contiguous().numpy()
    nti = nt * su.ratio
    sc, rc = su.cells(x_s, loc_mode), su.cells(x_r, loc_mode)
    nq = -(-nti // su.save_every)
    q = _buffer(workspace, "q", (nq, S, su.nzp * su.nyp), su.npdtype) if save else None
    rec = run(su.npdtype, su.nzp, su.nyp, order, pml_width, S, su.dz, su.dy, su.dti, su.w,
              su.sigz, su.sigy, nti, 0, sc, amp, 0, rc, 0, q_write=q, save_every=su.save_every)
    rec = torch.from_numpy(rec[::su.ratio].copy())
    if save:
        return rec, (su, q, sc, rc, amp)
    return rec


def acoustic_adjoint(state, grad_rec, source_amplitudes, dtype=torch.float64):
    """Hand-written adjoint of acoustic_forward(save=True): `state` is what that call returned next to the traces.
    Returns (grad_v [nz, ny], grad_src [nt, S, Ns]) for J = <grad_rec, rec>."""
    su, q, sc, rc, amp = state
    nt, S, Ns = source_amplitudes.shape
    nti = nt * su.ratio
    g = grad_rec.to(dtype)
    g_in = np.zeros((nti, S, g.shape[2]), dtype=su.npdtype)
    g_in[::su.ratio] = g.numpy()
    G = np.zeros((S, su.nzp * su.nyp), dtype=su.npdtype)
    gs = run(su.npdtype, su.nzp, su.nyp, su.order, su.pw, S, su.dz, su.dy, su.dti, su.w,
             su.sigz, su.sigy, nti, 1, rc, g_in, 1, sc, 1, q_read=q, G=G, save_every=getattr(su, "save_every", 1))
    # dJ/dv_pad = (2 / v_pad) * (G + scatter_src(sum_n gs_n * s_n)), then fold the replicate pad
    # (save_every = k: the image holds every k-th product; rectangle rule, weight k)
    Gt = torch.from_numpy(G).sum(0) * getattr(su, "save_every", 1)
    gs_t = torch.from_numpy(gs)                                     # [nti, S, Ns]
    src_term = (gs_t * torch.from_numpy(amp)).sum(0).reshape(-1)    # [S*Ns]
    Gt = Gt.index_add(0, torch.from_numpy(sc.reshape(-1).astype(np.int64)), src_term)
    gpad = (2.0 / su.v_pad) * Gt.reshape(su.nzp, su.nyp)
    grad_v = fold_replicate_pad(gpad, su.P)
    # adjoint of the FFT upsampling: apply autograd to the linear map
    with torch.enable_grad():            # (callable from inside an autograd Function's backward)
        s = source_amplitudes.detach().to(dtype).clone().requires_grad_(True)
        (gs_s,) = torch.autograd.grad((T.upsample(s, su.ratio) * gs_t).sum(), s)
    return grad_v, gs_s


def acoustic_gradient(v, dx, source_amplitudes, x_s, x_r, dt, grad_rec=None, observed=None, *,
                      order=4, pml_width=10, vpmax=None, loc_mode="trunc", dtype=torch.float64, workspace=None,
                      save_every=1):
    """Returns (rec, grad_v [nz, ny], grad_src [nt, S, Ns]) with the hand-written adjoint.  save_every = k > 1: the
    decimated imaging condition of SURVEY.md 7.0 (snapshots and products every k-th inner step, weight k)."""
    rec, state = acoustic_forward(
        v, dx, source_amplitudes, x_s, x_r, dt, order=order, pml_width=pml_width, vpmax=vpmax,
        loc_mode=loc_mode, dtype=dtype, save=True, workspace=workspace, save_every=save_every)
    g = grad_rec.to(dtype) if grad_rec is not None else (rec - observed.to(dtype))
    grad_v, grad_s = acoustic_adjoint(state, g, source_amplitudes, dtype)
    return rec, grad_v, grad_s


def fold_replicate_pad(gpad: torch.Tensor, P: int) -> torch.Tensor:
    """Adjoint of replicate padding by P on all four sides of a 2-D array."""
    if P == 0:
        return gpad
    g = gpad.clone()
    g[P, :] += g[:P, :].sum(0)
    g[-P - 1, :] += g[-P:, :].sum(0)
    g = g[P:-P, :]
    g[:, P] += g[:, :P].sum(1)
    g[:, -P - 1] += g[:, -P:].sum(1)
    return g[:, P:-P].contiguous()
