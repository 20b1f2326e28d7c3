"""deepwave-style scalar (constant-density acoustic) propagator on the B200 kernels.

Drop-in for the call the reference makes every training iteration
(models/networks.py:5408-5411, :5449, :5464, backward at :5491):

    prop = Propagator({'vp': model}, dx)
    receiver_amplitudes = prop(source_amplitudes, x_s, x_r, dt)     # [nt, shots, receivers]
    loss(receiver_amplitudes, ...).backward()                        # -> model.grad, src.grad

The time loops run in libfwi_b200.so (sm_100a kernels, include/fwi_b200.h).  There is no CPU
path: tensors must live on a CUDA device and the library must be built.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np
import torch
import torch.nn.functional as F

from . import _host as H
from . import _lib


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


class _Plan:
    """Everything about one call that does not depend on the direction of time."""

    def __init__(self, v: torch.Tensor, dx, dt: float, order: int, pml_width: int,
                 vpmax: Optional[float], n_shots: int, chunk_rows: int = 0, window=None):
        """`window` = (c0, c1): restrict the plan to the padded columns [c0, c1) of the global grid
        (an x-slab, slab.py); model, PML profiles and point coordinates stay global."""
        if v.dim() != 2:
            raise RuntimeError(f"model['vp'] must be 2-D [nz, ny], got shape {tuple(v.shape)}")
        if not v.is_cuda:
            raise RuntimeError("model['vp'] must be a CUDA tensor: this propagator has no CPU path")
        if v.dtype != torch.float32:
            raise RuntimeError(f"model['vp'] must be float32, got {v.dtype}")
        if order not in (2, 4, 6, 8):
            raise RuntimeError(f"accuracy must be 2, 4, 6 or 8, got {order}")
        self.dev = v.device
        self.dz, self.dy = H.spacing(dx)
        self.nz, self.ny = int(v.shape[0]), int(v.shape[1])
        self.order, self.h, self.pw = order, order // 2, int(pml_width)
        self.P = self.pw + self.h
        self.nzp, self.nyp_global = self.nz + 2 * self.P, self.ny + 2 * self.P
        self.c0, self.c1 = (0, self.nyp_global) if window is None else (int(window[0]), int(window[1]))
        if not (0 <= self.c0 < self.c1 <= self.nyp_global) or self.c1 - self.c0 <= 2 * self.P:
            raise RuntimeError(f"bad column window {window} for a padded width of {self.nyp_global}")
        self.nyp = self.c1 - self.c0
        self.vmax = float(vpmax) if vpmax is not None else float(v.detach().abs().max())
        self.ratio, self.dti = H.inner_step(float(dt), self.dz, self.dy, self.vmax)
        self.S = n_shots
        self.chunk_rows = chunk_rows

        lay = _lib.AcousticLayout()
        d = self.desc(nsteps=0, adjoint=0, n_inj=0, n_smp=0)
        _lib.check(_lib.lib().fwi_acoustic_get_layout(C.byref(d), C.byref(lay)), "fwi_acoustic_get_layout")
        self.lay = lay
        self.pitch, self.col0 = int(lay.pitch), int(lay.col0)
        self.stride = int(lay.shot_stride)

        # padded copy of the model: the x-slab driver folds its gradient through the padding itself (slab.py)
        self.v_pad_global = F.pad(v.detach()[None, None], (self.P,) * 4, mode="replicate")[0, 0]
        self.v_pad = self.v_pad_global[:, self.c0:self.c1]
        # w = v^2 dt^2 on the padded window and the PML profiles: one native kernel (csrc/host_kernels.cu), nothing cached
        self.w = torch.empty(self.nzp, self.pitch, dtype=torch.float32, device=self.dev)
        self.sigz = torch.empty(self.nzp, dtype=torch.float32, device=self.dev)
        self.sigy = torch.empty(self.pitch, dtype=torch.float32, device=self.dev)
        m = _lib.ModelDesc(self.nz, self.ny, self.order, self.pw, self.c0, self.c1, self.dz, self.dy, self.dti, self.vmax)
        vc = v.detach().contiguous()
        with torch.cuda.device(self.dev):
            _lib.check(_lib.lib().fwi_acoustic_prepare_model(C.byref(m), _ptr(vc), _ptr(self.w), _ptr(self.sigz), _ptr(self.sigy),
                                                             C.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)),
                       "fwi_acoustic_prepare_model")
        self.n_fields, self.n_images = int(lay.wave_fields), int(lay.n_images)

    def desc(self, nsteps, adjoint, n_inj, n_smp) -> _lib.AcousticDesc:
        return _lib.AcousticDesc(self.nzp, self.nyp, self.order, self.pw, self.S, self.chunk_rows, nsteps,
                                 adjoint, n_inj, 1 if adjoint else 0, n_smp, 1 if adjoint else 0,
                                 self.dz, self.dy, self.dti)

    def points(self, x: torch.Tensor, what: str, loc_mode: str):
        """[S, n, 2] metres -> (cell [S, n] int32, z [S, n], ycol [S, n]) in device layout + bad flag."""
        c = H.to_cells(x.to(self.dev), self.dz, self.dy, loc_mode)
        bad = ((c[..., 0] < 0) | (c[..., 0] >= self.nz) | (c[..., 1] < 0) | (c[..., 1] >= self.ny)).any()
        z = c[..., 0].clamp(0, self.nz - 1) + self.P
        ycol = (c[..., 1].clamp(0, self.ny - 1) + self.P - self.c0).clamp(self.h, self.nyp - self.h - 1) + self.col0
        cell = z * self.pitch + ycol
        return cell, z, ycol, bad

    def padded_columns(self, x: torch.Tensor, loc_mode: str) -> torch.Tensor:
        """[S, n, 2] metres -> global padded column [S, n] of each point."""
        return H.to_cells(x.to(self.dev), self.dz, self.dy, loc_mode)[..., 1] + self.P

    def lists(self, cell, keep=None):
        return H.injection_lists(cell, self.nzp, self.pitch, keep)

    def new_wave(self) -> torch.Tensor:
        return torch.zeros(int(self.lay.wave_fields) * self.S * self.stride, dtype=torch.float32, device=self.dev)

    def settle(self, wave: torch.Tensor, nsteps: int) -> torch.Tensor:
        """The library starts every call with (u, u-) = fields (0, 1) and rotates the pairs each
        step: after an odd number of steps put the fields back in that order so that the state
        can be carried into another call."""
        if nsteps % 2 == 0:
            return wave
        nf = int(self.lay.wave_fields)
        perm = [1, 0, 3, 2, 5, 4][:nf]
        return wave.view(nf, -1)[perm].contiguous().view(-1)

    def run(self, *, nsteps, adjoint, inj, inj_amp, smp_cell, n_smp, q_write=None, q_read=None,
            image=None, wave=None, link=None) -> torch.Tensor:
        """One time loop.  inj = (start, cell, col) lists; returns smp_out [nsteps, S, n_smp].
        `wave` carries the state of a previous call (see settle()); None starts from rest.
        `link` (a _lib.SlabLink): the loop exchanges ghost columns with the neighbouring slabs."""
        n_inj = 0 if inj_amp is None else int(inj_amp.shape[2])
        d = self.desc(nsteps, adjoint, n_inj, n_smp)
        if wave is None:
            wave = self.new_wave()
        smp_out = torch.empty(nsteps, self.S, n_smp, dtype=torch.float32, device=self.dev)
        start, icell, icol = inj
        table = torch.empty(int(self.lay.unit_table_bytes), dtype=torch.uint8, device=self.dev)
        b = _lib.AcousticBuffers(_ptr(self.w), _ptr(self.sigz), _ptr(self.sigy), _ptr(wave), _ptr(start),
                                 _ptr(icell), _ptr(icol), _ptr(inj_amp), _ptr(smp_cell), _ptr(smp_out),
                                 _ptr(q_write), _ptr(q_read), _ptr(image), _ptr(table))
        cur = torch.cuda.current_stream(self.dev)
        with torch.cuda.device(self.dev):
            fn = _lib.lib().fwi_acoustic_backward if adjoint else _lib.lib().fwi_acoustic_forward
            ev = _lib.time_loop_begin(cur)
            if link is not None:
                _lib.check(_lib.lib().fwi_acoustic_run_slab(C.byref(d), C.byref(b), C.byref(link),
                                                            C.c_void_p(cur.cuda_stream)), "fwi_acoustic_run_slab")
            else:
                _lib.check(fn(C.byref(d), C.byref(b), C.c_void_p(cur.cuda_stream)),
                           "fwi_acoustic_backward" if adjoint else "fwi_acoustic_forward")
            # tags: forward with snapshots (MODE 1), forward only (MODE 0: direct wave, checkpoint pass), adjoint
            tag = "acoustic_adj" if adjoint else ("acoustic_fwd" if q_write is not None else "acoustic_fwd0")
            _lib.time_loop_end(cur, ev, tag, nsteps, self.S)
        return smp_out


def _settle(wave: torch.Tensor, nsteps: int, n_fields: int) -> torch.Tensor:
    """The library starts every call with (u, u-) = fields (0, 1) and rotates the pairs each step: after an odd number
    of steps put the fields back in that order so that the state can be carried into another call."""
    if nsteps % 2 == 0:
        return wave
    perm = [1, 0, 3, 2, 5, 4][:n_fields]
    return wave.view(n_fields, -1)[perm].contiguous().view(-1)


def _forward_groups(n_shots: int, shot_stride: int, env=None):
    """Shot groups [s0, s1) of the forward time loop: one group unless at least two more shots than fit L2 with their
    two wavefields (64 MB budget of the 126 MB L2; FWI_B200_FWD_SHOTS overrides the group size, 0 = never split)."""
    per_shot = 2 * shot_stride * 4
    nmax = max(1, int(64e6 // per_shot))
    env = env if env is not None else os.environ.get("FWI_B200_FWD_SHOTS")
    if env is not None:
        nmax = int(env)
        if nmax <= 0 or n_shots <= nmax:
            return [(0, n_shots)]
    elif n_shots < nmax + 2 or per_shot < (1 << 20):
        return [(0, n_shots)]
    k = -(-n_shots // nmax)
    base, extra = divmod(n_shots, k)
    out, lo = [], 0
    for g in range(k):
        hi = lo + base + (1 if g < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def _adjoint_groups(n_shots: int, shot_stride: int):
    """Shot groups of the adjoint time loop, run concurrently on their own streams: two groups when there are at least four
    shots on a grid large enough for the step kernel to be bandwidth-bound (FWI_B200_ADJ_GROUPS overrides the number of
    groups; 1 = one launch over all shots)."""
    k = 2 if (n_shots >= 4 and shot_stride * 4 >= (4 << 20)) else 1
    env = os.environ.get("FWI_B200_ADJ_GROUPS")
    if env is not None:
        k = max(1, min(int(env), n_shots))
    base, extra = divmod(n_shots, k)
    out, lo = [], 0
    for g in range(k):
        hi = lo + base + (1 if g < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


class _ScalarFunction(torch.autograd.Function):
    """forward / backward of the scalar propagator on the TORCH_LIBRARY(b200fwi) operators (csrc/torch_ext.cpp):
    plan -> prepare (model pad + scale + PML profiles, one kernel) -> plan_points (cells + sorted lists, one kernel per
    point set) -> run (the native time loop) [-> fold (image -> dJ/dv) in backward]."""

    @staticmethod
    def forward(ctx, v, source_amplitudes, x_s, x_r, dt, dx, order, pml_width, vpmax, loc_mode,
                chunk_rows, checkpoint_every=0, save_every=1):
        if source_amplitudes.dim() != 3:
            raise RuntimeError("source_amplitudes must be [nt, n_shots, n_sources_per_shot]")
        nt, S, Ns = source_amplitudes.shape
        if x_s.dim() != 3 or x_s.shape[0] != S or x_s.shape[1] != Ns or x_s.shape[2] != 2:
            raise RuntimeError(f"source_locations must be [{S}, {Ns}, 2], got {tuple(x_s.shape)}")
        if x_r.dim() != 3 or x_r.shape[0] != S or x_r.shape[2] != 2:
            raise RuntimeError(f"receiver_locations must be [{S}, n_receivers, 2], got {tuple(x_r.shape)}")
        if loc_mode not in ("trunc", "round"):
            raise RuntimeError(f"unknown loc_mode {loc_mode!r}")
        if not v.is_cuda:
            raise RuntimeError("model['vp'] must be a CUDA tensor: this propagator has no CPU path")
        O = _lib.ops()
        R = int(x_r.shape[1])
        dz, dy = H.spacing(dx)
        dev = v.device
        geo = (int(order), int(pml_width), dz, dy)
        meta = O.scalar_plan(v, float(dt), dz, dy, int(order), int(pml_width), float(vpmax) if vpmax is not None else -1.0,
                             S, int(chunk_rows), 0, 0)
        ratio, n_fields, stride = int(meta[1]), int(meta[7]), int(meta[8])
        vd = v.detach()
        w, sigz, sigy = O.scalar_prepare(vd, meta, *geo, 0, 0)
        bad = torch.zeros(1, dtype=torch.int32, device=dev)
        rnd = loc_mode == "round"
        s_cell, s_start, s_sc, s_so = O.scalar_plan_points(x_s, vd, meta, bad, *geo, rnd, 0, 0)
        r_cell, r_start, r_sc, r_so = O.scalar_plan_points(x_r, vd, meta, bad, *geo, rnd, 0, 0)
        # The out-of-model flag travels to pinned memory ahead of the time loop and is read after the loop is queued:
        # the host only waits for work queued before this call, never for this call's own kernels
        flag = O.flag_begin(bad)
        amp = H.resample_up(source_amplitudes.detach().to(dev, torch.float32), ratio).contiguous()
        nti = nt * ratio
        need_v, need_s = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        cur = torch.cuda.current_stream(dev)

        # Forward time loops run per GROUP of shots whose two wavefields fit L2 (one group on small problems): a forward
        # launch is fastest when its wavefields stay in L2 from one step to the next (B200: 126 MB; measured on C2: 3 shots
        # 18.2 us per step, 6 shots 41.2).  Groups run one after the other - concurrently on two streams they evict each
        # other (38.2 us; FWI_B200_MULTI=1 to try) - and write into the snapshot array of all shots (q_step_stride).
        fgroups = _forward_groups(S, stride)
        one = len(fgroups) == 1
        fplans = [(s_start, s_sc, s_so)] if one else [O.scalar_plan_points(x_s[s0:s1], vd, meta, bad, *geo, rnd, 0, 0)[1:]
                                                      for s0, s1 in fgroups]
        famps = [amp] if one else [amp[:, s0:s1].contiguous() for s0, s1 in fgroups]

        def run(gi, t0, nsteps, wave, q_write):
            """Steps [t0, t0 + nsteps) of the forward time loop of group gi; q_write: snapshot array of ALL shots for these steps."""
            s0, s1 = fgroups[gi]
            ev = _lib.time_loop_begin(cur)
            out = O.scalar_run(w, sigz, sigy, wave, *fplans[gi], famps[gi][t0:t0 + nsteps], r_cell[s0:s1],
                               None if q_write is None else q_write[s0 * stride:], None, None, meta, s1 - s0, nsteps, False,
                               *geo, int(chunk_rows), 0 if (one or q_write is None) else S * stride, kq)
            _lib.time_loop_end(cur, ev, "acoustic_fwd" if q_write is not None else "acoustic_fwd0", nsteps, s1 - s0)
            return out

        def new_wave(gi):
            return torch.zeros(n_fields * (fgroups[gi][1] - fgroups[gi][0]) * stride, dtype=torch.float32, device=dev)

        q, K = None, int(checkpoint_every)
        kq = max(1, int(save_every))            # snapshot decimation: every kq-th inner step is stored and imaged
        ctx.ckpt = None
        with torch.cuda.device(dev):
            if need_v and 0 < K < nti:
                if kq > 1:
                    raise RuntimeError("save_every and checkpoint_every cannot be combined")
                # wavefield checkpointing: keep the state every K steps instead of a snapshot per step;
                # backward() recomputes the snapshots of one segment at a time
                ckpts, recs = [], []
                for gi in range(len(fgroups)):
                    wave, parts, cg = new_wave(gi), [], []
                    for t0 in range(0, nti, K):
                        n = min(K, nti - t0)
                        cg.append(wave.clone())
                        parts.append(run(gi, t0, n, wave, None))
                        wave = _settle(wave, n, n_fields)
                    ckpts.append(cg)
                    recs.append(torch.cat(parts, 0))
                rec = recs[0] if one else torch.cat(recs, 1)
                ctx.ckpt = (K, ckpts)
            else:
                if need_v:
                    q = torch.empty(-(-nti // kq) * S * stride, dtype=torch.float32, device=dev)
                if one or os.environ.get("FWI_B200_MULTI", "0") != "1":
                    recs = [run(gi, 0, nti, new_wave(gi), q) for gi in range(len(fgroups))]
                else:
                    ev = _lib.time_loop_begin(cur)
                    recs = list(O.scalar_run_multi(w, sigz, sigy, [new_wave(gi) for gi in range(len(fgroups))],
                                                   [p[0] for p in fplans], [p[1] for p in fplans], [p[2] for p in fplans], famps,
                                                   [r_cell[s0:s1] for s0, s1 in fgroups], q, None, None,
                                                   [s0 for s0, _ in fgroups], meta, S, nti, False, *geo, int(chunk_rows), kq))
                    _lib.time_loop_end(cur, ev, "acoustic_fwd" if q is not None else "acoustic_fwd0", nti, S)
                rec = recs[0] if one else torch.cat(recs, 1)
        O.flag_end(flag)
        out = rec[::ratio].contiguous() if ratio > 1 else rec
        ctx.saved = (vd, meta, w, sigz, sigy, s_cell, r_cell, r_start, r_sc, r_so, s_start, s_sc, s_so)
        ctx.q = q
        ctx.amp = amp if need_v else None
        ctx.dims = (nt, S, Ns, R, nti, ratio, n_fields, stride, geo, int(chunk_rows), kq)
        ctx.need = (need_v, need_s)
        ctx.src_like = source_amplitudes.detach()
        ctx.x_r, ctx.rnd = x_r, rnd
        ctx.fwd = (fgroups, fplans, famps)
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, grad_rec):
        O = _lib.ops()
        vd, meta, w, sigz, sigy, s_cell, r_cell, r_start, r_sc, r_so, s_start, s_sc, s_so = ctx.saved
        nt, S, Ns, R, nti, ratio, n_fields, stride, geo, chunk_rows, kq = ctx.dims
        need_v, need_s = ctx.need
        dev = vd.device
        g = grad_rec.to(dev, torch.float32)
        if ratio > 1:
            g_in = torch.zeros(nti, S, R, dtype=torch.float32, device=dev)
            g_in[::ratio] = g
        else:
            g_in = g.contiguous()
        image = torch.zeros(S * stride, dtype=torch.float32, device=dev) if need_v else None
        cur = torch.cuda.current_stream(dev)

        fgroups, fplans, famps = ctx.fwd
        # The adjoint does not fit L2 (wavefields + image + two snapshots per shot), so small groups gain it nothing as such -
        # but every launch has a ramp (CTAs start over ~3 us, every warp waits for its first TMA tile) and a tail (the
        # slowest warp), ~13 % of a 6-shot launch.  Two groups on two streams interleave their launches, each covering the
        # other's ramp and tail (fwi_acoustic_run_multi): measured on C2 46.9 us per step against 55.4 in one launch.
        agroups = _adjoint_groups(S, stride) if need_v else [(0, S)]
        a_one = len(agroups) == 1
        if a_one:
            aplans = [(r_start, r_sc, r_so)]
        else:
            nobad = torch.zeros(1, dtype=torch.int32, device=dev)
            aplans = [O.scalar_plan_points(ctx.x_r[s0:s1], vd, meta, nobad, *geo, ctx.rnd, 0, 0)[1:] for s0, s1 in agroups]
        awaves = [torch.zeros(n_fields * (s1 - s0) * stride, dtype=torch.float32, device=dev) for s0, s1 in agroups]

        def run_adjoint(t0, nsteps, q_read):
            """Adjoint steps [t0, t0 + nsteps) (in reverse) of every group; q_read: snapshots of ALL shots for these steps."""
            amps = [g_in[t0:t0 + nsteps] if a_one else g_in[t0:t0 + nsteps, s0:s1].contiguous() for s0, s1 in agroups]
            ev = _lib.time_loop_begin(cur)
            if a_one:
                out = O.scalar_run(w, sigz, sigy, awaves[0], *aplans[0], amps[0].contiguous(), s_cell, None, q_read,
                                   image if q_read is not None else None, meta, S, nsteps, True, *geo, chunk_rows, 0, kq)
            else:
                parts = O.scalar_run_multi(w, sigz, sigy, awaves, [p[0] for p in aplans], [p[1] for p in aplans],
                                           [p[2] for p in aplans], amps, [s_cell[s0:s1] for s0, s1 in agroups], None, q_read,
                                           image, [s0 for s0, _ in agroups], meta, S, nsteps, True, *geo, chunk_rows, kq)
                out = torch.cat(list(parts), 1)
            _lib.time_loop_end(cur, ev, "acoustic_adj", nsteps, S)
            return out

        with torch.cuda.device(dev):
            if ctx.ckpt is not None:
                K, ckpts = ctx.ckpt
                qseg = torch.empty(K * S * stride, dtype=torch.float32, device=dev)
                f_one = len(fgroups) == 1
                parts = []
                for t0 in reversed(range(0, nti, K)):
                    n = min(K, nti - t0)
                    for gi, (s0, s1) in enumerate(fgroups):           # recompute this segment's snapshots, group by group
                        ev = _lib.time_loop_begin(cur)
                        O.scalar_run(w, sigz, sigy, ckpts[gi].pop(), *fplans[gi], famps[gi][t0:t0 + n], r_cell[s0:s1, :0],
                                     qseg[s0 * stride:], None, None, meta, s1 - s0, n, False, *geo, chunk_rows,
                                     0 if f_one else S * stride)
                        _lib.time_loop_end(cur, ev, "acoustic_fwd", n, s1 - s0)
                    parts.append(run_adjoint(t0, n, qseg))
                    awaves = [_settle(a, n, n_fields) for a in awaves]
                gs = torch.cat(parts[::-1], 0)
                ctx.ckpt = None
            else:
                gs = run_adjoint(0, nti, ctx.q if need_v else None)
            grad_v = grad_s = None
            if need_v:
                if kq > 1:
                    image.mul_(float(kq))       # rectangle rule over the decimated products
                # dJ/dv = fold((2 / v_pad) (sum_s image_s + source term)): one native call (csrc/host_kernels.cu)
                grad_v = O.scalar_fold(image, gs, ctx.amp, s_cell, vd, meta, S, *geo)
        if need_s:
            grad_s = H.resample_up_adjoint(gs, ratio, ctx.src_like.to(dev, torch.float32))
            grad_s = grad_s.to(ctx.src_like.device, ctx.src_like.dtype)
        ctx.q = None
        return grad_v, grad_s, None, None, None, None, None, None, None, None, None, None, None


class Propagator(torch.nn.Module):
    """``deepwave.scalar.Propagator`` (v0.0.x class API) on the B200 kernels.

    Args follow the reference's usage (models/networks.py:5408, :5449): ``model`` is a dict with
    key ``'vp'`` holding a float32 CUDA tensor [nz, ny]; ``dx`` a float or (dz, dy).
    ``pml_width`` (int, default 10), ``survey_pad`` (None or cells) and ``vpmax`` as in deepwave.
    Extensions: ``accuracy`` (2/4/6/8, default 4), ``loc_mode`` ('trunc' | 'round'),
    ``chunk_rows`` (cap on the rows of one work segment, kernel tuning / tests, 0 = none),
    ``checkpoint_every`` (K > 0: wavefield checkpointing - the forward pass keeps the state every K
    inner steps instead of one snapshot per step, and the backward pass recomputes the snapshots
    of one K-step segment at a time: memory O(nt/K + K) fields instead of O(nt), one extra forward),
    ``save_every`` (k > 1: snapshot decimation - only every k-th inner step is stored and imaged, with weight k; a
    performance mode, k = 1 is the exact discrete gradient),
    ``slabs`` (True or a process group: x-slab domain decomposition of every shot over the ranks of
    the group with ghost-zone exchange every ``halo_steps`` inner steps - 'auto' picks the interval
    from the slab width, slab.auto_halo_steps - see slab.py; every rank
    passes identical arguments and receives the full traces and the full gradient; ``exchange`` =
    'peer': ghost columns move by peer stores + flags inside the native time loop, 'p2p': by
    batched NCCL send/recv between halo_steps-long launches).
    """

    def __init__(self, model: dict, dx, pml_width=None, survey_pad=None, vpmax=None, *,
                 accuracy: int = 4, loc_mode: str = "trunc", chunk_rows: int = 0, checkpoint_every: int = 0,
                 save_every: int = 1, slabs=False, halo_steps=8, exchange: str = "peer"):
        super().__init__()
        if not isinstance(model, dict) or "vp" not in model:
            raise RuntimeError("model must be a dict with key 'vp'")
        extra = set(model) - {"vp"}
        if extra:
            raise RuntimeError(f"scalar propagator takes only 'vp'; got extra keys {sorted(extra)} "
                               "(use physicsbasedfwi2_b200.Propagator for vp/vs/rho)")
        if not isinstance(model["vp"], torch.Tensor):
            raise RuntimeError("model['vp'] must be a torch.Tensor")
        self.model = model
        self.dx = dx
        if pml_width is None:
            pml_width = 10
        if not isinstance(pml_width, int):
            vals = [int(p) for p in (pml_width.tolist() if isinstance(pml_width, torch.Tensor) else pml_width)]
            if len(set(vals)) != 1:
                raise RuntimeError("pml_width must be the same on every side")
            pml_width = vals[0]
        if pml_width < 0:
            raise RuntimeError("pml_width must be >= 0")
        self.pml_width = pml_width
        self.survey_pad = survey_pad
        self.vpmax = vpmax
        self.accuracy = accuracy
        self.loc_mode = loc_mode
        self.chunk_rows = chunk_rows
        if checkpoint_every < 0:
            raise RuntimeError("checkpoint_every must be >= 0")
        self.checkpoint_every = int(checkpoint_every)
        if save_every < 1:
            raise RuntimeError("save_every must be >= 1")
        self.save_every = int(save_every)
        if self.save_every > 1 and (checkpoint_every or (slabs is not False and slabs is not None)):
            raise RuntimeError("save_every cannot be combined with checkpoint_every or slabs")
        if slabs is not False and slabs is not None and checkpoint_every:
            raise RuntimeError("slabs and checkpoint_every cannot be combined")
        self.slabs = None if slabs is False else slabs
        self.halo_steps = 0 if halo_steps in (None, 'auto') else int(halo_steps)     # 0: slab.auto_halo_steps
        if self.halo_steps < 0:
            raise RuntimeError("halo_steps must be 'auto' or >= 1")
        if exchange not in ("peer", "p2p"):
            raise RuntimeError("exchange must be 'peer' (in-loop peer stores over NVLink) or 'p2p' (NCCL send/recv)")
        self.exchange = exchange

    def forward(self, source_amplitudes: torch.Tensor, source_locations: torch.Tensor,
                receiver_locations: torch.Tensor, dt: float) -> torch.Tensor:
        v = self.model["vp"]
        x_s, x_r = source_locations, receiver_locations
        if self.survey_pad is not None:
            v, x_s, x_r = _crop_to_survey(v, self.dx, x_s, x_r, self.survey_pad)
        if self.slabs is not None:
            from .slab import _SlabFunction
            return _SlabFunction.apply(v, source_amplitudes, x_s, x_r, float(dt), self.dx, self.accuracy,
                                       self.pml_width, self.vpmax, self.loc_mode, self.halo_steps,
                                       None if self.slabs is True else self.slabs, _Plan, self.exchange)
        return _ScalarFunction.apply(v, source_amplitudes, x_s, x_r, float(dt), self.dx, self.accuracy,
                                     self.pml_width, self.vpmax, self.loc_mode, self.chunk_rows, self.checkpoint_every,
                                     self.save_every)


def _crop_to_survey(v, dx, x_s, x_r, survey_pad):
    """Restricts the model to the bounding box of the survey plus `survey_pad` cells."""
    dz, dy = H.spacing(dx)
    pad = survey_pad if isinstance(survey_pad, (list, tuple)) else [survey_pad] * 4
    pad = [0 if p is None else int(p) for p in pad]
    if len(pad) == 2:
        pad = [pad[0], pad[0], pad[1], pad[1]]
    allx = torch.cat([x_s.reshape(-1, 2), x_r.reshape(-1, 2)], 0).double().cpu()
    zc, yc = allx[:, 0] / dz, allx[:, 1] / dy
    z0 = max(0, int(torch.floor(zc.min() + 1e-6)) - pad[0])
    z1 = min(v.shape[0], int(torch.floor(zc.max() + 1e-6)) + pad[1] + 1)
    y0 = max(0, int(torch.floor(yc.min() + 1e-6)) - pad[2])
    y1 = min(v.shape[1], int(torch.floor(yc.max() + 1e-6)) + pad[3] + 1)
    off = torch.tensor([z0 * dz, y0 * dy], dtype=x_s.dtype, device=x_s.device)
    return v[z0:z1, y0:y1], x_s - off, x_r - off.to(x_r.device, x_r.dtype)
