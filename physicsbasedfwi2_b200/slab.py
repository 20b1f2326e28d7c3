"""x-slab domain decomposition of the scalar propagator: one shot spread over the GPUs of a box.

Second multi-GPU mode of the path (BASELINE config 4 variant) for grids whose
single-shot working set (one snapshot per step) does not fit one GPU.  The reference has no such
mode: it runs every shot whole on one hard-coded GPU (models/networks.py:5316, :5464), so this
is an extension behind the same call -

    prop = Propagator({'vp': v}, dx, slabs=True)          # every rank, identical arguments
    rec = prop(source_amplitudes, x_s, x_r, dt)           # full [nt, S, R] on every rank
    loss(rec).backward()                                  # full v.grad on every rank

Decomposition.  The padded grid [nzp, nyp] is cut along y (the fast axis is kept whole per row
pitch; slabs are column ranges) into `world` slabs of near-equal width.  Rank r owns the padded
columns [lo_r, hi_r) and works on the window [lo_r - G, hi_r + G) (clipped at the two ends of
the grid), G = halo_steps * order/2 ghost columns.  A window is advanced `halo_steps` steps at a
time with the ordinary step kernel (csrc/acoustic.cu, through the C ABI): the outermost order/2
columns of a window are never updated, so after k steps the k*order/2 outer columns of an inner
edge are stale, and after halo_steps steps exactly the ghost zone is.  Then every rank refreshes
its ghost zones from the neighbours' owned columns - u, u-, pz, py, G columns each - with one
batched send/recv pair per neighbour (exchange='p2p': NCCL P2P over NVLink on GPUs; gloo in the CPU
tests) - or, the default on GPUs (exchange='peer'), inside ONE native time loop: a pack kernel
stores the columns straight into the neighbour's staging buffer through a CUDA-IPC peer mapping
and raises a flag, an unpack kernel waits for its own flags and fills the ghost columns (csrc/
acoustic.cu, fwi_acoustic_run_slab); no host synchronisation, no collective.
Owned columns therefore always hold the values the undecomposed grid would, and the arithmetic
per cell is that of the single-GPU kernel.  The adjoint is the same recurrence, so it is
decomposed the same way; snapshots and the image stay local to a window and only the owned
columns of the image are kept.

Points.  Sources are injected by every rank whose window contains them (ghost zones included -
the ghost field must evolve like the neighbour's); receivers are sampled by the owning rank only
and the traces are summed over ranks (exact: the other ranks contribute zeros).  The adjoint swaps
the roles.  The model gradient is assembled from the owned image columns with one all_reduce of
the padded image (every element has exactly one non-zero contributor).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from . import _host as H
from .distributed import shard_bounds


def auto_halo_steps(nyp: int, order: int, world: int) -> int:
    """Steps between ghost exchanges when the caller does not fix them: the ghost columns of both
    edges (2 * T * order/2) are held to about 15 % of a slab's width, T within [4, 32].  Measured on
    the SEAM-sized c4dd grid (3549 padded columns, order 8): 2 GPUs prefer T = 32 (310 against 296
    Gcell-updates/s at T = 8), 8 GPUs T = 8 (837 against 802 at T = 16); profiles/r01e_bench_c4dd_*."""
    width = nyp // max(int(world), 1)
    return max(4, min(32, int(0.15 * width / max(int(order), 2))))


class SlabGeometry:
    """Column bookkeeping of one rank's slab (pure integers; tested on CPU)."""

    def __init__(self, nyp: int, order: int, pml_width: int, world: int, rank: int, halo_steps: int):
        if halo_steps < 1:
            raise RuntimeError("halo_steps must be >= 1")
        self.h = order // 2
        self.P = pml_width + self.h
        self.G = halo_steps * self.h
        self.world, self.rank, self.T = world, rank, halo_steps
        self.lo, self.hi = shard_bounds(nyp, rank, world)
        widths = [shard_bounds(nyp, r, world) for r in range(world)]
        need = max(self.G, 2 * self.P + 1)
        if world > 1 and min(b - a for a, b in widths) < need:
            raise RuntimeError(f"slabs of {min(b - a for a, b in widths)} columns are narrower than the "
                               f"{need} columns {halo_steps} halo steps need: use fewer ranks or halo steps")
        self.has_left, self.has_right = rank > 0, rank < world - 1
        self.c0 = self.lo - (self.G if self.has_left else 0)
        self.c1 = self.hi + (self.G if self.has_right else 0)
        self.nyp = nyp

    # window-local column ranges
    @property
    def own(self) -> Tuple[int, int]:
        return self.lo - self.c0, self.hi - self.c0

    def send_left(self) -> Tuple[int, int]:      # my first G owned columns -> left neighbour's right ghost
        a = self.lo - self.c0
        return a, a + self.G

    def send_right(self) -> Tuple[int, int]:     # my last G owned columns -> right neighbour's left ghost
        b = self.hi - self.c0
        return b - self.G, b

    def ghost_left(self) -> Tuple[int, int]:
        return 0, self.G

    def ghost_right(self) -> Tuple[int, int]:
        b = self.hi - self.c0
        return b, b + self.G


def exchange_ghosts(wave: torch.Tensor, geo: SlabGeometry, n_fields: int, n_shots: int, nzp: int,
                    pitch: int, col0: int, group=None) -> None:
    """Refreshes the ghost columns of the live fields (u, u-, pz, py) in place.

    `wave` is the flat workspace [n_fields][S][nzp][pitch] in settled order.  One packed message per
    neighbour and direction; the four P2P operations are batched (a single NCCL group)."""
    if geo.world == 1:
        return
    live = [0, 1, 2, 4] if n_fields >= 6 else [0, 1]
    W = wave.view(n_fields, n_shots, nzp, pitch)
    ops, recvs = [], []

    def ranks(r):
        return r if group is None else dist.get_global_rank(group, r)

    def plan(has, peer, send_cols, ghost_cols):
        if not has:
            return
        a, b = send_cols
        out = W[live, :, :, col0 + a:col0 + b].contiguous()
        buf = torch.empty_like(out)
        ops.append(dist.P2POp(dist.isend, out, ranks(peer), group))
        ops.append(dist.P2POp(dist.irecv, buf, ranks(peer), group))
        recvs.append((buf, ghost_cols))

    plan(geo.has_left, geo.rank - 1, geo.send_left(), geo.ghost_left())
    plan(geo.has_right, geo.rank + 1, geo.send_right(), geo.ghost_right())
    for req in dist.batch_isend_irecv(ops):
        req.wait()
    for buf, (a, b) in recvs:
        W[live, :, :, col0 + a:col0 + b] = buf


class PeerComm:
    """Exchange memory of one rank for the in-loop ghost exchange (include/fwi_b200.h, fwi_slab_link):
    a peer-visible comm block per rank, the two neighbours' blocks mapped through CUDA IPC, and the
    running exchange count.  Creation is collective over the group."""

    _cache = {}

    def __init__(self, group, slot_floats: int, dev: torch.device):
        import ctypes as C
        from . import _lib
        self.group, self.dev = group, dev
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        self.bytes = _lib.SLAB_HEADER_BYTES + 4 * slot_floats * 4
        L = _lib.lib()
        with torch.cuda.device(dev):
            ptr, handle = C.c_void_p(), C.create_string_buffer(64)
            _lib.check(L.fwi_peer_alloc(self.bytes, C.byref(ptr), handle), "fwi_peer_alloc")
            self.ptr = ptr.value
            handles = [None] * self.world
            dist.all_gather_object(handles, bytes(handle.raw), group=group)
            self.peer = [None, None]
            for side, r in ((0, self.rank - 1), (1, self.rank + 1)):
                if 0 <= r < self.world:
                    q = C.c_void_p()
                    _lib.check(L.fwi_peer_open(handles[r], C.byref(q)), "fwi_peer_open")
                    self.peer[side] = q.value
        self.epoch = 0

    @classmethod
    def get(cls, group, slot_floats: int, dev: torch.device) -> "PeerComm":
        key = (id(group) if group is not None else 0, slot_floats, dev.index)
        if key not in cls._cache:
            cls._cache[key] = cls(group, slot_floats, dev)
        return cls._cache[key]

    def link(self, geo: "SlabGeometry", col0: int, nsteps: int):
        """fwi_slab_link for one native time loop of `nsteps` steps; advances the exchange count."""
        from . import _lib
        sl, sr, gl, gr = geo.send_left()[0], geo.send_right()[0], geo.ghost_left()[0], geo.ghost_right()[0]
        ln = _lib.SlabLink(geo.T, geo.G, int(geo.has_left), int(geo.has_right), col0 + sl, col0 + sr, col0 + gl,
                           col0 + gr, self.ptr, self.peer[0], self.peer[1], self.epoch & 0xFFFFFFFF, 0)
        self.epoch += int(_lib.lib().fwi_slab_exchanges(nsteps, geo.T))
        return ln

    def check(self):
        """Raises if a flag wait timed out on this rank (a neighbour died or the calls diverged)."""
        import ctypes as C
        from . import _lib
        st = C.c_uint32(0)
        cur = torch.cuda.current_stream(self.dev)
        with torch.cuda.device(self.dev):
            _lib.check(_lib.lib().fwi_slab_status(self.ptr, C.c_void_p(cur.cuda_stream), C.byref(st)), "fwi_slab_status")
        if st.value != 0:
            type(self)._cache = {k: v for k, v in type(self)._cache.items() if v is not self}
            raise RuntimeError("slab ghost exchange timed out waiting for a neighbouring rank")


class _SlabFunction(torch.autograd.Function):
    """Propagator call on an x-slab decomposition; see the module docstring.  `plan_cls` is the
    per-window time-loop runner: scalar._Plan (the CUDA kernels) in the product."""

    @staticmethod
    def forward(ctx, v, source_amplitudes, x_s, x_r, dt, dx, order, pml_width, vpmax, loc_mode,
                halo_steps, group, plan_cls, exchange="p2p"):
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("slab decomposition needs an initialised torch.distributed process group")
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        if source_amplitudes.dim() != 3:
            raise RuntimeError("source_amplitudes must be [nt, n_shots, n_sources_per_shot]")
        nt, S, Ns = source_amplitudes.shape
        if x_s.dim() != 3 or tuple(x_s.shape) != (S, Ns, 2):
            raise RuntimeError(f"source_locations must be [{S}, {Ns}, 2], got {tuple(x_s.shape)}")
        if x_r.dim() != 3 or x_r.shape[0] != S or x_r.shape[2] != 2:
            raise RuntimeError(f"receiver_locations must be [{S}, n_receivers, 2], got {tuple(x_r.shape)}")
        R = int(x_r.shape[1])
        h = order // 2
        nyp = int(v.shape[1]) + 2 * (pml_width + h)
        if not halo_steps:
            halo_steps = auto_halo_steps(nyp, order, world)
        geo = SlabGeometry(nyp, order, pml_width, world, rank, halo_steps)
        if vpmax is None:
            vpmax = float(v.detach().abs().max())        # the model is replicated: same on every rank
        plan = plan_cls(v, dx, dt, order, pml_width, vpmax, S, window=(geo.c0, geo.c1))
        dev = plan.dev
        s_cell, _, _, bad_s = plan.points(x_s, "source", loc_mode)
        r_cell, _, _, bad_r = plan.points(x_r, "receiver", loc_mode)
        bad = (bad_s | bad_r).reshape(1)
        flag = None
        if bad.is_cuda:
            # read back through pinned memory AFTER this call's time loop is queued (as scalar._ScalarFunction does):
            # the host does not wait for the planning kernels before it launches
            from . import _lib as _L
            flag = _L.ops().flag_begin(bad.to(torch.int32))
        elif bool(bad.item()):
            raise RuntimeError("source/receiver location outside the model")
        s_col, r_col = plan.padded_columns(x_s, loc_mode), plan.padded_columns(x_r, loc_mode)

        def masks(col):
            inside = (col >= geo.c0 + h) & (col < geo.c1 - h)       # the window can inject there
            owned = (col >= geo.lo) & (col < geo.hi)                 # this rank reports it
            return inside.to(torch.float32), owned.to(torch.float32)

        s_in, s_own = masks(s_col)
        r_in, r_own = masks(r_col)
        amp = H.resample_up(source_amplitudes.detach().to(dev, torch.float32), plan.ratio).contiguous()
        nti = nt * plan.ratio
        need_v, need_s = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        q = torch.empty(nti * S * plan.stride, dtype=torch.float32, device=dev) if need_v else None
        r32 = r_cell.to(torch.int32).contiguous()
        lists = plan.lists(s_cell, s_in > 0)
        amp_in = (amp * s_in[None]).contiguous()
        comm = None
        if exchange == "peer" and world > 1:
            live = 4 if plan.n_fields >= 6 else 2
            comm = PeerComm.get(group, live * S * plan.nzp * geo.G, dev)
        T = geo.T
        if comm is not None or world == 1:
            # one native time loop; ghost columns move by peer stores + flags inside it
            rec = plan.run(nsteps=nti, adjoint=0, inj=lists, inj_amp=amp_in, smp_cell=r32, n_smp=R, q_write=q,
                           **({"link": comm.link(geo, plan.col0, nti)} if comm is not None else {}))
            if comm is not None:
                comm.check()
        else:
            wave, parts = plan.new_wave(), []
            for t0 in range(0, nti, T):
                n = min(T, nti - t0)
                qv = None if q is None else q[t0 * S * plan.stride:(t0 + n) * S * plan.stride]
                parts.append(plan.run(nsteps=n, adjoint=0, inj=lists, inj_amp=amp_in[t0:t0 + n], smp_cell=r32,
                                      n_smp=R, q_write=qv, wave=wave))
                wave = plan.settle(wave, n)
                if t0 + n < nti:
                    exchange_ghosts(wave, geo, plan.n_fields, S, plan.nzp, plan.pitch, plan.col0, group)
            rec = torch.cat(parts, 0)
        rec = rec * r_own[None]
        if world > 1:
            dist.all_reduce(rec, op=dist.ReduceOp.SUM, group=group)
        if flag is not None:
            _L.ops().flag_end(flag)
        out = rec[::plan.ratio].contiguous() if plan.ratio > 1 else rec
        ctx.plan, ctx.geo, ctx.group, ctx.q = plan, geo, group, q
        ctx.comm = comm
        ctx.amp = amp
        ctx.cells = (s_cell, r_cell, s_col)
        ctx.masks = (s_own, r_in)
        ctx.dims = (nt, S, Ns, R, nti)
        ctx.need = (need_v, need_s)
        ctx.src_like = source_amplitudes.detach()
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, grad_rec):
        plan, geo, group = ctx.plan, ctx.geo, ctx.group
        s_cell, r_cell, s_col = ctx.cells
        s_own, r_in = ctx.masks
        nt, S, Ns, R, nti = ctx.dims
        need_v, need_s = ctx.need
        dev = plan.dev
        g = grad_rec.to(dev, torch.float32)
        if plan.ratio > 1:
            g_in = torch.zeros(nti, S, R, dtype=torch.float32, device=dev)
            g_in[::plan.ratio] = g
        else:
            g_in = g
        g_in = (g_in * r_in[None]).contiguous()
        image = torch.zeros(plan.n_images * plan.stride, dtype=torch.float32, device=dev) if need_v else None
        s32 = s_cell.to(torch.int32).contiguous()
        rec_lists = plan.lists(r_cell, r_in > 0)
        T = geo.T
        comm = ctx.comm
        if comm is not None or geo.world == 1:
            gs = plan.run(nsteps=nti, adjoint=1, inj=rec_lists, inj_amp=g_in, smp_cell=s32, n_smp=Ns,
                          q_read=ctx.q if need_v else None, image=image,
                          **({"link": comm.link(geo, plan.col0, nti)} if comm is not None else {}))
            if comm is not None:
                comm.check()
        else:
            awave, parts = plan.new_wave(), []
            # segments aligned to the END of the record, like the native loop's adjoint schedule
            ends = list(range(nti, 0, -T))
            for t1 in ends:
                t0 = max(0, t1 - T)
                n = t1 - t0
                qv = None if not need_v else ctx.q[t0 * S * plan.stride:(t0 + n) * S * plan.stride]
                parts.append(plan.run(nsteps=n, adjoint=1, inj=rec_lists, inj_amp=g_in[t0:t0 + n], smp_cell=s32,
                                      n_smp=Ns, q_read=qv, image=image, wave=awave))
                awave = plan.settle(awave, n)
                if t0 > 0:
                    exchange_ghosts(awave, geo, plan.n_fields, S, plan.nzp, plan.pitch, plan.col0, group)
            gs = torch.cat(parts[::-1], 0)
        gs = gs * s_own[None]                                          # [nti, S, Ns], owner's samples only
        if geo.world > 1:
            dist.all_reduce(gs, op=dist.ReduceOp.SUM, group=group)
        grad_v = grad_s = None
        if need_v:
            a, b = geo.own
            Gw = image.view(plan.n_images, plan.nzp, plan.pitch).sum(0)[:, plan.col0 + a:plan.col0 + b]
            Gp = torch.zeros(plan.nzp, geo.nyp, dtype=torch.float32, device=dev)
            Gp[:, geo.lo:geo.hi] = Gw
            if geo.world > 1:
                dist.all_reduce(Gp, op=dist.ReduceOp.SUM, group=group)
            # source term of dJ/dw at the source cells (global padded coordinates, every rank alike)
            z = torch.div(s_cell, plan.pitch, rounding_mode="floor")
            Gp.view(-1).index_add_(0, (z * geo.nyp + s_col).reshape(-1), (gs * ctx.amp).sum(0).reshape(-1))
            grad_v = H.replicate_pad_adjoint((2.0 / plan.v_pad_global) * Gp, plan.P)
        if need_s:
            grad_s = H.resample_up_adjoint(gs, plan.ratio, ctx.src_like.to(dev, torch.float32))
            grad_s = grad_s.to(ctx.src_like.device, ctx.src_like.dtype)
        ctx.q = None
        return (grad_v, grad_s) + (None,) * 12


def slab_columns(nyp: int, world: int) -> List[Tuple[int, int]]:
    """Owned padded-column range of every rank."""
    return [shard_bounds(nyp, r, world) for r in range(world)]
