"""x-slab domain decomposition (physicsbasedfwi2_b200/slab.py): geometry on one process and the
full forward + adjoint on world_size 2 / 3 over gloo, with the C oracle as the per-window time
loop (tests/_cpu_plan.py).  The decomposed gradient must equal the undecomposed oracle gradient."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from physicsbasedfwi2_b200.slab import SlabGeometry, auto_halo_steps, slab_columns


def test_geometry_tiles_the_padded_grid():
    for nyp, world, T, order, pw in ((324, 2, 4, 4, 10), (2349, 8, 8, 8, 20), (200, 3, 3, 2, 5)):
        cols = slab_columns(nyp, world)
        assert cols[0][0] == 0 and cols[-1][1] == nyp
        assert all(cols[i][1] == cols[i + 1][0] for i in range(world - 1))
        for r in range(world):
            g = SlabGeometry(nyp, order, pw, world, r, T)
            assert g.G == T * order // 2
            assert g.c0 == max(0, g.lo - g.G) and g.c1 == min(nyp, g.hi + g.G)
            a, b = g.own
            assert b - a == g.hi - g.lo
            if g.has_right:   # what I send right lands exactly in the neighbour's left ghost
                n = SlabGeometry(nyp, order, pw, world, r + 1, T)
                s0, s1 = g.send_right()
                g0, g1 = n.ghost_left()
                assert (s0 + g.c0, s1 + g.c0) == (g0 + n.c0, g1 + n.c0)
                s0, s1 = n.send_left()
                g0, g1 = g.ghost_right()
                assert (s0 + n.c0, s1 + n.c0) == (g0 + g.c0, g1 + g.c0)


def test_geometry_rejects_slabs_narrower_than_the_ghost_zone():
    with pytest.raises(RuntimeError):
        SlabGeometry(120, 8, 4, 4, 0, 16)
    with pytest.raises(RuntimeError):
        SlabGeometry(120, 8, 4, 2, 0, 0)


def test_auto_halo_steps_follows_the_slab_width():
    # the two measured optima on the SEAM-sized grid (profiles/r01e_bench_c4dd_*): 2 ranks -> 32, 8 ranks -> 8
    assert auto_halo_steps(3549, 8, 2) == 32
    assert auto_halo_steps(3549, 8, 8) == 8
    assert auto_halo_steps(3549, 8, 4) == 16
    assert auto_halo_steps(104, 4, 2) == 4          # floor
    assert auto_halo_steps(100000, 2, 1) == 32      # ceiling
    for world in (1, 2, 3, 5, 8):                   # the choice always fits the geometry it is made for
        T = auto_halo_steps(3549, 8, world)
        SlabGeometry(3549, 8, 20, world, world - 1, T)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, T, order, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        from oracle import cpu_ref as O
        from physicsbasedfwi2_b200.slab import _SlabFunction
        from tests import _cases as K
        from tests._cpu_plan import CpuPlan
        nz, ny, S, R, nt, pw = 20, 90, 3, 12, 70, 5
        v = K.layered_model(nz, ny, seed=3)
        x_s, x_r = K.survey(nz, ny, 10.0, S, R, 1, seed=4)
        x_s[1, 0, 0] = 5 * 10.0                                     # one buried source
        x_r[:, ::3, 0] = 7 * 10.0
        src = K.ricker(20.0, nt, 1e-3, 0.05).reshape(-1, 1, 1).repeat(1, S, 1).contiguous()
        g = torch.randn(nt, S, R, generator=torch.Generator().manual_seed(5))
        vg = v.clone().requires_grad_(True)
        sg = src.clone().requires_grad_(True)
        rec = _SlabFunction.apply(vg, sg, x_s, x_r, 1e-3, 10.0, order, pw, None, "trunc", T, None, CpuPlan)
        (rec * g).sum().backward()
        r0, gv0, gs0 = O.acoustic_gradient(v, 10.0, src, x_s, x_r, 1e-3, grad_rec=g, order=order, pml_width=pw,
                                           dtype=torch.float32)
        out[rank] = (K.rel_l2(rec, r0), K.rel_l2(vg.grad, gv0), K.rel_l2(sg.grad, gs0))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,T,order", [(2, 4, 4), (3, 3, 8), (2, 1, 2), (1, 4, 4), (2, 0, 4)])
def test_slab_gradient_equals_undecomposed(world, T, order):       # T = 0: the interval is chosen by auto_halo_steps
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), T, order, out), nprocs=world, join=True)
    assert len(out) == world
    for r in range(world):
        et, ev, es = out[r]
        assert et < 2e-6 and ev < 2e-5 and es < 2e-5, (r, out[r])
