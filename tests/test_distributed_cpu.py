"""World-size-2 gloo test of the multi-GPU host logic: shot sharding + the single gradient
allreduce reproduce the single-process gradient.  The per-rank 'propagator' here is the CPU
oracle (tests may use it); on GPUs the same helpers wrap the CUDA Propagator (bench.py)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from physicsbasedfwi2_b200.distributed import allreduce_gradient, shard_bounds, shard_shots


def test_shard_bounds_cover_all_shots_once():
    for n in (1, 7, 10, 256):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                lo, hi = shard_bounds(n, r, world)
                assert 0 <= lo <= hi <= n
                seen += list(range(lo, hi))
            assert seen == list(range(n))
            sizes = [shard_bounds(n, r, world)[1] - shard_bounds(n, r, world)[0] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(RuntimeError):
        shard_bounds(4, 2, 2)


def test_allreduce_is_identity_without_process_group():
    g = torch.ones(3, 4)
    out, loss = allreduce_gradient(g, torch.tensor(2.0))
    assert out is g and float(loss) == 2.0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(1)
        from oracle import cpu_ref as O
        from tests import _cases as K
        v = K.layered_model(24, 50, seed=1)
        x_s, x_r = K.survey(24, 50, 10.0, 5, 8, 1, seed=2)
        src = K.ricker(20.0, 50, 1e-3, 0.05).reshape(-1, 1, 1).repeat(1, 5, 1).contiguous()
        obs = torch.randn(50, 5, 8, generator=torch.Generator().manual_seed(9)).double() * 1e-2
        src_l, xs_l, xr_l, obs_l = shard_shots([src, x_s, x_r, obs], [1, 0, 0, 1], rank, world)
        rec, gv, _ = O.acoustic_gradient(v, 10.0, src_l.contiguous(), xs_l, xr_l, 1e-3,
                                         observed=obs_l, order=4, pml_width=6, dtype=torch.float64)
        loss = 0.5 * ((rec - obs_l) ** 2).sum()
        g, l = allreduce_gradient(gv, loss)
        if rank == 0:
            rec_a, gv_a, _ = O.acoustic_gradient(v, 10.0, src, x_s, x_r, 1e-3, observed=obs, order=4,
                                                 pml_width=6, dtype=torch.float64)
            loss_a = 0.5 * ((rec_a - obs) ** 2).sum()
            out["grad_err"] = float((g - gv_a).norm() / gv_a.norm())
            out["loss_err"] = abs(float(l) - float(loss_a)) / float(loss_a)
    finally:
        dist.destroy_process_group()


def test_two_rank_shot_sharding_equals_single_process():
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert out["grad_err"] < 1e-12, dict(out)
    assert out["loss_err"] < 1e-12, dict(out)
