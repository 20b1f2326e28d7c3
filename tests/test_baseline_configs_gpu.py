"""Oracle parity on the BASELINE.json configurations themselves (not on toy grids).

* C1 in full (100 x 300, dx 10 m, 15 Hz, 10 shots x 101 receivers, 1000 steps, order 4, pml 10, MSE against the
  data of the unperturbed layered model) against oracle/cpu_ref (fp32 build), on both time-loop implementations.
* C2 on the real 751 x 2301 grid (order 8, pml 20): 2 shots x 600 of the 3000 steps, residual = data of the true
  model - data of the smoothed model, against the C oracle.
* C3 on the real 376 x 1151 elastic grid: 1 shot x 120 steps against oracle/elastic_ref.py (fp32).
* a committed fp64 torch-AUTOGRAD golden vector at C1 size (tests/golden/c1_autograd_*.npz: 2 shots x 600
  steps), the derivation of the adjoint that is independent of the hand-written one.

Tolerances are the north star's: relative L2 <= 1e-5 on traces, <= 1e-4 on gradients (fp32 against the fp32
oracle).  Against fp64 references the bound is the fp32 round-off of the recurrence itself (the fp32 and fp64
builds of the oracle differ by that much), stated next to each check.

Call sequence followed: models/networks.py:5445-5476 (propagate, misfit) and :5491-5493 (backward)."""
import os

import numpy as np
import pytest
import torch

from tests import _baseline as B
from tests import _cases as K

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TRACE_TOL, GRAD_TOL = 1e-5, 1e-4


@pytest.mark.parametrize("small", ["1", "0"], ids=["cluster-resident", "step-kernel"])
def test_c1_full_matches_the_oracle(cuda_lib, small, monkeypatch):
    from oracle import cpu_ref as O
    from physicsbasedfwi2_b200.scalar import Propagator
    monkeypatch.setenv("FWI_B200_SMALL", small)
    c = B.c1()
    kw = dict(order=c["order"], pml_width=c["pml"], dtype=torch.float32)
    obs = O.acoustic_forward(c["v_true"], c["dx"], c["src"], c["x_s"], c["x_r"], c["dt"], **kw)
    n = obs.numel()
    rec_o = O.acoustic_forward(c["v"], c["dx"], c["src"], c["x_s"], c["x_r"], c["dt"], **kw)
    g = 2.0 * (rec_o - obs) / n                                  # d(MSE)/d(rec) on the oracle's own traces
    _, gv_o, gs_o = O.acoustic_gradient(c["v"], c["dx"], c["src"], c["x_s"], c["x_r"], c["dt"], grad_rec=g, **kw)

    vg = c["v"].to(DEV).requires_grad_(True)
    sg = c["src"].to(DEV).requires_grad_(True)
    rec = Propagator({"vp": vg}, c["dx"])(sg, c["x_s"].to(DEV), c["x_r"].to(DEV), c["dt"])   # deepwave defaults: order 4, pml 10
    loss = torch.nn.functional.mse_loss(rec, obs.to(DEV))
    loss.backward()
    assert tuple(rec.shape) == (1000, 10, 101)
    et, ev, es = K.rel_l2(rec, rec_o), K.rel_l2(vg.grad, gv_o), K.rel_l2(sg.grad, gs_o)
    assert et <= TRACE_TOL, f"C1 traces rel-L2 {et:.3e}"
    assert ev <= GRAD_TOL, f"C1 grad_v rel-L2 {ev:.3e}"
    assert es <= GRAD_TOL, f"C1 grad_src rel-L2 {es:.3e}"
    assert float(loss) > 0 and float(vg.grad.abs().max()) > 0


def test_c2_grid_two_shots_match_the_oracle(cuda_lib):
    from oracle import cpu_ref as O
    from physicsbasedfwi2_b200.scalar import Propagator
    c = B.c2(nt=600)
    kw = dict(order=c["order"], pml_width=c["pml"], vpmax=c["vmax"], dtype=torch.float32)
    recs, obss, gv_o = [], [], 0
    for s in range(c["S"]):                                       # one shot at a time: 4.5 GB of snapshots each
        sl = slice(s, s + 1)
        args = (c["dx"], c["src"][:, sl], c["x_s"][sl], c["x_r"][sl], c["dt"])
        obs = O.acoustic_forward(c["v"], *args, **kw)
        r, gv, _ = O.acoustic_gradient(c["v_smooth"], *args, observed=obs, **kw)     # grad_rec = rec - obs
        recs.append(r)
        obss.append(obs)
        gv_o = gv_o + gv
    rec_o, obs = torch.cat(recs, 1), torch.cat(obss, 1)

    vg = c["v_smooth"].to(DEV).requires_grad_(True)
    prop = Propagator({"vp": vg}, c["dx"], pml_width=c["pml"], accuracy=c["order"], vpmax=c["vmax"])
    rec = prop(c["src"].to(DEV), c["x_s"].to(DEV), c["x_r"].to(DEV), c["dt"])
    (0.5 * ((rec - obs.to(DEV)) ** 2).sum()).backward()
    et, ev = K.rel_l2(rec, rec_o), K.rel_l2(vg.grad, gv_o)
    assert float(rec_o.abs().max()) > 0 and float((rec_o - obs).abs().max()) > 0
    assert et <= TRACE_TOL, f"C2 traces rel-L2 {et:.3e}"
    assert ev <= GRAD_TOL, f"C2 grad_v rel-L2 {ev:.3e}"


def test_c3_grid_one_shot_matches_the_elastic_oracle(cuda_lib):
    from oracle import elastic_ref as E
    import physicsbasedfwi2_b200 as pkg
    c = B.c3(nt=120)
    g = torch.randn(2, c["nt"], 1, c["R"], generator=torch.Generator().manual_seed(2))
    o = E.elastic_gradient(c["vp"], c["vs"], c["rho"], c["dx"], c["src"], c["x_s"], c["x_r"], c["dt"], g,
                           order=c["order"], pml_width=c["pml"], vpmax=c["vmax"])
    leaves = [t.to(DEV).requires_grad_(True) for t in (c["vp"], c["vs"], c["rho"], c["src"])]
    prop = pkg.Propagator({"vp": leaves[0], "vs": leaves[1], "rho": leaves[2]}, c["dx"], pml_width=c["pml"],
                          accuracy=c["order"], vpmax=c["vmax"])
    rec = prop(leaves[3], c["x_s"].to(DEV), c["x_r"].to(DEV), c["dt"])
    (rec * g.to(DEV)).sum().backward()
    assert float(o[0].abs().max()) > 0
    assert K.rel_l2(rec, o[0]) <= TRACE_TOL, K.rel_l2(rec, o[0])
    for name, leaf, want in zip(("vp", "vs", "rho", "src"), leaves, o[1:]):
        assert K.rel_l2(leaf.grad, want) <= GRAD_TOL, (name, K.rel_l2(leaf.grad, want))


def _c1_golden():
    import glob
    return sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "c1_autograd_*.npz")))


@pytest.mark.parametrize("small", ["1", "0"], ids=["cluster-resident", "step-kernel"])
@pytest.mark.parametrize("path", _c1_golden(), ids=lambda p: p.split("/")[-1][:-4])
def test_c1_size_autograd_golden(cuda_lib, path, small, monkeypatch):
    """fp64 torch-autograd vectors (tests/golden/make_golden.py main_c1): fp32 kernels against an fp64 reference
    over 600 steps, so the bounds are the fp32 round-off of the recurrence (5e-5 traces / 2e-4 gradients, the
    same bounds tests/test_acoustic_gpu.py uses against the fp64 oracle)."""
    from physicsbasedfwi2_b200.scalar import Propagator
    monkeypatch.setenv("FWI_B200_SMALL", small)
    z = np.load(path)
    vg = torch.from_numpy(z["v"]).to(DEV).requires_grad_(True)
    sg = torch.from_numpy(z["src"]).to(DEV).requires_grad_(True)
    prop = Propagator({"vp": vg}, float(z["dx"]), pml_width=int(z["pml"]), accuracy=int(z["order"]))
    rec = prop(sg, torch.from_numpy(z["x_s"]).to(DEV), torch.from_numpy(z["x_r"]).to(DEV), float(z["dt"]))
    (rec * torch.from_numpy(z["grad_rec"]).to(DEV)).sum().backward()     # J = <g, rec>, g = the fp64 MSE residual
    mse = float(torch.nn.functional.mse_loss(rec.detach(), torch.from_numpy(z["obs"]).to(DEV)))
    assert abs(mse - float(z["mse"])) <= 1e-3 * float(z["mse"])
    assert K.rel_l2(rec, torch.from_numpy(z["rec"])) <= 5e-5
    assert K.rel_l2(vg.grad, torch.from_numpy(z["grad_v"])) <= 2e-4
    assert K.rel_l2(sg.grad, torch.from_numpy(z["grad_src"])) <= 2e-4
