"""GPU parity: CUDA propagator (through the C ABI) vs the CPU oracle on the same seeded inputs.

Tolerances are the ones BASELINE.json's north_star states: relative L2 <= 1e-5 on receiver
traces and <= 1e-4 on gradients, fp32.  The oracle is oracle/cpu_ref.c (fp32 build, same
arithmetic order) cross-checked against its fp64 build.
"""
import pytest
import torch

from tests import _cases as K

pytestmark = pytest.mark.gpu

TRACE_TOL = 1e-5
GRAD_TOL = 1e-4

CASES = [
    # nz, ny, dx, order, pml, S, R, Ns, nt, dt, freq, scattered
    dict(nz=40, ny=150, dx=10.0, order=4, pml=10, S=3, R=17, Ns=1, nt=120, dt=1e-3, f=15.0, scattered=False),
    dict(nz=40, ny=150, dx=10.0, order=8, pml=12, S=2, R=20, Ns=2, nt=100, dt=2e-3, f=12.0, scattered=True),
    dict(nz=70, ny=70, dx=15.0, order=2, pml=10, S=5, R=70, Ns=1, nt=150, dt=1e-3, f=25.0, scattered=False),
    dict(nz=33, ny=261, dx=10.0, order=6, pml=7, S=4, R=31, Ns=1, nt=90, dt=1.5e-3, f=15.0, scattered=True),
    dict(nz=50, ny=131, dx=10.0, order=4, pml=0, S=2, R=9, Ns=1, nt=80, dt=1e-3, f=15.0, scattered=True),
]


def _run_case(c, spg=0):
    from oracle import cpu_ref as O
    from physicsbasedfwi2_b200.scalar import Propagator
    v = K.layered_model(c["nz"], c["ny"], seed=1)
    x_s, x_r = K.survey(c["nz"], c["ny"], c["dx"], c["S"], c["R"], c["Ns"], seed=2, scattered=c["scattered"])
    src = K.ricker(c["f"], c["nt"], c["dt"], 1.0 / c["f"]).reshape(-1, 1, 1).repeat(1, c["S"], c["Ns"]).contiguous()
    src = src * (1.0 + 0.1 * torch.arange(c["S"] * c["Ns"]).reshape(1, c["S"], c["Ns"]))
    g = torch.Generator().manual_seed(3)
    grad_rec = torch.randn(c["nt"], c["S"], c["R"], generator=g)

    dev = torch.device("cuda:0")
    vg = v.to(dev).requires_grad_(True)
    sg = src.to(dev).requires_grad_(True)
    prop = Propagator({"vp": vg}, c["dx"], pml_width=c["pml"], accuracy=c["order"], shots_per_group=spg)
    rec = prop(sg, x_s.to(dev), x_r.to(dev), c["dt"])
    (rec * grad_rec.to(dev)).sum().backward()

    out = {}
    for name, dt_ in (("f32", torch.float32), ("f64", torch.float64)):
        r, gv, gs = O.acoustic_gradient(v, c["dx"], src, x_s, x_r, c["dt"], grad_rec=grad_rec,
                                        order=c["order"], pml_width=c["pml"], dtype=dt_)
        out[name] = (K.rel_l2(rec, r), K.rel_l2(vg.grad, gv), K.rel_l2(sg.grad, gs))
    return out


@pytest.mark.parametrize("idx", range(len(CASES)))
def test_traces_and_gradients_match_oracle(cuda_lib, idx):
    errs = _run_case(CASES[idx])
    for name, (et, ev, es) in errs.items():
        assert et <= TRACE_TOL, f"{name}: trace rel-L2 {et:.3e} > {TRACE_TOL}"
        assert ev <= GRAD_TOL, f"{name}: grad_v rel-L2 {ev:.3e} > {GRAD_TOL}"
        assert es <= GRAD_TOL, f"{name}: grad_src rel-L2 {es:.3e} > {GRAD_TOL}"


@pytest.mark.parametrize("spg", [1, 2, 3])
def test_shot_grouping_does_not_change_results(cuda_lib, spg):
    errs = _run_case(CASES[0], spg=spg)
    et, ev, es = errs["f32"]
    assert et <= TRACE_TOL and ev <= GRAD_TOL and es <= GRAD_TOL, errs


def _golden():
    import glob
    import os
    return sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")))


@pytest.mark.parametrize("path", _golden(), ids=lambda p: p.split("/")[-1][:-4])
def test_cuda_matches_committed_golden_vectors(cuda_lib, path):
    """Golden vectors from tests/golden/make_golden.py (torch-autograd oracle, fp64)."""
    import numpy as np
    from physicsbasedfwi2_b200.scalar import Propagator
    z = np.load(path)
    dev = torch.device("cuda:0")
    dx = z["dx"].tolist()
    vg = torch.from_numpy(z["v"]).to(dev).requires_grad_(True)
    sg = torch.from_numpy(z["src"]).to(dev).requires_grad_(True)
    prop = Propagator({"vp": vg}, dx, pml_width=int(z["pml"]), accuracy=int(z["order"]))
    rec = prop(sg, torch.from_numpy(z["x_s"]).to(dev), torch.from_numpy(z["x_r"]).to(dev), float(z["dt"]))
    (rec * torch.from_numpy(z["grad_rec"]).to(dev)).sum().backward()
    assert K.rel_l2(rec, torch.from_numpy(z["rec"])) <= TRACE_TOL
    assert K.rel_l2(vg.grad, torch.from_numpy(z["grad_v"])) <= GRAD_TOL
    assert K.rel_l2(sg.grad, torch.from_numpy(z["grad_src"])) <= GRAD_TOL
