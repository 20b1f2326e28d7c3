"""CPU tests of the oracle itself (no GPU): the C/OpenMP restatement with the hand-derived adjoint
vs the torch-autograd restatement (golden vectors made by tests/golden/make_golden.py), plus the
property / known-answer tests the reference's own tree suggests for this path:

* adjoint dot-product test (implied by the forward/adjoint operator pair at
  seisgan/fwi/pde/seismic/acoustic/acoustic_example.py:56-85),
* gradient Taylor test, slopes 1 and 2 within rtol 0.1
  (seisgan/fwi/pde/seismic/acoustic/gradient_example.py:115-146),
* analytic 2-D Green's function -i*pi*H0^(2)(kr) (seisgan/fwi/pde/seismic/acoustic/accuracy.ipynb:168-188).

PARITY UNPINNED: none of these pins deepwave's numbers (the reference holds none); see DESIGN.md.
"""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import cpu_ref as O
from oracle import torch_ref as T
from tests import _cases as K

GOLDEN = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "o*.npz")))


def _load(path):
    z = np.load(path)
    dx = z["dx"].tolist()
    dx = tuple(dx) if isinstance(dx, list) else float(dx)
    t = {k: torch.from_numpy(z[k]) for k in ("v", "x_s", "x_r", "src", "grad_rec", "rec", "grad_v", "grad_src")}
    return t, dx, float(z["dt"]), int(z["order"]), int(z["pml"])


def test_golden_vectors_exist():
    assert len(GOLDEN) >= 3


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_c_oracle_matches_golden_fp64(path):
    t, dx, dt, order, pml = _load(path)
    rec, gv, gs = O.acoustic_gradient(t["v"], dx, t["src"], t["x_s"], t["x_r"], dt, grad_rec=t["grad_rec"],
                                      order=order, pml_width=pml, dtype=torch.float64)
    # hand-written adjoint == autograd of the time loop, to round-off
    assert K.rel_l2(rec, t["rec"]) < 1e-12
    assert K.rel_l2(gv, t["grad_v"]) < 1e-10
    assert K.rel_l2(gs, t["grad_src"]) < 1e-10


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_c_oracle_fp32_within_parity_tolerance(path):
    t, dx, dt, order, pml = _load(path)
    rec, gv, gs = O.acoustic_gradient(t["v"], dx, t["src"], t["x_s"], t["x_r"], dt, grad_rec=t["grad_rec"],
                                      order=order, pml_width=pml, dtype=torch.float32)
    assert K.rel_l2(rec, t["rec"]) < 1e-5      # north_star trace tolerance
    assert K.rel_l2(gv, t["grad_v"]) < 1e-4    # north_star gradient tolerance
    assert K.rel_l2(gs, t["grad_src"]) < 1e-4


def test_torch_oracle_reproduces_golden():
    """The committed vectors are what oracle/torch_ref.py produces today (smallest case)."""
    t, dx, dt, order, pml = _load(GOLDEN[-1])
    rec = T.acoustic_forward(t["v"].double(), dx, t["src"].double(), t["x_s"], t["x_r"], dt,
                             order=order, pml_width=pml)
    assert K.rel_l2(rec, t["rec"]) < 1e-12


def _small_problem(order=4, pml=8, nz=28, ny=40, S=2, R=6, nt=70):
    v = K.layered_model(nz, ny, seed=4)
    x_s, x_r = K.survey(nz, ny, 10.0, S, R, 1, seed=5, scattered=True)
    src = K.ricker(20.0, nt, 1e-3, 0.05).reshape(-1, 1, 1).repeat(1, S, 1).contiguous()
    return v, x_s, x_r, src


@pytest.mark.parametrize("order", [2, 4, 6, 8])
def test_adjoint_dot_product(order):
    """<F s, r> == <s, F^T r> for the linear map source amplitudes -> receiver traces."""
    v, x_s, x_r, src, = _small_problem(order=order)
    g = torch.Generator().manual_seed(order)
    s = torch.randn(src.shape, generator=g, dtype=torch.float64)
    r = torch.randn(src.shape[0], x_r.shape[0], x_r.shape[1], generator=g, dtype=torch.float64)
    rec, _, gs = O.acoustic_gradient(v, 10.0, s, x_s, x_r, 1e-3, grad_rec=r, order=order, pml_width=8,
                                     dtype=torch.float64)
    lhs, rhs = float((rec * r).sum()), float((s * gs).sum())
    assert abs(lhs - rhs) <= 1e-10 * max(abs(lhs), abs(rhs))


def test_gradient_taylor_slopes():
    """gradient_example.py:115-146: first-order error slope 1, second-order remainder slope 2."""
    v, x_s, x_r, src = _small_problem(order=4, nt=90)
    v_true = v.double()
    v0 = torch.full_like(v_true, float(v_true.mean()))
    v0 = 0.7 * v0 + 0.3 * v_true
    obs = O.acoustic_forward(v_true, 10.0, src, x_s, x_r, 1e-3, order=4, pml_width=8, vpmax=4000.0)
    dm = v_true - v0

    def phi(vm):
        rec = O.acoustic_forward(vm, 10.0, src, x_s, x_r, 1e-3, order=4, pml_width=8, vpmax=4000.0)
        return 0.5 * float(((rec - obs) ** 2).sum())

    rec0, gv, _ = O.acoustic_gradient(v0, 10.0, src, x_s, x_r, 1e-3, observed=obs, order=4, pml_width=8,
                                      vpmax=4000.0, dtype=torch.float64)
    F0 = 0.5 * float(((rec0 - obs) ** 2).sum())
    G = float((gv * dm).sum())
    H = [0.5, 0.25, 0.125, 0.0625, 0.0312, 0.015625, 0.0078125]
    e1 = [abs(phi(v0 + h * dm) - F0) for h in H]
    e2 = [abs(phi(v0 + h * dm) - F0 - h * G) for h in H]
    p1 = np.polyfit(np.log10(H), np.log10(e1), 1)[0]
    p2 = np.polyfit(np.log10(H), np.log10(e2), 1)[0]
    assert np.isclose(p1, 1.0, rtol=0.1), p1
    assert np.isclose(p2, 2.0, rtol=0.1), p2


def test_analytic_green_function_homogeneous():
    """accuracy.ipynb:168-188: u(r, w) = -(i/4) H0^(2)(k r) F(w) for (1/c^2) u_tt - Lap u = f delta(x).
    The scheme injects v^2 dt^2 s(t) into one cell, i.e. f = s * dz * dy.  Also checks the O(dt^2)
    convergence the notebook records (accuracy.ipynb:308-368)."""
    from scipy.special import hankel2
    c0, dx, f0, n = 1500.0, 5.0, 15.0, 161
    errs = []
    for dt, nt in ((5e-4, 600), (2.5e-4, 1200)):
        v = torch.full((n, n), c0, dtype=torch.float64)
        src = T.ricker(f0, nt, dt, 1.2 / f0, dtype=torch.float64).reshape(-1, 1, 1)
        x_s = torch.tensor([[[80 * dx, 80 * dx]]])
        x_r = torch.tensor([[[80 * dx, 120 * dx]]])          # 200 m away, on the grid
        rec = O.acoustic_forward(v, dx, src, x_s, x_r, dt, order=8, pml_width=20,
                                 dtype=torch.float64)[:, 0, 0].numpy()
        N = 8 * nt                                            # long periodic window: no wrap-around
        s = np.zeros(N)
        s[:nt] = src[:, 0, 0].numpy()
        Fw = np.fft.rfft(s) * dx * dx
        w = 2 * np.pi * np.fft.rfftfreq(N, dt)
        U = np.zeros_like(Fw)
        U[1:] = -0.25j * hankel2(0, w[1:] * (40 * dx) / c0) * Fw[1:]
        ua = np.fft.irfft(U, N)[:nt]
        errs.append(np.linalg.norm(rec - ua) / np.linalg.norm(ua))
    assert errs[0] < 5e-3, errs
    assert 3.0 < errs[0] / errs[1] < 5.0, errs               # second order in time


def test_shot_permutation_invariance():
    v, x_s, x_r, src = _small_problem(S=3)
    src = src * torch.tensor([1.0, 0.5, 2.0]).reshape(1, 3, 1)
    g = torch.randn(src.shape[0], 3, x_r.shape[1], generator=torch.Generator().manual_seed(1))
    perm = torch.tensor([2, 0, 1])
    r1, gv1, gs1 = O.acoustic_gradient(v, 10.0, src, x_s, x_r, 1e-3, grad_rec=g, order=4, pml_width=8, dtype=torch.float64)
    r2, gv2, gs2 = O.acoustic_gradient(v, 10.0, src[:, perm], x_s[perm], x_r[perm], 1e-3, grad_rec=g[:, perm],
                                       order=4, pml_width=8, dtype=torch.float64)
    assert K.rel_l2(r2, r1[:, perm]) < 1e-14
    assert K.rel_l2(gv2, gv1) < 1e-12
    assert K.rel_l2(gs2, gs1[:, perm]) < 1e-14


def test_location_outside_model_raises():
    v, x_s, x_r, src = _small_problem()
    bad = x_r.clone()
    bad[0, 0, 1] = 1e6
    with pytest.raises(RuntimeError):
        O.acoustic_forward(v, 10.0, src, x_s, bad, 1e-3)


def test_c_oracle_matches_c1_size_autograd_golden():
    """The hand-written adjoint of oracle/cpu_ref.c against fp64 torch autograd at BASELINE config 1's size
    (tests/golden/c1_autograd_2shots.npz: 100 x 300, 2 of the 10 shots, 600 of the 1000 steps; J = <g, rec> with
    g the MSE residual of the fp64 traces).  Traces and the source gradient are stored in fp32 (hence 3e-7), the
    model gradient in fp64."""
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "c1_autograd_2shots.npz"))
    t = {k: torch.from_numpy(z[k]) for k in ("v", "x_s", "x_r", "src", "obs", "grad_rec", "rec", "grad_v", "grad_src")}
    kw = dict(order=int(z["order"]), pml_width=int(z["pml"]))
    args = (t["v"], float(z["dx"]), t["src"], t["x_s"], t["x_r"], float(z["dt"]))
    rec, gv, gs = O.acoustic_gradient(*args, grad_rec=t["grad_rec"], dtype=torch.float64, **kw)
    assert K.rel_l2(rec, t["rec"]) < 3e-7
    assert K.rel_l2(gv, t["grad_v"]) < 1e-9
    assert K.rel_l2(gs, t["grad_src"]) < 3e-7
    assert abs(float(torch.nn.functional.mse_loss(rec, t["obs"].double())) - float(z["mse"])) < 1e-9 * float(z["mse"])
    # the fp32 build against the fp64 vector: fp32 round-off of 600 steps of the recurrence
    rec32, gv32, gs32 = O.acoustic_gradient(*args, grad_rec=t["grad_rec"], dtype=torch.float32, **kw)
    assert K.rel_l2(rec32, t["rec"]) < 5e-5 and K.rel_l2(gv32, t["grad_v"]) < 2e-4 and K.rel_l2(gs32, t["grad_src"]) < 2e-4
