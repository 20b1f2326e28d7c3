"""Generates the golden input/output vectors under tests/golden/.

PARITY UNPINNED: the reference's propagator (deepwave, models/networks.py:10) cannot be imported
in this container, so these vectors are produced by oracle/torch_ref.py in float64 (the autograd
restatement of SURVEY.md 7.0), not by the reference.  They pin the hand-derived adjoint of
oracle/cpu_ref.c and the CUDA kernels to that independent statement, and guard against
regressions.  Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import torch_ref as T  # noqa: E402
from tests import _cases as K      # noqa: E402

CASES = {
    # the call shape of models/networks.py:5449-5464 at a size autograd finishes in seconds
    "o4_pml10": dict(nz=24, ny=60, dx=10.0, order=4, pml=10, S=2, R=9, Ns=1, nt=60, dt=1e-3, f=20.0, scattered=False),
    # 8th order, sub-stepped (dt above the CFL limit -> ratio 2), two sources per shot, scattered geometry
    "o8_pml6_sub": dict(nz=20, ny=44, dx=10.0, order=8, pml=6, S=2, R=7, Ns=2, nt=40, dt=2.5e-3, f=12.0, scattered=True),
    # 2nd order, anisotropic spacing
    "o2_pml8_aniso": dict(nz=30, ny=30, dx=(12.0, 8.0), order=2, pml=8, S=3, R=10, Ns=1, nt=50, dt=1e-3, f=25.0, scattered=True),
}


def inputs(c):
    dx0 = c["dx"][1] if isinstance(c["dx"], tuple) else c["dx"]
    v = K.layered_model(c["nz"], c["ny"], seed=1).double()
    x_s, x_r = K.survey(c["nz"], c["ny"], dx0, c["S"], c["R"], c["Ns"], seed=2, scattered=c["scattered"])
    if isinstance(c["dx"], tuple):   # survey() used dy for both axes; rescale z to stay on cells
        x_s[..., 0] *= c["dx"][0] / dx0
        x_r[..., 0] *= c["dx"][0] / dx0
    src = K.ricker(c["f"], c["nt"], c["dt"], 1.0 / c["f"]).reshape(-1, 1, 1).repeat(1, c["S"], c["Ns"])
    src = (src * (1.0 + 0.1 * torch.arange(c["S"] * c["Ns"]).reshape(1, c["S"], c["Ns"]))).double()
    g = torch.Generator().manual_seed(3)
    grad_rec = torch.randn(c["nt"], c["S"], c["R"], generator=g).double()
    return v, x_s, x_r, src, grad_rec


def main():
    here = os.path.dirname(os.path.abspath(__file__))
    for name, c in CASES.items():
        v, x_s, x_r, src, grad_rec = inputs(c)
        rec, gv, gs = T.acoustic_gradient(v, c["dx"], src, x_s, x_r, c["dt"], grad_rec=grad_rec,
                                          order=c["order"], pml_width=c["pml"])
        np.savez_compressed(
            os.path.join(here, name + ".npz"), v=v.float().numpy(), x_s=x_s.numpy(), x_r=x_r.numpy(),
            src=src.float().numpy(), grad_rec=grad_rec.float().numpy(),
            dx=np.asarray(c["dx"], dtype=np.float64), dt=c["dt"], order=c["order"], pml=c["pml"],
            rec=rec.numpy(), grad_v=gv.numpy(), grad_src=gs.numpy())
        print(name, "rec", tuple(rec.shape), float(rec.abs().max()), "grad_v", float(gv.abs().max()))


ELASTIC = {
    # the elastic call shape (vp, vs, rho; explosive source; two-component receivers), incl. a water layer
    "el_o4_pml6": dict(nz=20, ny=28, dx=10.0, order=4, pml=6, S=2, R=5, nt=50, dt=1e-3, f=20.0),
    "el_o8_pml5_sub": dict(nz=18, ny=24, dx=10.0, order=8, pml=5, S=1, R=4, nt=30, dt=2.5e-3, f=15.0),
}


def elastic_inputs(c):
    vp = K.layered_model(c["nz"], c["ny"], seed=4).double()
    vs = vp / 1.732
    vs[:3] = 0.0
    rho = 310.0 * vp ** 0.25
    g = torch.Generator().manual_seed(11)
    x_s = torch.zeros(c["S"], 1, 2)
    x_s[..., 0] = torch.randint(1, c["nz"] - 1, (c["S"], 1), generator=g).float() * c["dx"]
    x_s[..., 1] = torch.randint(1, c["ny"] - 1, (c["S"], 1), generator=g).float() * c["dx"]
    x_r = torch.zeros(c["S"], c["R"], 2)
    x_r[..., 0] = torch.randint(0, c["nz"], (c["S"], c["R"]), generator=g).float() * c["dx"]
    x_r[..., 1] = torch.randint(0, c["ny"], (c["S"], c["R"]), generator=g).float() * c["dx"]
    src = (K.ricker(c["f"], c["nt"], c["dt"], 1.0 / c["f"]).reshape(-1, 1, 1).repeat(1, c["S"], 1) * 1e6).double()
    grad_rec = torch.randn(2, c["nt"], c["S"], c["R"], generator=g).double()
    return vp, vs, rho, x_s, x_r, src, grad_rec


def main_elastic():
    from oracle import elastic_ref as E
    here = os.path.dirname(os.path.abspath(__file__))
    for name, c in ELASTIC.items():
        vp, vs, rho, x_s, x_r, src, grad_rec = elastic_inputs(c)
        rec, gvp, gvs, grho, gs = E.elastic_gradient(vp, vs, rho, c["dx"], src, x_s, x_r, c["dt"], grad_rec,
                                                     order=c["order"], pml_width=c["pml"])
        np.savez_compressed(
            os.path.join(here, name + ".npz"), vp=vp.float().numpy(), vs=vs.float().numpy(), rho=rho.float().numpy(),
            x_s=x_s.numpy(), x_r=x_r.numpy(), src=src.float().numpy(), grad_rec=grad_rec.float().numpy(),
            dx=np.asarray(c["dx"], dtype=np.float64), dt=c["dt"], order=c["order"], pml=c["pml"], rec=rec.numpy(),
            grad_vp=gvp.numpy(), grad_vs=gvs.numpy(), grad_rho=grho.numpy(), grad_src=gs.numpy())
        print(name, "rec", tuple(rec.shape), float(rec.abs().max()), "grad_vp", float(gvp.abs().max()))


def main_c1(shots=(2, 7), nt=600):
    """C1-sized vector (100 x 300, dx 10 m, 15 Hz, order 4, pml 10, MSE against the data of the
    unperturbed layered model): 2 of the 10 shots, 600 of the 1000 steps, fp64 torch autograd."""
    from tests import _baseline as B
    here = os.path.dirname(os.path.abspath(__file__))
    c = B.c1(nt=nt)
    idx = list(shots)
    x_s, x_r, src = c["x_s"][idx], c["x_r"][idx], c["src"][:, idx].double()
    with torch.no_grad():
        obs = T.acoustic_forward(c["v_true"].double(), c["dx"], src, x_s, x_r, c["dt"], order=4, pml_width=10)
        obs = obs.float().double()      # the data is stored (and handed to the kernels) in fp32
    v = c["v"].double().requires_grad_(True)
    s = src.clone().requires_grad_(True)
    rec = T.acoustic_forward(v, c["dx"], s, x_s, x_r, c["dt"], order=4, pml_width=10)
    # J = <grad_rec, rec> with grad_rec = d(MSE)/d(rec) of the fp64 traces, rounded to fp32 as it is stored: the
    # gradient check then tests the adjoint, not the fp32 cancellation in (rec - obs)
    grad_rec = (2.0 * (rec.detach() - obs) / rec.numel()).float().double()
    (rec * grad_rec).sum().backward()
    np.savez_compressed(
        os.path.join(here, "c1_autograd_2shots.npz"), v=c["v"].numpy(), x_s=x_s.numpy(), x_r=x_r.numpy(),
        src=src.float().numpy(), obs=obs.float().numpy(), grad_rec=grad_rec.float().numpy(),
        mse=float(torch.nn.functional.mse_loss(rec.detach(), obs)),
        dx=np.float64(c["dx"]), dt=c["dt"], order=4, pml=10,
        rec=rec.detach().float().numpy(), grad_v=v.grad.numpy(), grad_src=s.grad.float().numpy())
    print("c1_autograd_2shots rec", tuple(rec.shape), float(rec.abs().max()), "grad_v", float(v.grad.abs().max()))


if __name__ == "__main__":
    main()
    main_elastic()
    main_c1()
