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


if __name__ == "__main__":
    main()
