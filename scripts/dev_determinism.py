"""Run-to-run determinism and exact power-of-two linearity of the C2-size forward pass (dev tool).
FWI_B200_PDL=0 python scripts/dev_determinism.py   # without programmatic dependent launch"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests import _cases as K  # noqa: E402
from physicsbasedfwi2_b200.scalar import Propagator  # noqa: E402

DEV = "cuda:0"
nz, ny, nt, dx, dt = 751, 2301, int(os.environ.get("NT", "1000")), 4.0, 4e-4
v = K.layered_model(nz, ny, seed=5, vmin=1500.0, vmax=4500.0, noise=30.0).to(DEV)
w = K.ricker(15.0, nt, dt, 1 / 15.0)
a, b = (230 * dx, 900 * dx), (290 * dx, 960 * dx)
x_s = torch.tensor([[a], [b]], device=DEV)
x_r = torch.tensor([[b], [a]], device=DEV)
src = w.reshape(-1, 1, 1).repeat(1, 2, 1).contiguous().to(DEV)
prop = Propagator({"vp": v}, dx, pml_width=20, accuracy=8, vpmax=float(os.environ.get("VPMAX", "4600")))
runs = [prop(src, x_s, x_r, dt) for _ in range(3)]
print("PDL", os.environ.get("FWI_B200_PDL", "1"), "peak", float(runs[0].abs().max()))
for i in (1, 2):
    print("run", i, "equal to run 0:", bool(torch.equal(runs[i], runs[0])), K.rel_l2(runs[i], runs[0]))
r4 = prop(4.0 * src, x_s, x_r, dt)
print("4x: equal", bool(torch.equal(r4, 4.0 * runs[0])), K.rel_l2(r4, 4.0 * runs[0]))
print("reciprocity", K.rel_l2(runs[0][:, 0], runs[0][:, 1]))
