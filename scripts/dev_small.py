# dev aid: per-step time of the cluster-resident kernel on C1 / C5 grids, forward without / with snapshots, adjoint
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench as B
from physicsbasedfwi2_b200 import _lib
from physicsbasedfwi2_b200.scalar import Propagator
from physicsbasedfwi2_b200.wavelets import ricker
for wl in ("c1", "c5"):
    w = dict(B.WORKLOADS[wl]); nt = 1000; S = w["survey_shots"]
    dev = torch.device("cuda:0")
    v = B.synth_model(w).to(dev)
    x_s, x_r = B.synth_survey(w, S)
    src = ricker(w["freq"], nt, w["dt"], 1.0 / w["freq"]).reshape(-1, 1, 1).repeat(1, S, 1).contiguous().to(dev)
    kw = dict(pml_width=w["pml"], accuracy=w["order"], vpmax=w["vmax"])
    for rep in range(2):
        _lib.LOOP_TIMES = []
        with torch.no_grad():
            Propagator({"vp": v}, w["dx"], **kw)(src, x_s.to(dev), x_r.to(dev), w["dt"])
        vv = v.clone().requires_grad_(True)
        rec = Propagator({"vp": vv}, w["dx"], **kw)(src, x_s.to(dev), x_r.to(dev), w["dt"])
        rec.square().sum().backward()
        torch.cuda.synchronize()
        t = [(tag, a.elapsed_time(b) / n * 1e3) for tag, a, b, n, _ in _lib.LOOP_TIMES]
    print(wl, "us/step: fwd no-snapshot %.2f, fwd snapshot %.2f, adjoint %.2f" % (t[0][1], t[1][1], t[2][1]), flush=True)
