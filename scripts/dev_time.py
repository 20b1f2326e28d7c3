# dev aid: times the native time loops on the C2 grid (run under gpurun).  usage: dev_time.py shots nt [reps]
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench as B
from physicsbasedfwi2_b200 import _lib
from physicsbasedfwi2_b200.scalar import Propagator
from physicsbasedfwi2_b200.wavelets import ricker
S = int(sys.argv[1]); nt = int(sys.argv[2]); reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
w = dict(B.WORKLOADS[os.environ.get("WL", "c2")]); w["nt"] = nt
if "PML" in os.environ: w["pml"] = int(os.environ["PML"])
if "NY" in os.environ: w["ny"] = int(os.environ["NY"])
if "NZ" in os.environ: w["nz"] = int(os.environ["NZ"])
w.pop("slabs", None)
if "ORDER" in os.environ: w["order"] = int(os.environ["ORDER"])
dev = torch.device("cuda:0")
v = B.synth_model(w).to(dev)
x_s, x_r = B.synth_survey(w, S)
src = ricker(w["freq"], nt, w["dt"], 1.0 / w["freq"]).reshape(-1, 1, 1).repeat(1, S, 1).contiguous().to(dev)
obs = torch.randn(nt, S, w["n_rec"], device=dev) * 1e-3
for r in range(reps):
    _lib.LOOP_TIMES = []
    t0 = time.time()
    vv = v.detach().requires_grad_(True)
    rec = Propagator({"vp": vv}, w["dx"], pml_width=w["pml"], accuracy=w["order"], vpmax=w["vmax"])(src, x_s.to(dev), x_r.to(dev), w["dt"])
    (0.5 * ((rec - obs) ** 2).sum()).backward()
    torch.cuda.synchronize()
    wall = time.time() - t0
    out = {tag: a.elapsed_time(b) / n * 1e3 for tag, a, b, n, _ in _lib.LOOP_TIMES}
    cells = w["nz"] * w["ny"] * S
    print(f"rep {r} shots {S} nt {nt} wall {wall:.3f}s", {k: f"{x:.1f} us/launch" for k, x in out.items()},
          "GB/s fwd %.0f adj %.0f" % (cells * 20 / out["acoustic_fwd"] / 1e3, cells * 28 / out["acoustic_adj"] / 1e3),
          "finite", bool(torch.isfinite(vv.grad).all()), flush=True)
