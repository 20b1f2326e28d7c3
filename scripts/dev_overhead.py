# dev aid: wall time of one gradient vs the time inside the native loops, on a small reference-sized case
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from physicsbasedfwi2_b200 import _lib
from physicsbasedfwi2_b200.scalar import Propagator
from physicsbasedfwi2_b200.wavelets import ricker
dev = torch.device("cuda:0")
for (nz, ny, S, R, nt) in ((70, 70, 5, 70, 1000), (100, 300, 10, 101, 1000), (151, 200, 18, 200, 4001)):
    v = (2000 + 1000 * torch.rand(nz, ny)).to(dev)
    x_s = torch.zeros(S, 1, 2); x_s[:, 0, 1] = torch.linspace(0, (ny - 1) * 10.0, S)
    x_r = torch.zeros(S, R, 2); x_r[:, :, 1] = (torch.arange(R) * (ny // R) * 10.0)[None]
    x_s, x_r = x_s.to(dev), x_r.to(dev)
    src = ricker(10.0, nt, 1e-3, 0.1).reshape(-1, 1, 1).repeat(1, S, 1).to(dev)
    obs = torch.randn(nt, S, R, device=dev)
    def grad():
        vv = v.detach().requires_grad_(True)
        rec = Propagator({"vp": vv}, 10.0)(src, x_s, x_r, 1e-3)
        ((rec - obs) ** 2).mean().backward()
        return vv.grad
    for _ in range(3): grad()
    torch.cuda.synchronize()
    _lib.LOOP_TIMES = []
    t0 = time.perf_counter()
    for _ in range(5): grad()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / 5
    loops = sum(a.elapsed_time(b) for _, a, b, _, _ in _lib.LOOP_TIMES) / 5
    _lib.LOOP_TIMES = None
    print(f"{nz}x{ny} S={S} nt={nt}: wall {wall*1e3:.2f} ms per gradient, native loops {loops:.2f} ms, rest {wall*1e3-loops:.2f} ms", flush=True)
