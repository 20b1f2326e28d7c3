# dev aid: cProfile of the host side of one gradient on the reference-sized 100x300 case
import cProfile, os, pstats, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from physicsbasedfwi2_b200.scalar import Propagator
from physicsbasedfwi2_b200.wavelets import ricker
dev = torch.device("cuda:0")
nz, ny, S, R, nt = 100, 300, 10, 101, 1000
v = (2000 + 1000 * torch.rand(nz, ny)).to(dev)
x_s = torch.zeros(S, 1, 2); x_s[:, 0, 1] = torch.linspace(0, (ny - 1) * 10.0, S)
x_r = torch.zeros(S, R, 2); x_r[:, :, 1] = (torch.arange(R) * (ny // R) * 10.0)[None]
x_s, x_r = x_s.to(dev), x_r.to(dev)
src = ricker(10.0, nt, 1e-3, 0.1).reshape(-1, 1, 1).repeat(1, S, 1).to(dev)
obs = torch.randn(nt, S, R, device=dev)
def grad():
    vv = v.detach().requires_grad_(True)
    rec = Propagator({"vp": vv}, 10.0, vpmax=3000.0)(src, x_s, x_r, 1e-3)
    ((rec - obs) ** 2).mean().backward()
    return vv.grad
for _ in range(5): grad()
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(20): grad()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr); st.sort_stats("cumulative").print_stats(45)
