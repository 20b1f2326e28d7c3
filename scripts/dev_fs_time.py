"""Times the elastic gradient on the C3 grid with and without the free surface (dev aid): python scripts/dev_fs_time.py"""
import time
import torch
import physicsbasedfwi2_b200 as pkg
from physicsbasedfwi2_b200.wavelets import ricker

dev = torch.device("cuda:0")
nz, ny, dx, nt, dt, S, R = 376, 1151, 10.0, 300, 1e-3, 5, 200
vp = (2000.0 + 1500.0 * torch.linspace(0, 1, nz).reshape(-1, 1).repeat(1, ny)).to(dev)
vs, rho = vp / 1.732, 310.0 * vp ** 0.25
x_s = torch.zeros(S, 1, 2); x_s[:, 0, 0] = 2 * dx; x_s[:, 0, 1] = torch.linspace(100, (ny - 10) * dx, S)
x_r = torch.zeros(S, R, 2); x_r[:, :, 0] = dx; x_r[:, :, 1] = torch.linspace(50, (ny - 5) * dx, R)[None]
src = ricker(8.0, nt, dt, 0.15).reshape(-1, 1, 1).repeat(1, S, 1).to(dev)
for fs in (False, True, False, True):
    ts = []
    for it in range(4):
        leaves = [t.clone().requires_grad_(True) for t in (vp, vs, rho)]
        torch.cuda.synchronize(); t0 = time.perf_counter()
        prop = pkg.Propagator({"vp": leaves[0], "vs": leaves[1], "rho": leaves[2]}, dx, pml_width=20, accuracy=4, free_surface=fs)
        rec = prop(src, x_s.to(dev), x_r.to(dev), dt)
        (rec ** 2).sum().backward()
        torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
    inner = 2 * S * nz * ny * nt
    print(f"free_surface={fs}: {min(ts) * 1e3:.1f} ms per gradient ({nt} steps, {S} shots)")
