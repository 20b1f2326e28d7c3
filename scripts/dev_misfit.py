# dev aid: fused trace misfit vs the same expression in eager torch on the GPU (run under gpurun)
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from physicsbasedfwi2_b200.misfit import trace_misfit
from oracle import misfit_ref as M
dev = torch.device("cuda:0")
for shape in ((3000, 6, 576), (4001, 18, 200), (3000, 32, 576)):
    pred, direct, obs = (torch.randn(*shape, device=dev) for _ in range(3))
    def fused():
        p = pred.detach().requires_grad_(True)
        trace_misfit(p, obs, direct).backward()
        return p.grad
    def eager():
        p = pred.detach().requires_grad_(True)
        M.trace_misfit(p, obs, direct).backward()
        return p.grad
    out = {}
    for name, fn in (("fused", fused), ("eager", eager)):
        for _ in range(3): g = fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): g = fn()
        e1.record(); torch.cuda.synchronize()
        out[name] = (e0.elapsed_time(e1) / 10, g)
    n = pred.numel()
    err = float((out["fused"][1] - out["eager"][1]).norm() / out["eager"][1].norm())
    print(f"misfit {shape}: fused {out['fused'][0]*1e3:.0f} us ({n*4*8/out['fused'][0]/1e6:.0f} GB/s at 8 passes), "
          f"eager torch {out['eager'][0]*1e3:.0f} us, x{out['eager'][0]/out['fused'][0]:.1f}, rel diff {err:.1e}", flush=True)
