# dev aid: what the memory system sustains for the forward step's read/write mix (run under gpurun)
import sys, os, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from physicsbasedfwi2_b200 import _lib
L = _lib.lib()
L.fwi_debug_stream_step.argtypes = [C.c_void_p] * 4 + [C.c_int64, C.c_int64, C.c_int, C.c_void_p]
dev = torch.device("cuda:0")
S, nzp, pitch, nt = 6, 799, 2400, 40
n = S * nzp * pitch
u = torch.randn(n, device=dev); um = torch.randn(n, device=dev); w = torch.rand(nzp * pitch, device=dev) * 0.1
q = torch.empty(nt, n, device=dev)
st = torch.cuda.current_stream().cuda_stream
for blocks in (148 * 2, 148 * 4, 148 * 8, 148 * 16, 148 * 32):
    for rep in range(2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        a, b = u, um
        for t in range(nt):
            L.fwi_debug_stream_step(a.data_ptr(), b.data_ptr(), w.data_ptr(), q[t].data_ptr(), n, nzp * pitch, blocks, st)
            a, b = b, a
        e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / nt * 1e3
    print(f"blocks {blocks}: {us:.1f} us/launch, algorithmic (5 x 4 B/cell, padded cells) {n * 20 / us / 1e3:.0f} GB/s", flush=True)
