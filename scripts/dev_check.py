# dev aid: prints CUDA-vs-oracle errors per case (run under gpurun)
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tests.test_acoustic_gpu import CASES, _run_case
for i, c in enumerate(CASES):
    for spg in (0, 6):
        try:
            t = time.time(); e = _run_case(c, spg); torch.cuda.synchronize()
            print(i, "chunk_rows", spg, {k: tuple(f"{x:.2e}" for x in v) for k, v in e.items()}, f"{time.time()-t:.1f}s", flush=True)
        except Exception as ex:
            print(i, "chunk_rows", spg, "FAILED", repr(ex)[:300], flush=True)
