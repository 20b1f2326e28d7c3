"""Fixtures produced by the REFERENCE'S OWN code: `AutoMarmousi22_Net.prop` (models/networks.py:5314-5510) and
`Auto_Net.prop` (:2711-2850), imported unmodified from /root/reference and run on seeded inputs
(tests/_prop_cases.py) with the oracle-backed deepwave stand-in of tests/_ref_harness.py.

    python tests/golden/make_prop_golden.py [auto] [auto22]        (auto22 takes a few minutes of CPU)

Stored per case: what `prop` returned (the masked / weighted gradient, the loss), a subsample of the
traces of its gradient-carrying deepwave call, and the arguments it passed to deepwave (locations after the
reference's own shuffle, strides, shapes) - the call-site contract tests/test_dropin_gpu.py replays on the B200.
The numbers come from the CPU oracle (PARITY UNPINNED for the propagator arithmetic, see oracle/cpu_ref.c); what
these fixtures pin is everything the reference does around the call.
"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import _prop_cases as PC   # noqa: E402
from tests import _ref_harness as RH  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def _save(name, grad, loss, log, sub=(16, 1)):
    call = log[-1]                     # the gradient-carrying call (the first is the constant direct-wave model)
    st, sr = sub                       # traces are kept every st-th sample, every sr-th receiver (fixtures stay small)
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"), grad=grad.detach().cpu().numpy(), loss=np.float64(loss if loss is not None else np.nan),
        pred_sub=call["rec"][::st, :, ::sr].numpy(), x_s=call["x_s"].cpu().numpy(), x_r=call["x_r"].cpu().numpy(),
        src_shape=np.asarray(call["src_shape"]), dt=call["dt"], dx=np.float64(call["dx"]),
        n_calls=len(log), direct_sub=log[0]["rec"][::st, :, ::sr].numpy(), direct_vp=float(log[0]["v"].flatten()[0]),
        contiguous=np.asarray(call["contiguous"]))
    print(name, "grad", tuple(grad.shape), float(grad.abs().max()), "loss", loss, "calls", len(log))


def run_auto22():
    log = []
    N = RH.load_networks(RH.oracle_deepwave(log))
    net = N.AutoMarmousi22_Net(18, 1)
    c = PC.auto22_inputs()
    torch.manual_seed(c["seed"])
    t0 = time.time()
    grad, loss = net.prop(c["inputs"], c["vel"], 0, 1, c["mintrue"], c["maxtrue"], c["true"])
    print(f"AutoMarmousi22_Net.prop: {time.time() - t0:.0f} s")
    return grad, loss, log


def run_auto():
    log = []
    N = RH.load_networks(RH.oracle_deepwave(log))
    net = N.Auto_Net(5, 1)
    c = PC.auto_inputs()
    torch.manual_seed(c["seed"])
    grad = net.prop(c["inputs"], c["vel"].as_subclass(PC.CpuPinned), 0, 1, c["mintrue"], c["maxtrue"])
    return grad.as_subclass(torch.Tensor), None, log


if __name__ == "__main__":
    which = sys.argv[1:] or ["auto", "auto22"]
    if "auto" in which:
        _save("prop_auto", *run_auto())
    if "auto22" in which:
        _save("prop_auto22", *run_auto22(), sub=(64, 4))
