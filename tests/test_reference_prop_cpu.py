"""The reference's own code on the CPU (needs /root/reference: skipped on the GPU box, where it does not exist).

* `Auto_Net.prop` (models/networks.py:2711), imported UNMODIFIED, run on the seeded inputs with the oracle-backed
  deepwave stand-in: must reproduce the committed fixture tests/golden/prop_auto.npz - the fixture is what the
  reference code computes, not what we think it computes.
* The line-by-line replays of tests/_prop_cases.py (what tests/test_dropin_gpu.py runs on the B200 with the
  product) must give the same numbers as the reference methods when fed the same stand-in.
* `FullBatchLBFGS` (functions/LBFGS.py:956-1072, unmodified) with a closure shaped like
  models/AutoElMar22LBFGS_model.py:477-498 minimises a small acoustic misfit through a deepwave-style
  autograd.Function: several closure evaluations per step (Wolfe line search), i.e. the re-entrancy pattern
  the product's Function must support (tests/test_dropin_gpu.py runs the same loop on the GPU with torch.optim.LBFGS,
  the reference's optimiser cannot travel).
"""
import os

import numpy as np
import pytest
import torch

from tests import _cases as K
from tests import _prop_cases as PC
from tests import _ref_harness as RH

pytestmark = pytest.mark.skipif(not RH.available(), reason="/root/reference is not present on this machine")
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_reference_auto_net_prop_reproduces_the_fixture():
    log = []
    N = RH.load_networks(RH.oracle_deepwave(log))
    c = PC.auto_inputs()
    net = N.Auto_Net(5, 1)                                         # (weight init draws random numbers: seed after it)
    torch.manual_seed(c["seed"])                                   # prop shuffles the shots with torch.randperm (:2803)
    grad = net.prop(c["inputs"], c["vel"].as_subclass(PC.CpuPinned), 0, 1, c["mintrue"], c["maxtrue"])
    z = np.load(os.path.join(GOLD, "prop_auto.npz"))
    assert len(log) == int(z["n_calls"]) == 2                      # direct wave + the gradient-carrying call
    assert K.rel_l2(grad.as_subclass(torch.Tensor), torch.from_numpy(z["grad"])) < 1e-6
    assert K.rel_l2(log[-1]["rec"][::16], torch.from_numpy(z["pred_sub"])) < 1e-6
    # the call-site contract the drop-in has to accept: [nt, shots, 1] amplitudes, [shots, 1|R, 2] metres, python dt
    assert log[-1]["src_shape"] == (1000, 5, 1) and tuple(log[-1]["x_r"].shape) == (5, 70, 2) and log[-1]["dt"] == 0.001

    # the replay used on the GPU is the same computation
    g2, _, pred2 = PC.replay_auto(RH.oracle_deepwave(), c, "cpu")
    assert K.rel_l2(g2, torch.from_numpy(z["grad"])) < 1e-6
    assert K.rel_l2(pred2[::16], torch.from_numpy(z["pred_sub"])) < 1e-6


def test_auto22_fixture_is_committed_and_sane():
    """The AutoMarmousi22_Net.prop fixture takes minutes of CPU to regenerate (tests/golden/make_prop_golden.py auto22);
    here only its shape and the reference's post-processing are checked: water rows zeroed (:5493), z^2 weight (:5492)."""
    z = np.load(os.path.join(GOLD, "prop_auto22.npz"))
    c = PC.auto22_inputs()
    grad = torch.from_numpy(z["grad"])
    assert tuple(grad.shape) == (151, 200) and np.isfinite(z["loss"]) and float(z["loss"]) > 0
    assert float(grad[c["true"][0, 0] == 1500].abs().max()) == 0.0
    assert float(grad.abs().max()) > 0
    assert tuple(z["src_shape"]) == (4001, 9, 1) and tuple(z["x_s"].shape) == (9, 1, 2) and tuple(z["x_r"].shape) == (9, 200, 2)
    assert int(z["n_calls"]) == 2 and float(z["direct_vp"]) == 1500.0


def test_reference_fullbatch_lbfgs_drives_a_deepwave_style_function():
    LB = RH.load_lbfgs()
    dw = RH.oracle_deepwave()
    nz, ny, dx, nt, dt = 24, 40, 10.0, 120, 1e-3
    v_true = K.layered_model(nz, ny, seed=3, noise=0.0)
    x_s, x_r = K.survey(nz, ny, dx, 2, 12, 1, seed=1)
    src = K.ricker(20.0, nt, dt, 0.05).reshape(-1, 1, 1).repeat(1, 2, 1).contiguous()
    with torch.no_grad():
        obs = dw.scalar.Propagator({"vp": v_true}, dx)(src, x_s, x_r, dt)
    v = torch.nn.Parameter(torch.full((nz, ny), 2300.0))
    opt = LB.FullBatchLBFGS([v], lr=1.0, history_size=10, line_search="Wolfe", debug=False)
    calls = [0]

    def closure():            # models/AutoElMar22LBFGS_model.py:477-498: forward AND backward inside, a detached loss out
        opt.zero_grad()       # (the reference's LBFGS has its own F_new.backward() commented out, functions/LBFGS.py:816,848)
        calls[0] += 1
        loss = torch.nn.functional.mse_loss(dw.scalar.Propagator({"vp": v}, dx)(src, x_s, x_r, dt), obs)
        loss.backward()
        return loss.detach()

    loss = closure()
    first = float(loss)
    for _ in range(3):                                            # :518-523
        loss, _, _, _, _, _, _, _ = opt.step({"closure": closure, "current_loss": loss, "damping": True})
    assert float(loss) < first and calls[0] >= 4
