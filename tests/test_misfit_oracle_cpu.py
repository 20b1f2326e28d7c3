"""The misfit oracle (oracle/misfit_ref.py) restates reference lines models/networks.py:5467-5476;
here: hand-computed values on a tiny case and a finite-difference check of its gradient."""
import torch

from oracle import misfit_ref as M


def test_hand_computed_l1_and_mse():
    pred = torch.tensor([[[2.0]], [[-4.0]], [[1.0]]])            # one trace, max |.| = 4
    direct = torch.tensor([[[0.0]], [[0.0]], [[1.0]]])           # d = 2, -4, 0
    obs = torch.tensor([[[0.0]], [[-1.0]], [[0.5]]])
    l1 = M.trace_misfit(pred, obs, direct, kind="l1")
    assert abs(float(l1) - (0.5 + 0.0 + 0.5) / 3) < 1e-6         # n = 0.5, -1, 0
    mse = M.trace_misfit(pred, obs, direct, kind="mse")
    assert abs(float(mse) - (0.25 + 0.0 + 0.25) / 3) < 1e-6
    raw = M.trace_misfit(pred, obs, None, kind="mse", normalize=False)
    assert abs(float(raw) - (4.0 + 9.0 + 0.25) / 3) < 1e-6


def test_gradient_matches_finite_differences():
    g = torch.Generator().manual_seed(0)
    pred = torch.randn(20, 2, 3, generator=g, dtype=torch.float64)
    direct = torch.randn(20, 2, 3, generator=g, dtype=torch.float64) * 0.2
    obs = torch.randn(20, 2, 3, generator=g, dtype=torch.float64)
    l0, gr = M.trace_misfit_grad(pred, obs, direct, kind="mse")
    dp = torch.randn(20, 2, 3, generator=g, dtype=torch.float64)
    h = 1e-6
    lp = M.trace_misfit(pred + h * dp, obs, direct, kind="mse")
    lm = M.trace_misfit(pred - h * dp, obs, direct, kind="mse")
    assert abs(float((lp - lm) / (2 * h)) - float((gr * dp).sum())) < 1e-7
