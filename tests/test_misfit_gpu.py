"""Fused trace misfit (csrc/misfit.cu through the C ABI) vs the torch restatement of the reference
lines models/networks.py:5467-5476 (oracle/misfit_ref.py)."""
import pytest
import torch

from tests import _cases as K

pytestmark = pytest.mark.gpu


def _data(nt, S, R, seed=0, dead=False):
    g = torch.Generator().manual_seed(seed)
    pred = torch.randn(nt, S, R, generator=g) * torch.rand(1, S, R, generator=g) * 3.0
    direct = 0.3 * torch.randn(nt, S, R, generator=g)
    obs = torch.randn(nt, S, R, generator=g).clamp(-1, 1)
    if dead:
        pred[:, 0, 0] = 0.0
        direct[:, 0, 0] = 0.0       # a dead trace: max = 0, normalised to 0 / 1e-10 = 0
    return pred, direct, obs


@pytest.mark.parametrize("kind", ["l1", "mse"])
@pytest.mark.parametrize("normalize", [True, False])
@pytest.mark.parametrize("shape", [(200, 3, 17), (1000, 6, 101), (37, 1, 1), (3000, 2, 576)])
def test_loss_and_gradient_match_reference_expression(cuda_lib, kind, normalize, shape):
    from oracle import misfit_ref as M
    from physicsbasedfwi2_b200.misfit import trace_misfit
    pred, direct, obs = _data(*shape, seed=1, dead=True)
    l0, g0 = M.trace_misfit_grad(pred.double(), obs.double(), direct.double(), kind=kind, normalize=normalize)
    dev = torch.device("cuda:0")
    p = pred.to(dev).requires_grad_(True)
    loss = trace_misfit(p, obs.to(dev), direct.to(dev), kind=kind, normalize=normalize)
    (2.5 * loss).backward()
    assert abs(float(loss) - float(l0)) <= 2e-6 * abs(float(l0))
    assert K.rel_l2(p.grad, 2.5 * g0) <= 1e-5
    if kind == "l1" and normalize:
        l1, g1 = M.trace_misfit_grad(pred, obs, direct, kind=kind, normalize=normalize)   # fp32 torch, same ops
        assert K.rel_l2(p.grad, 2.5 * g1) <= 1e-5


def test_without_direct_wave_and_deterministic(cuda_lib):
    from oracle import misfit_ref as M
    from physicsbasedfwi2_b200.misfit import trace_misfit, normalize_traces
    pred, _, obs = _data(500, 4, 33, seed=2)
    dev = torch.device("cuda:0")
    outs = []
    for _ in range(2):
        p = pred.to(dev).requires_grad_(True)
        loss = trace_misfit(p, normalize_traces(obs.to(dev)), kind="mse")
        loss.backward()
        outs.append((loss.detach().clone(), p.grad.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    l0, g0 = M.trace_misfit_grad(pred.double(), (obs / (obs.abs().amax(0, keepdim=True) + 1e-10)).double(), None, kind="mse")
    assert abs(float(outs[0][0]) - float(l0)) <= 2e-6 * abs(float(l0)) and K.rel_l2(outs[0][1], g0) <= 1e-5


def test_rejects_cpu_tensors_and_bad_shapes(cuda_lib):
    from physicsbasedfwi2_b200.misfit import trace_misfit
    with pytest.raises(RuntimeError):
        trace_misfit(torch.zeros(4, 1, 1), torch.zeros(4, 1, 1))
    dev = torch.device("cuda:0")
    with pytest.raises(RuntimeError):
        trace_misfit(torch.zeros(4, 1, 2, device=dev), torch.zeros(4, 1, 1, device=dev))
    with pytest.raises(RuntimeError):
        trace_misfit(torch.zeros(4, 1, 1, device=dev), torch.zeros(4, 1, 1, device=dev), kind="huber")


def test_direct_wave_is_computed_once_per_geometry(cuda_lib):
    import physicsbasedfwi2_b200 as pkg
    from physicsbasedfwi2_b200.misfit import DirectWave
    from physicsbasedfwi2_b200.scalar import Propagator
    dev = torch.device("cuda:0")
    x_s, x_r = K.survey(30, 60, 10.0, 2, 7, 1, seed=1)
    src = K.ricker(20.0, 50, 1e-3, 0.05).reshape(-1, 1, 1).repeat(1, 2, 1).contiguous().to(dev)
    dw = DirectWave()
    a = dw((30, 60), 10.0, src, x_s, x_r, 1e-3)
    n0 = pkg.kernel_launches()
    b = dw((30, 60), 10.0, src, x_s, x_r, 1e-3)
    assert pkg.kernel_launches() == n0 and b is a                          # cached: no kernel ran
    ref = Propagator({"vp": torch.full((30, 60), 1500.0, device=dev)}, 10.0)(src, x_s.to(dev), x_r.to(dev), 1e-3)
    assert torch.equal(a, ref)
    c = dw((30, 60), 10.0, src, x_s + 10.0, x_r, 1e-3)                     # moved sources: recomputed
    assert pkg.kernel_launches() > n0 and not torch.equal(c, a)
