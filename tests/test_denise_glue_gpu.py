"""DENISE-side glue of the reference's elastic `prop` on the GPU, against the restatement of
models/networks.py:9194-9240 (oracle/denise_glue_ref.py, pinned by those lines) and scipy's Butterworth sections."""
import numpy as np
import pytest
import torch

from tests import _cases as K

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("order,zero_phase,fc,dt", [(6, True, 10.0, 1e-3), (4, False, 10.0, 2e-3), (2, True, 25.0, 1e-3)])
def test_lowpass_matches_scipy_sections_and_its_backward_is_the_adjoint(cuda_lib, order, zero_phase, fc, dt):
    from oracle import denise_glue_ref as G
    from physicsbasedfwi2_b200.denise_glue import lowpass
    g = torch.Generator().manual_seed(1)
    x = torch.randn(700, 3, 37, generator=g)
    y = lowpass(x.to(DEV), dt, fc, order, zero_phase)
    want = torch.from_numpy(G.lowpass(x.numpy(), dt, fc, order, zero_phase).copy())
    assert K.rel_l2(y, want) < 2e-6
    # <F x, g> = <x, F^T g>
    xg = x.to(DEV).requires_grad_(True)
    gg = torch.randn(700, 3, 37, generator=g).to(DEV)
    (lowpass(xg, dt, fc, order, zero_phase) * gg).sum().backward()
    lhs = float((y.double() * gg.double()).sum())
    rhs = float((x.to(DEV).double() * xg.grad.double()).sum())
    assert abs(lhs - rhs) <= 1e-5 * max(abs(lhs), 1e-12), (lhs, rhs)


def test_gradient_post_matches_the_reference_lines(cuda_lib):
    from oracle import denise_glue_ref as G
    from physicsbasedfwi2_b200.denise_glue import gradient_post
    nz, ny = 170, 396                                               # the AutoElFullRho grid (models/networks.py:8661)
    g = torch.Generator().manual_seed(2)
    grads = [torch.randn(nz, ny, generator=g) * s for s in (1e-9, 3e-10, 2e-12)]
    vp = K.layered_model(nz, ny, seed=3, vmin=1500.0, vmax=4700.0)
    vs, rho = vp / 1.73, 310.0 * vp ** 0.25
    out = gradient_post(*[t.to(DEV) for t in grads], vp.to(DEV), vs.to(DEV), rho.to(DEV))
    want = G.gradient_post(*[t.numpy() for t in grads], vp.numpy(), vs.numpy(), rho.numpy())
    for a, b, name in zip(out, want, ("vp", "vs", "rho")):
        assert K.rel_l2(a, torch.from_numpy(np.asarray(b))) < 1e-6, name
        assert float(a[:24].abs().max()) <= 1e-10 + 1e-12
    assert abs(float(out[0].abs().max()) - float(vp.max())) < 1e-3 * float(vp.max())      # scaled to max|model|


def test_elastic_inner_gradient_against_the_torch_oracle(cuda_lib):
    """The whole chain at a size torch autograd handles: elastic forward, zero-phase low-pass of synthetic and observed
    seismograms, 0.5 |.|^2, adjoint - the raw gradients against oracle/elastic_ref.py with the filter as a dense matrix -
    then the post-processing against the reference lines."""
    from oracle import denise_glue_ref as G
    from oracle import elastic_ref as E
    from physicsbasedfwi2_b200.denise_glue import elastic_inner_gradient
    nz, ny, dx, dt, nt, S, R = 34, 60, 20.0, 2e-3, 160, 2, 9
    unit = 100.0
    vp = K.layered_model(nz, ny, seed=5, vmin=1500.0, vmax=3800.0) / unit
    vs = vp / 1.8
    vs[:4] = 0.0
    rho = 3.1 * (vp * unit) ** 0.25 / 1.0
    x_s = torch.tensor([[[40.0, 300.0]], [[40.0, 820.0]]])
    x_r = torch.zeros(S, R, 2)
    x_r[:, :, 0] = 460.0
    x_r[:, :, 1] = (200.0 + 100.0 * torch.arange(R))[None]
    obs = 1e-9 * torch.randn(2, nt, S, R, generator=torch.Generator().manual_seed(7))
    kw = dict(dx=dx, dt=dt, fsource=12.0, fc_high=20.0, filter_order=4, unit=unit, source_scale=1e6, pml_width=8)
    gvp, gvs, grho, loss = elastic_inner_gradient(vp.to(DEV), vs.to(DEV), rho.to(DEV), obs.to(DEV), x_s, x_r,
                                                  postprocess=False, **kw)
    # oracle: same chain in float64 torch on the CPU
    Fm = torch.from_numpy(G.lowpass_matrix(nt, dt, 20.0, 4, True).copy())
    leaves = [(t.double() * unit).requires_grad_(True) for t in (vp, vs, rho)]
    src = (K.ricker(12.0, nt, dt, 1 / 12.0).double() * 1e6).reshape(-1, 1, 1).repeat(1, S, 1)
    rec = E.elastic_forward(leaves[0], leaves[1], leaves[2], dx, src, x_s, x_r, dt, order=4, pml_width=8)
    res = torch.einsum("ts,csab->ctab", Fm, rec - obs.double())
    l0 = 0.5 * (res ** 2).sum()
    l0.backward()
    assert abs(float(loss) - float(l0)) <= 2e-4 * float(l0)
    for got, leaf, name in zip((gvp, gvs, grho), leaves, ("vp", "vs", "rho")):
        assert K.rel_l2(got, leaf.grad) <= 5e-4, (name, K.rel_l2(got, leaf.grad))
    # and with the post-processing
    out = elastic_inner_gradient(vp.to(DEV), vs.to(DEV), rho.to(DEV), obs.to(DEV), x_s, x_r, water_rows=4, **kw)
    want = G.gradient_post(gvp.cpu().numpy(), gvs.cpu().numpy(), grho.cpu().numpy(), (vp * unit).numpy(), (vs * unit).numpy(),
                           (rho * unit).numpy(), water_rows=4)
    for a, b in zip(out[:3], want):
        assert K.rel_l2(a, torch.from_numpy(np.asarray(b))) < 1e-5
