"""Size-independent properties at the full grid sizes of BASELINE.json (the oracle is too slow there):

* C2 acoustic 751 x 2301, order 8, PML 20: source-receiver reciprocity (the scheme is symmetric, the
  oracle shows it to 4e-15 in fp64), the adjoint identity <L s, g> = <s, L^T g> between the forward
  and the adjoint kernel, linearity, and stored-snapshot against checkpointed model gradients.
* C3 elastic 376 x 1151, order 4, C-PML 10: the adjoint identity and linearity.

The time axis is shortened (the properties hold step by step); grids, orders, PML widths, spacings and
time steps are the full configurations, so the work table, the TMA tiling and the PML frame are the
ones the bench runs."""
import pytest
import torch

from tests import _cases as K

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _dot(a, b):
    return float((a.detach().double().cpu() * b.detach().double().cpu()).sum())


def _c2(nt=500):
    nz, ny, dx, dt = 751, 2301, 4.0, 4e-4
    v = K.layered_model(nz, ny, seed=5, vmin=1500.0, vmax=4500.0, noise=30.0).to(DEV)
    src = K.ricker(15.0, nt, dt, 1 / 15.0)
    return nz, ny, nt, dx, dt, v, src


def test_c2_reciprocity_and_linearity(cuda_lib):
    from physicsbasedfwi2_b200.scalar import Propagator
    nz, ny, nt, dx, dt, v, w = _c2(nt=1000)
    # shot 0: a -> b, shot 1: b -> a (inside the model, either side of the first interface, 340 m apart)
    a, b = (230 * dx, 900 * dx), (290 * dx, 960 * dx)
    x_s = torch.tensor([[a], [b]], device=DEV)
    x_r = torch.tensor([[b], [a]], device=DEV)
    src = w.reshape(-1, 1, 1).repeat(1, 2, 1).contiguous().to(DEV)
    prop = Propagator({"vp": v}, dx, pml_width=20, accuracy=8, vpmax=4600.0)
    rec = prop(src, x_s, x_r, dt)
    assert float(rec[:, 0].abs().max()) > 1e-3 * float(src.abs().max()) * float(v.min()) ** 2 * dt ** 2    # the wave has arrived
    assert K.rel_l2(rec[:, 0], rec[:, 1]) <= 1e-4, K.rel_l2(rec[:, 0], rec[:, 1])
    # Linearity.  Not bit-exact even for a power of two: ahead of the wavefront the field passes through the
    # fp32 denormal range, where scaling by 4 does not commute with rounding, and a leapfrog recurrence carries
    # that on (a plain torch fp32 leapfrog loop shows 2e-5 after 1500 steps; with 2^60-scaled sources it is exact).
    rec4 = prop(4.0 * src, x_s, x_r, dt)
    assert K.rel_l2(rec4, 4.0 * rec) <= 3e-5
    assert torch.equal(prop(src, x_s, x_r, dt), rec)                      # run-to-run bit-identical


def test_c2_adjoint_identity_and_checkpointed_gradient(cuda_lib):
    from physicsbasedfwi2_b200.scalar import Propagator
    nz, ny, nt, dx, dt, v, w = _c2()
    S, R = 3, 96
    x_s, x_r = K.survey(nz, ny, dx, S, R, 1, z_src=2, z_rec=2, seed=3)
    gen = torch.Generator().manual_seed(9)
    src = (w.reshape(-1, 1, 1) * (1.0 + 0.3 * torch.rand(1, S, 1, generator=gen))).contiguous()
    x_s, x_r = x_s.to(DEV), x_r.to(DEV)

    def gradient(**kw):
        vg = v.clone().requires_grad_(True)
        sg = src.to(DEV).requires_grad_(True)
        rec = Propagator({"vp": vg}, dx, pml_width=20, accuracy=8, vpmax=4600.0, **kw)(sg, x_s, x_r, dt)
        g = rec.detach() * wgt
        (rec * g).sum().backward()
        return rec.detach(), g, vg.grad, sg.grad

    wgt = (0.5 + torch.rand(nt, S, R, generator=gen)).to(DEV)
    rec, g, gv, gs = gradient()
    lhs, rhs = _dot(rec, g), _dot(src, gs)                 # <L s, g> and <s, L^T g>
    assert lhs > 0 and abs(lhs - rhs) <= 1e-4 * lhs, (lhs, rhs)
    rec_c, _, gv_c, gs_c = gradient(checkpoint_every=50)
    assert K.rel_l2(rec_c, rec) <= 1e-6
    assert K.rel_l2(gv_c, gv) <= 2e-5, K.rel_l2(gv_c, gv)
    assert K.rel_l2(gs_c, gs) <= 2e-5, K.rel_l2(gs_c, gs)
    assert float(gv.abs().max()) > 0 and bool(torch.isfinite(gv).all())


def test_c3_elastic_adjoint_identity_and_linearity(cuda_lib):
    import physicsbasedfwi2_b200 as pkg
    nz, ny, nt, dx, dt = 376, 1151, 300, 10.0, 1e-3
    vp = K.layered_model(nz, ny, seed=6, vmin=1500.0, vmax=4000.0, noise=30.0)
    vs = vp / 1.732
    vs[:24] = 0.0                                          # water layer
    rho = 310.0 * vp ** 0.25
    S, R = 2, 74
    x_s, x_r = K.survey(nz, ny, dx, S, R, 1, z_src=4, z_rec=30, seed=4)
    src = (K.ricker(8.0, nt, dt, 1 / 8.0).reshape(-1, 1, 1).repeat(1, S, 1) * 1e6).contiguous()
    x_s, x_r = x_s.to(DEV), x_r.to(DEV)
    leaves = [t.to(DEV).requires_grad_(True) for t in (vp, vs, rho, src)]
    prop = pkg.Propagator({"vp": leaves[0], "vs": leaves[1], "rho": leaves[2]}, dx, pml_width=10, accuracy=4)
    rec = prop(leaves[3], x_s, x_r, dt)
    assert tuple(rec.shape) == (2, nt, S, R)
    g = rec.detach() * (0.5 + torch.rand(rec.shape, generator=torch.Generator().manual_seed(2))).to(DEV)
    (rec * g).sum().backward()
    lhs, rhs = _dot(rec, g), _dot(src, leaves[3].grad)
    assert lhs > 0 and abs(lhs - rhs) <= 1e-4 * lhs, (lhs, rhs)
    for t in leaves[:3]:
        assert bool(torch.isfinite(t.grad).all()) and float(t.grad.abs().max()) > 0
    with torch.no_grad():
        rec2 = pkg.Propagator({"vp": vp.to(DEV), "vs": vs.to(DEV), "rho": rho.to(DEV)}, dx, pml_width=10,
                              accuracy=4)(2.0 * src.to(DEV), x_s, x_r, dt)
    assert K.rel_l2(rec2, 2.0 * rec) <= 1e-6
