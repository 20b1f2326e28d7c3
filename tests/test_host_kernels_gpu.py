"""The device-side host preparation (csrc/host_kernels.cu, through the C ABI) against the torch statements of the same
steps in physicsbasedfwi2_b200/_host.py: padded + scaled model and PML profiles, metres -> cells -> sorted injection
lists, and the fold of the image back to a model gradient."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

from tests import _cases as K

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _setup(nz, ny, order, pml, dz, dy, dt, vmax):
    from physicsbasedfwi2_b200 import _lib
    h = order // 2
    P = pml + h
    nzp, nyp = nz + 2 * P, ny + 2 * P
    lay = _lib.AcousticLayout()
    d = _lib.AcousticDesc(nzp, nyp, order, pml, 1, 0, 0, 0, 0, 0, 0, 0, dz, dy, dt)
    _lib.check(_lib.lib().fwi_acoustic_get_layout(C.byref(d), C.byref(lay)), "layout")
    m = _lib.ModelDesc(nz, ny, order, pml, 0, 0, dz, dy, dt, vmax)
    return m, lay, P, nzp, nyp


@pytest.mark.parametrize("order,pml,dz,dy", [(4, 10, 10.0, 10.0), (8, 20, 4.0, 4.0), (2, 7, 12.0, 8.0), (6, 0, 10.0, 10.0)])
def test_prepare_model_matches_the_torch_statement(cuda_lib, order, pml, dz, dy):
    import numpy as np
    from physicsbasedfwi2_b200 import _host as H, _lib
    nz, ny, dt, vmax = 37, 141, 7.3e-4, 3911.0
    v = K.layered_model(nz, ny, seed=2).to(DEV)
    m, lay, P, nzp, nyp = _setup(nz, ny, order, pml, dz, dy, dt, vmax)
    w = torch.full((nzp, lay.pitch), -1.0, device=DEV)
    sigz = torch.empty(nzp, device=DEV)
    sigy = torch.empty(lay.pitch, device=DEV)
    _lib.check(cuda_lib.fwi_acoustic_prepare_model(C.byref(m), _p(v), _p(w), _p(sigz), _p(sigy), None), "prepare")
    vp = F.pad(v[None, None], (P,) * 4, mode="replicate")[0, 0]
    dt32 = np.float32(dt)
    want = torch.zeros(nzp, lay.pitch, device=DEV)
    want[:, lay.col0:lay.col0 + nyp] = (vp * vp) * float(np.float32(dt32 * dt32))
    assert torch.equal(w, want)
    assert torch.equal(sigz.cpu(), H.sigma_profile(nzp, order // 2, pml, dz, vmax).float())
    sy = torch.zeros(lay.pitch)
    sy[lay.col0:lay.col0 + nyp] = H.sigma_profile(nyp, order // 2, pml, dy, vmax).float()
    assert torch.equal(sigy.cpu(), sy)


@pytest.mark.parametrize("mode", ["trunc", "round"])
def test_plan_points_matches_the_torch_lists(cuda_lib, mode):
    from physicsbasedfwi2_b200 import _host as H, _lib
    nz, ny, order, pml, dz, dy = 41, 77, 4, 10, 10.0, 12.5
    m, lay, P, nzp, nyp = _setup(nz, ny, order, pml, dz, dy, 1e-3, 3000.0)
    S, n = 5, 23
    g = torch.Generator().manual_seed(4)
    x = torch.stack([torch.rand(S, n, generator=g) * (nz - 1) * dz, torch.rand(S, n, generator=g) * (ny - 1) * dy], -1)
    x[1, 3] = x[1, 7]                                  # duplicates keep their order
    x[2, :4, :] = torch.tensor([0.0, 0.0])
    x[3, 5] = torch.tensor([(nz - 1) * dz, (ny - 1) * dy])
    x = x.to(DEV)
    cell = torch.empty(S, n, dtype=torch.int32, device=DEV)
    rs = torch.empty(S, nzp + 1, dtype=torch.int32, device=DEV)
    sc = torch.empty(S * n, dtype=torch.int32, device=DEV)
    so = torch.empty(S * n, dtype=torch.int32, device=DEV)
    bad = torch.zeros(1, dtype=torch.int32, device=DEV)
    _lib.check(cuda_lib.fwi_acoustic_plan_points(C.byref(m), _p(x), S, n, 1 if mode == "round" else 0, _p(cell), _p(rs),
                                                 _p(sc), _p(so), _p(bad), None), "plan")
    c = H.to_cells(x, dz, dy, mode)
    want_cell = (c[..., 0] + P) * lay.pitch + (c[..., 1] + P + lay.col0)
    assert torch.equal(cell.long(), want_cell)
    start, wc, wo = H.injection_lists(want_cell, nzp, lay.pitch)
    assert torch.equal(rs.reshape(-1), start) and torch.equal(sc, wc) and torch.equal(so, wo)
    assert int(bad) == 0
    x[0, 0, 1] = ny * dy + 1.0                          # outside: flagged, clamped
    _lib.check(cuda_lib.fwi_acoustic_plan_points(C.byref(m), _p(x), S, n, 0, _p(cell), _p(rs), _p(sc), _p(so), _p(bad), None), "plan")
    assert int(bad) == 1 and int(cell.max()) < nzp * lay.pitch


def test_fold_gradient_matches_the_torch_statement(cuda_lib):
    from physicsbasedfwi2_b200 import _host as H, _lib
    nz, ny, order, pml, dz, dy = 29, 53, 8, 6, 10.0, 10.0
    m, lay, P, nzp, nyp = _setup(nz, ny, order, pml, dz, dy, 1e-3, 3000.0)
    S, Ns, nsteps = 3, 2, 37
    g = torch.Generator().manual_seed(6)
    v = K.layered_model(nz, ny, seed=3).to(DEV)
    image = torch.randn(S, nzp, lay.pitch, generator=g).to(DEV)
    gs = torch.randn(nsteps, S, Ns, generator=g).to(DEV)
    amp = torch.randn(nsteps, S, Ns, generator=g).to(DEV)
    zc = torch.randint(0, nz, (S, Ns), generator=g)
    yc = torch.randint(0, ny, (S, Ns), generator=g)
    zc[1], yc[1] = zc[0], yc[0]                          # the same source cells in two shots, and twice within shot 2
    zc[2, 1], yc[2, 1] = zc[2, 0], yc[2, 0]
    cell = ((zc + P) * lay.pitch + yc + P + lay.col0).to(torch.int32).to(DEV)
    ws = torch.empty(int(cuda_lib.fwi_acoustic_fold_workspace_floats(C.byref(m), S, Ns)), device=DEV)
    out = torch.empty(nz, ny, device=DEV)
    _lib.check(cuda_lib.fwi_acoustic_fold_gradient(C.byref(m), S, _p(image), _p(gs), _p(amp), _p(cell), Ns, nsteps, _p(v), _p(ws),
                                                   _p(out), None), "fold")
    G = image.double().sum(0)[:, lay.col0:lay.col0 + nyp].contiguous()
    pc = ((zc + P) * nyp + yc + P).reshape(-1).to(DEV)
    G.view(-1).index_add_(0, pc, (gs.double() * amp.double()).sum(0).reshape(-1))
    vp = F.pad(v[None, None], (P,) * 4, mode="replicate")[0, 0].double()
    want = H.replicate_pad_adjoint((2.0 / vp) * G, P)
    assert K.rel_l2(out, want) < 5e-7
