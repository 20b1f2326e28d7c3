"""Properties of the elastic CPU oracle (oracle/elastic_ref.py): it is the checker of the CUDA
elastic path, so it is itself pinned by physics it must satisfy (the reference has no fixture for
this path: PARITY UNPINNED, see the oracle header)."""
import math

import torch

from oracle import elastic_ref as E
from tests import _cases as K


def _setup(nz=24, ny=30, nt=60, order=4, pml=6):
    vp = K.layered_model(nz, ny, seed=3).double()
    vs = vp / 1.8
    rho = 310.0 * vp ** 0.25
    x_s = torch.tensor([[[80.0, 140.0]]])
    x_r = torch.zeros(1, 6, 2)
    x_r[0, :, 0] = 30.0
    x_r[0, :, 1] = torch.arange(6) * 40.0 + 30.0
    src = K.ricker(20.0, nt, 1e-3, 1 / 20.0).reshape(-1, 1, 1).double() * 1e6
    return vp, vs, rho, x_s, x_r, src, dict(order=order, pml_width=pml)


def test_gradient_taylor_slopes():
    """|J(m + h dm) - J(m) - h <grad, dm>| must fall as h^2 (the test the reference's unused Devito
    tree applies to its acoustic gradient: seisgan/fwi/pde/seismic/acoustic/gradient_example.py:115-146)."""
    vp, vs, rho, x_s, x_r, src, kw = _setup()
    obs = E.elastic_forward(vp * 1.03, vs * 1.02, rho, 10.0, src, x_s, x_r, 1e-3, **kw)

    def J(vp_, vs_, rho_):
        r = E.elastic_forward(vp_, vs_, rho_, 10.0, src, x_s, x_r, 1e-3, **kw)
        return 0.5 * ((r - obs) ** 2).sum()
    leaves = [t.clone().requires_grad_(True) for t in (vp, vs, rho)]
    J0 = J(*leaves)
    g = torch.autograd.grad(J0, leaves)
    gen = torch.Generator().manual_seed(0)
    dm = [torch.randn(t.shape, generator=gen, dtype=torch.float64) * s for t, s in zip((vp, vs, rho), (30.0, 20.0, 10.0))]
    lin = sum((a * b).sum() for a, b in zip(g, dm))
    errs = []
    hs = [1.0, 0.5, 0.25, 0.125]
    for h in hs:
        Jh = J(vp + h * dm[0], vs + h * dm[1], rho + h * dm[2])
        errs.append(abs(float(Jh - J0 - h * lin)))
    slopes = [math.log(errs[i] / errs[i + 1]) / math.log(2.0) for i in range(len(hs) - 1)]
    assert all(abs(s - 2.0) < 0.2 for s in slopes), (errs, slopes)


def test_cpml_absorbs():
    """With the C-PML the late-time wavefield energy at the receivers is far below the no-PML run
    (which reflects from the rigid edge of the padded grid)."""
    vp, vs, rho, x_s, x_r, src, kw = _setup(nt=420, pml=10)
    vp = torch.full_like(vp, 2500.0); vs = vp / 1.8; rho = torch.full_like(vp, 2200.0)
    with torch.no_grad():
        r_pml = E.elastic_forward(vp, vs, rho, 10.0, src, x_s, x_r, 1e-3, order=4, pml_width=10)
        r_ref = E.elastic_forward(vp, vs, rho, 10.0, src, x_s, x_r, 1e-3, order=4, pml_width=0)
    late = slice(300, 420)
    assert float(r_pml[:, late].abs().max()) < 0.1 * float(r_ref[:, late].abs().max())
    assert float(r_pml[:, :150].abs().max()) > 0


def test_water_layer_has_no_shear_stress_parameters():
    vp = torch.full((10, 12), 1500.0, dtype=torch.float64)
    vs = torch.zeros_like(vp); vs[5:] = 800.0
    rho = torch.full_like(vp, 1000.0)
    l2m, lam, mu, by, bz = E.effective_parameters(vp, vs, rho)
    assert float(mu[:5].abs().max()) == 0.0 and float(mu[4].abs().max()) == 0.0   # any fluid corner -> 0
    assert torch.allclose(lam[:5], l2m[:5]) and float(mu[6:9].min()) > 0
    assert torch.allclose(by, torch.full_like(by, 1e-3)) and torch.allclose(bz, by)


def test_stag_coefficients_are_consistent():
    for order, c in E.STAG.items():
        # exact for linear functions: sum c_k (2k - 1) = 1
        assert abs(sum(ck * (2 * k - 1) for k, ck in enumerate(c, 1)) - 1.0) < 1e-12


def test_oracle_reproduces_committed_golden_vectors():
    import glob
    import os
    import numpy as np
    paths = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "el_*.npz")))
    assert len(paths) >= 2
    for path in paths:
        z = np.load(path)
        t = lambda k: torch.from_numpy(z[k]).double()   # noqa: E731
        out = E.elastic_gradient(t("vp"), t("vs"), t("rho"), float(z["dx"]), t("src"), torch.from_numpy(z["x_s"]),
                                 torch.from_numpy(z["x_r"]), float(z["dt"]), t("grad_rec"), order=int(z["order"]),
                                 pml_width=int(z["pml"]))
        for got, key in zip(out, ("rec", "grad_vp", "grad_vs", "grad_rho", "grad_src")):
            assert K.rel_l2(got, torch.from_numpy(z[key])) < 1e-5, (path, key)   # inputs were rounded to fp32


def test_explosion_in_a_homogeneous_solid_is_a_pure_p_wave():
    """Analytic pin of the elastic scheme: an explosive source (sxx = szz) in a homogeneous solid
    radiates only a P wave.  It must travel at vp (not vs), decay as 1/sqrt(r) (2-D spreading), be
    radially symmetric, and leave nothing in the S-arrival window."""
    nz, ny, nt, dx, dt = 80, 120, 440, 10.0, 1e-3
    vp = torch.full((nz, ny), 3000.0, dtype=torch.float64)
    vs, rho = torch.full_like(vp, 1700.0), torch.full_like(vp, 2200.0)
    x_s = torch.tensor([[[400.0, 300.0]]])
    # vy lives at (i, j + 1/2): cells 50 / 80 / 9 of row 40 are 205 m, 505 m and -205 m from the source in y;
    # vz lives at (i + 1/2, j): cells 60 / 19 of column 30 are +205 m / -205 m from it in z
    cells = [(40, 50), (40, 80), (40, 9), (60, 30), (19, 30)]
    x_r = torch.tensor([[[z * dx, y * dx] for z, y in cells]])
    src = K.ricker(15.0, nt, dt, 1 / 15.0).reshape(-1, 1, 1).double() * 1e6
    with torch.no_grad():
        r = E.elastic_forward(vp, vs, rho, dx, src, x_s, x_r, dt, order=8, pml_width=10)
    vy, vz = r[0, :, 0], r[1, :, 0]
    near, far = vy[:, 0], vy[:, 1]
    lags = list(range(60, 200))
    cc = [float((near[:nt - l] * far[l:]).sum()) for l in lags]
    lag = lags[max(range(len(cc)), key=lambda i: cc[i])]
    assert abs(lag - 300.0 / 3000.0 / dt) <= 1, lag                       # (505 - 205) m at vp: 100 samples (176 at vs)
    ratio = float(near.abs().max() / far.abs().max()) ** 2
    assert abs(ratio - 505.0 / 205.0) < 0.1 * 505.0 / 205.0, ratio        # cylindrical spreading
    peak = float(near.abs().max())
    assert float((vy[:, 0] + vy[:, 2]).abs().max()) < 1e-3 * peak         # left / right mirror, opposite sign
    assert float((vz[:, 3] + vz[:, 4]).abs().max()) < 1e-3 * peak         # below / above
    assert float((vz[:, 3] - vy[:, 0]).abs().max()) < 1e-3 * peak         # same radial motion along z as along y
    s_arrival = int((505.0 / 1700.0 + 1 / 15.0) / dt)                     # where an S wave would peak at the far receiver
    assert float(far[s_arrival - 35:s_arrival + 65].abs().max()) < 0.02 * float(far.abs().max())
