"""GPU parity of the elastic propagator (through the C ABI) against oracle/elastic_ref.py.

Tolerances: relative L2 <= 1e-5 on traces and <= 1e-4 on gradients against the fp32 oracle (the
north star's figures); looser sanity bounds against the fp64 oracle (fp32 round-off of the
recurrence itself)."""
import pytest
import torch

from tests import _cases as K

pytestmark = pytest.mark.gpu

CASES = [
    dict(nz=30, ny=50, dx=10.0, order=4, pml=8, S=2, R=7, Ns=1, nt=120, dt=1e-3, f=15.0),
    dict(nz=26, ny=40, dx=10.0, order=8, pml=6, S=2, R=5, Ns=2, nt=90, dt=2.5e-3, f=12.0),     # sub-stepped
    dict(nz=24, ny=33, dx=(8.0, 12.0), order=2, pml=5, S=3, R=6, Ns=1, nt=100, dt=1e-3, f=20.0),
    dict(nz=28, ny=36, dx=10.0, order=4, pml=0, S=1, R=4, Ns=1, nt=60, dt=1e-3, f=20.0),
]


def _inputs(c):
    vp = K.layered_model(c["nz"], c["ny"], seed=4)
    vs = vp / 1.732
    vs[:3] = 0.0                                   # a water layer (mu = 0), cf. models/networks.py:8857-8858
    rho = 310.0 * vp ** 0.25
    dy = c["dx"][1] if isinstance(c["dx"], tuple) else c["dx"]
    dzz = c["dx"][0] if isinstance(c["dx"], tuple) else c["dx"]
    g = torch.Generator().manual_seed(11)
    x_s = torch.zeros(c["S"], c["Ns"], 2)
    x_s[..., 0] = torch.randint(1, c["nz"] - 1, (c["S"], c["Ns"]), generator=g).float() * dzz
    x_s[..., 1] = torch.randint(1, c["ny"] - 1, (c["S"], c["Ns"]), generator=g).float() * dy
    x_r = torch.zeros(c["S"], c["R"], 2)
    x_r[..., 0] = torch.randint(0, c["nz"], (c["S"], c["R"]), generator=g).float() * dzz
    x_r[..., 1] = torch.randint(0, c["ny"], (c["S"], c["R"]), generator=g).float() * dy
    src = K.ricker(c["f"], c["nt"], c["dt"], 1.0 / c["f"]).reshape(-1, 1, 1).repeat(1, c["S"], c["Ns"]).contiguous()
    src = src * (1.0 + 0.2 * torch.arange(c["S"] * c["Ns"]).reshape(1, c["S"], c["Ns"])) * 1e6
    grad_rec = torch.randn(2, c["nt"], c["S"], c["R"], generator=g)
    return vp, vs, rho, x_s, x_r, src, grad_rec


@pytest.mark.parametrize("fused_fwd", ["0", "1"], ids=["two-kernel-forward", "fused-forward"])
@pytest.mark.parametrize("idx", range(len(CASES)))
def test_elastic_traces_and_gradients_match_oracle(cuda_lib, idx, fused_fwd, monkeypatch):
    """Default kernels (two-kernel forward step, fused adjoint) and the opt-in fused forward step."""
    from oracle import elastic_ref as E
    import physicsbasedfwi2_b200 as pkg
    monkeypatch.setenv("FWI_B200_EL_FUSED_FWD", fused_fwd)
    c = CASES[idx]
    vp, vs, rho, x_s, x_r, src, grad_rec = _inputs(c)
    dev = torch.device("cuda:0")
    leaves = [t.to(dev).requires_grad_(True) for t in (vp, vs, rho, src)]
    prop = pkg.Propagator({"vp": leaves[0], "vs": leaves[1], "rho": leaves[2]}, c["dx"], pml_width=c["pml"],
                          accuracy=c["order"])
    rec = prop(leaves[3], x_s.to(dev), x_r.to(dev), c["dt"])
    assert tuple(rec.shape) == (2, c["nt"], c["S"], c["R"])
    (rec * grad_rec.to(dev)).sum().backward()
    for dtype, ttol, gtol in ((torch.float32, 1e-5, 1e-4), (torch.float64, 5e-5, 2e-4)):
        o = E.elastic_gradient(vp.to(dtype), vs.to(dtype), rho.to(dtype), c["dx"], src.to(dtype), x_s, x_r, c["dt"],
                               grad_rec.to(dtype), order=c["order"], pml_width=c["pml"])
        assert K.rel_l2(rec, o[0]) <= ttol, (dtype, K.rel_l2(rec, o[0]))
        for name, got, want in zip(("vp", "vs", "rho", "src"), [t.grad for t in leaves], o[1:]):
            assert K.rel_l2(got, want) <= gtol, (dtype, name, K.rel_l2(got, want))


def test_elastic_rejects_bad_input(cuda_lib):
    import physicsbasedfwi2_b200 as pkg
    v = torch.full((20, 20), 2000.0, device="cuda:0")
    with pytest.raises(RuntimeError, match="vp.*vs.*rho|keys"):
        pkg.elastic.Propagator({"vp": v, "vs": v}, 10.0)
    prop = pkg.Propagator({"vp": v, "vs": v / 2, "rho": v}, 10.0)
    with pytest.raises(RuntimeError, match="outside the model"):
        prop(torch.zeros(10, 1, 1, device="cuda:0"), torch.tensor([[[500.0, 0.0]]]), torch.zeros(1, 2, 2), 1e-3)


def _golden():
    import glob
    import os
    return sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "el_*.npz")))


@pytest.mark.parametrize("path", _golden(), ids=lambda p: p.split("/")[-1][:-4])
def test_elastic_cuda_matches_committed_golden_vectors(cuda_lib, path):
    """Golden vectors from tests/golden/make_golden.py (oracle/elastic_ref.py, fp64)."""
    import numpy as np
    import physicsbasedfwi2_b200 as pkg
    z = np.load(path)
    dev = torch.device("cuda:0")
    t = lambda k: torch.from_numpy(z[k]).to(dev)   # noqa: E731
    leaves = [t(k).requires_grad_(True) for k in ("vp", "vs", "rho", "src")]
    prop = pkg.Propagator({"vp": leaves[0], "vs": leaves[1], "rho": leaves[2]}, float(z["dx"]), pml_width=int(z["pml"]),
                          accuracy=int(z["order"]))
    rec = prop(leaves[3], t("x_s"), t("x_r"), float(z["dt"]))
    (rec * t("grad_rec")).sum().backward()
    assert K.rel_l2(rec, torch.from_numpy(z["rec"])) <= 5e-5
    for name, leaf in zip(("grad_vp", "grad_vs", "grad_rho", "grad_src"), leaves):
        assert K.rel_l2(leaf.grad, torch.from_numpy(z[name])) <= 2e-4, name


@pytest.mark.parametrize("source", ["fz", "fy"])
def test_force_sources_match_oracle(cuda_lib, source):
    """Vertical / horizontal point forces (DENISE QUELLTYPB 2 / 3; the reference sets 2 at models/networks.py:10452):
    the source term depends on the buoyancy, so the density gradient gets a source-cell contribution."""
    from oracle import elastic_ref as E
    import physicsbasedfwi2_b200 as pkg
    c = CASES[0]
    vp, vs, rho, x_s, x_r, src, grad_rec = _inputs(c)
    x_s[1] = x_s[0]                                      # two shots with the same source cell
    dev = torch.device("cuda:0")
    leaves = [t.to(dev).requires_grad_(True) for t in (vp, vs, rho, src)]
    prop = pkg.Propagator({"vp": leaves[0], "vs": leaves[1], "rho": leaves[2]}, c["dx"], pml_width=c["pml"],
                          accuracy=c["order"], source=source)
    rec = prop(leaves[3], x_s.to(dev), x_r.to(dev), c["dt"])
    (rec * grad_rec.to(dev)).sum().backward()
    o = E.elastic_gradient(vp, vs, rho, c["dx"], src, x_s, x_r, c["dt"], grad_rec, order=c["order"], pml_width=c["pml"],
                           source=source)
    assert float(o[0].abs().max()) > 0
    assert K.rel_l2(rec, o[0]) <= 1e-5, K.rel_l2(rec, o[0])
    for name, got, want in zip(("vp", "vs", "rho", "src"), [t.grad for t in leaves], o[1:]):
        assert K.rel_l2(got, want) <= 1e-4, (name, K.rel_l2(got, want))


@pytest.mark.parametrize("groups", ["1", "2"])
@pytest.mark.parametrize("K_", [16, 31, 50])
def test_elastic_checkpointing_reproduces_the_stored_snapshot_gradient(cuda_lib, K_, groups, monkeypatch):
    """checkpoint_every=K keeps the 13 live fields every K steps and recomputes the five snapshot fields of one segment
    at a time in the backward pass: the same kernels on the same values."""
    import physicsbasedfwi2_b200 as pkg
    monkeypatch.setenv("FWI_B200_EL_GROUPS", groups)     # "2": the two shots as concurrent groups on their own streams
    c = CASES[1]                                         # sub-stepped (ratio 2), two sources per shot
    vp, vs, rho, x_s, x_r, src, grad_rec = _inputs(c)
    dev = torch.device("cuda:0")
    outs = []
    for ck in (0, K_):
        if ck == 0:
            monkeypatch.setenv("FWI_B200_EL_GROUPS", "1")   # reference: one call over all shots, stored snapshots
        else:
            monkeypatch.setenv("FWI_B200_EL_GROUPS", groups)
        leaves = [t.to(dev).requires_grad_(True) for t in (vp, vs, rho, src)]
        prop = pkg.Propagator({"vp": leaves[0], "vs": leaves[1], "rho": leaves[2]}, c["dx"], pml_width=c["pml"],
                              accuracy=c["order"], checkpoint_every=ck)
        rec = prop(leaves[3], x_s.to(dev), x_r.to(dev), c["dt"])
        (rec * grad_rec.to(dev)).sum().backward()
        outs.append([rec.detach()] + [t.grad for t in leaves])
    # traces and the source gradient come straight from the kernels: bitwise.  The model gradients pass through torch's
    # replicate-pad backward (atomic adds, run-to-run round-off differences) on their way from the five images to vp, vs, rho.
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][4], outs[1][4])
    for a, b in zip(outs[0][1:4], outs[1][1:4]):
        assert K.rel_l2(a, b) <= 1e-6


def test_elastic_oversize_call_raises_a_sized_error(cuda_lib):
    import physicsbasedfwi2_b200 as pkg
    v = torch.full((376, 1151), 2500.0, device="cuda:0", requires_grad=True)
    prop = pkg.Propagator({"vp": v, "vs": v.detach() / 2, "rho": v.detach()}, 10.0)
    with pytest.raises(RuntimeError, match="GiB of snapshots.*checkpoint_every"):
        prop(torch.zeros(400000, 4, 1, device="cuda:0"), torch.zeros(4, 1, 2), torch.zeros(4, 3, 2), 1e-3)


@pytest.mark.parametrize("order,source", [(4, "explosive"), (8, "fz"), (2, "explosive")])
def test_free_surface_matches_oracle(cuda_lib, order, source):
    """Stress-imaging free surface at the first model row (DENISE FREE_SURF = 1, models/networks.py:9811): traces and all
    four gradients against oracle/elastic_ref.py (autograd), receivers on the surface, a source one cell below it."""
    from oracle import elastic_ref as E
    import physicsbasedfwi2_b200 as pkg
    c = dict(CASES[0], order=order, nt=140)
    vp, vs, rho, x_s, x_r, src, grad_rec = _inputs(c)
    vs = vp / 1.732                                      # solid up to the surface
    x_s[0, 0] = torch.tensor([c["dx"], 12 * c["dx"]])    # one cell below the surface
    x_r[:, :3, 0] = 0.0                                  # three receivers on the surface
    dev = torch.device("cuda:0")
    leaves = [t.to(dev).requires_grad_(True) for t in (vp, vs, rho, src)]
    prop = pkg.Propagator({"vp": leaves[0], "vs": leaves[1], "rho": leaves[2]}, c["dx"], pml_width=c["pml"],
                          accuracy=order, source=source, free_surface=True)
    rec = prop(leaves[3], x_s.to(dev), x_r.to(dev), c["dt"])
    (rec * grad_rec.to(dev)).sum().backward()
    o = E.elastic_gradient(vp, vs, rho, c["dx"], src, x_s, x_r, c["dt"], grad_rec, order=order, pml_width=c["pml"],
                           source=source, free_surface=True)
    o_pml = E.elastic_forward(vp, vs, rho, c["dx"], src, x_s, x_r, c["dt"], order=order, pml_width=c["pml"], source=source)
    assert K.rel_l2(o[0], o_pml) > 0.1                   # the surface matters in this set-up
    assert K.rel_l2(rec, o[0]) <= 1e-5, K.rel_l2(rec, o[0])
    for name, got, want in zip(("vp", "vs", "rho", "src"), [t.grad for t in leaves], o[1:]):
        assert K.rel_l2(got, want) <= 1e-4, (name, K.rel_l2(got, want))
