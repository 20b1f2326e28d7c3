"""Edge cases of the deepwave-style call on the CUDA path: ragged / degenerate inputs, survey
cropping, size-independent properties at full size (linearity in the source, shot-permutation
invariance), and the error behaviour the reference's callers rely on (RuntimeError)."""
import pytest
import torch

from tests import _cases as K

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _setup(nz=30, ny=80, S=2, R=9, nt=60, seed=1):
    v = K.layered_model(nz, ny, seed=seed)
    x_s, x_r = K.survey(nz, ny, 10.0, S, R, 1, seed=seed + 1)
    src = K.ricker(20.0, nt, 1e-3, 0.05).reshape(-1, 1, 1).repeat(1, S, 1).contiguous()
    return v, x_s, x_r, src


def test_single_time_sample_and_single_receiver(cuda_lib):
    from oracle import cpu_ref as O
    from physicsbasedfwi2_b200.scalar import Propagator
    v, x_s, x_r, src = _setup(S=1, R=1, nt=1)
    src = src + 1.0
    rec = Propagator({"vp": v.to(DEV)}, 10.0)(src.to(DEV), x_s.to(DEV), x_r.to(DEV), 1e-3)
    assert tuple(rec.shape) == (1, 1, 1)
    r0 = O.acoustic_forward(v, 10.0, src, x_s, x_r, 1e-3, dtype=torch.float32)
    assert torch.allclose(rec.cpu(), r0, atol=1e-12)


def test_no_receivers_gives_empty_traces(cuda_lib):
    from physicsbasedfwi2_b200.scalar import Propagator
    v, x_s, x_r, src = _setup()
    rec = Propagator({"vp": v.to(DEV)}, 10.0)(src.to(DEV), x_s.to(DEV), x_r[:, :0].to(DEV), 1e-3)
    assert tuple(rec.shape) == (src.shape[0], 2, 0)


def test_gradient_wrt_source_only_and_model_only(cuda_lib):
    """needs_input_grad combinations: the wavelet-only gradient (models/networks.py:13235-13237)
    must not need snapshots, the model-only gradient must not return a source gradient."""
    from oracle import cpu_ref as O
    from physicsbasedfwi2_b200.scalar import Propagator
    v, x_s, x_r, src = _setup()
    g = torch.randn(src.shape[0], 2, 9, generator=torch.Generator().manual_seed(3))
    _, gv0, gs0 = O.acoustic_gradient(v, 10.0, src, x_s, x_r, 1e-3, grad_rec=g, dtype=torch.float32)
    sg = src.to(DEV).requires_grad_(True)
    rec = Propagator({"vp": v.to(DEV)}, 10.0)(sg, x_s.to(DEV), x_r.to(DEV), 1e-3)
    (rec * g.to(DEV)).sum().backward()
    assert K.rel_l2(sg.grad, gs0) <= 1e-4
    vg = v.to(DEV).requires_grad_(True)
    rec = Propagator({"vp": vg}, 10.0)(src.to(DEV), x_s.to(DEV), x_r.to(DEV), 1e-3)
    (rec * g.to(DEV)).sum().backward()
    assert K.rel_l2(vg.grad, gv0) <= 1e-4


def test_survey_pad_crops_like_a_smaller_model(cuda_lib):
    from physicsbasedfwi2_b200.scalar import Propagator
    nz, ny = 40, 200
    v = K.layered_model(nz, ny, seed=4)
    x_s = torch.tensor([[[0.0, 900.0]]])
    x_r = torch.zeros(1, 11, 2)
    x_r[0, :, 1] = 800.0 + 20.0 * torch.arange(11)
    src = K.ricker(20.0, 80, 1e-3, 0.05).reshape(-1, 1, 1)
    pad = 12
    rec = Propagator({"vp": v.to(DEV)}, 10.0, survey_pad=pad)(src.to(DEV), x_s.to(DEV), x_r.to(DEV), 1e-3)
    y0, y1, z1 = 80 - pad, 100 + pad + 1, pad + 1        # the box around the survey, padded on every side
    off = torch.tensor([0.0, y0 * 10.0])
    ref = Propagator({"vp": v[:z1, y0:y1].contiguous().to(DEV)}, 10.0)(
        src.to(DEV), (x_s - off).to(DEV), (x_r - off).to(DEV), 1e-3)
    assert torch.equal(rec, ref)


def test_linearity_in_the_source_and_shot_permutation_at_c1_size(cuda_lib):
    """Size-independent properties on the reference's own 100x300 case (BASELINE config 1):
    traces are linear in the source amplitudes, and permuting the shots permutes the traces and
    leaves the model gradient unchanged up to summation order."""
    from physicsbasedfwi2_b200.scalar import Propagator
    nz, ny, S, R, nt = 100, 300, 10, 101, 400
    v = K.layered_model(nz, ny, seed=0).to(DEV)
    x_s, x_r = K.survey(nz, ny, 10.0, S, R, 1, seed=0)
    src = K.ricker(15.0, nt, 1e-3, 1 / 15.0).reshape(-1, 1, 1).repeat(1, S, 1).contiguous()
    g = torch.randn(nt, S, R, generator=torch.Generator().manual_seed(1)).to(DEV)

    def run(s, perm=None):
        xs, xr, gg = x_s, x_r, g
        if perm is not None:
            xs, xr, s, gg = x_s[perm], x_r[perm], s[:, perm], g[:, perm]
        vg = v.clone().requires_grad_(True)
        rec = Propagator({"vp": vg}, 10.0, vpmax=4000.0)(s.to(DEV), xs.to(DEV), xr.to(DEV), 1e-3)
        (rec * gg).sum().backward()
        return rec.detach(), vg.grad

    r1, g1 = run(src)
    r2, _ = run(-2.0 * src)                 # a power of two: exact in fp32 up to underflow
    assert K.rel_l2(r2, -2.0 * r1) <= 1e-6
    perm = torch.randperm(S, generator=torch.Generator().manual_seed(2))
    r3, g3 = run(src, perm)
    assert torch.equal(r3, r1[:, perm.to(DEV)])
    assert K.rel_l2(g3, g1) <= 1e-6


def test_error_behaviour_is_runtime_error(cuda_lib):
    from physicsbasedfwi2_b200.scalar import Propagator
    v, x_s, x_r, src = _setup()
    vd, xs, xr, sd = v.to(DEV), x_s.to(DEV), x_r.to(DEV), src.to(DEV)
    with pytest.raises(RuntimeError):
        Propagator({"vp": vd.double()}, 10.0)(sd, xs, xr, 1e-3)                  # dtype
    with pytest.raises(RuntimeError):
        Propagator({"vp": vd[None]}, 10.0)(sd, xs, xr, 1e-3)                     # 3-D model
    with pytest.raises(RuntimeError):
        Propagator({"vp": vd}, 10.0)(sd[:, :1], xs, xr, 1e-3)                    # shot count mismatch
    with pytest.raises(RuntimeError):
        bad = xr.clone()
        bad[0, 0, 1] = 1e6
        Propagator({"vp": vd}, 10.0)(sd, xs, bad, 1e-3)                          # receiver outside the model
    with pytest.raises(RuntimeError):
        Propagator({"vp": vd}, 10.0, accuracy=5)(sd, xs, xr, 1e-3)
    with pytest.raises(RuntimeError):
        Propagator({"vs": vd}, 10.0)
    with pytest.raises(RuntimeError):
        Propagator({"vp": vd}, 10.0, slabs=True)(sd, xs, xr, 1e-3)               # no process group


def test_devito_style_fwi_loss_adapter(cuda_lib):
    """seisgan/fwi/layers.py:145-201 contract: objective = sum 0.5||d - d_obs||^2, backward returns the
    squared-slowness gradient divided by its max |.| regardless of the incoming gradient."""
    from oracle import cpu_ref as O
    from physicsbasedfwi2_b200.fwiloss import fwi_loss
    v, x_s, x_r, src = _setup(nz=34, ny=90, S=3, R=12, nt=90, seed=6)
    obs = 0.05 * torch.randn(90, 3, 12, generator=torch.Generator().manual_seed(7))
    rec0, gv0, _ = O.acoustic_gradient(v, 10.0, src, x_s, x_r, 1e-3, observed=obs, dtype=torch.float32)
    gm0 = gv0 * (-0.5 * v.double() ** 3)
    x = v.to(DEV)[None, None].requires_grad_(True)
    J = fwi_loss(x, obs.to(DEV), src.to(DEV), x_s.to(DEV), x_r.to(DEV), 10.0, 1e-3)
    assert tuple(J.shape) == (1,)
    J0 = 0.5 * float(((rec0 - obs) ** 2).sum())
    assert abs(float(J) - J0) <= 1e-5 * J0
    (123.0 * J).sum().backward()                       # the incoming gradient is ignored, as in the reference
    assert tuple(x.grad.shape) == (1, 1, 34, 90)
    assert abs(float(x.grad.abs().max()) - 1.0) < 1e-6
    assert K.rel_l2(x.grad[0, 0], gm0 / gm0.abs().max()) <= 1e-4


def test_inner_loop_fwi_reduces_the_misfit_and_matches_the_torch_glue(cuda_lib):
    """models/networks.py:15372-15415 on the B200 kernels: the first misfit equals the one the plain torch
    expressions of the reference give on the propagator's output, and Adam(lr=10) steps reduce it."""
    from oracle import misfit_ref as M
    from physicsbasedfwi2_b200.inversion import inner_loop_fwi
    from physicsbasedfwi2_b200.scalar import Propagator
    nz, ny, S, R, nt = 40, 100, 4, 50, 300
    true = K.layered_model(nz, ny, seed=2, noise=0.0).clamp(1500.0, 4400.0)
    start = torch.full((nz, ny), 2500.0)
    x_s, x_r = K.survey(nz, ny, 10.0, S, R, 1, seed=3)
    wavelet = K.ricker(12.0, nt, 1e-3, 1 / 12.0).reshape(-1, 1, 1)
    src = wavelet.repeat(1, S, 1).contiguous().to(DEV)
    with torch.no_grad():
        observed = Propagator({"vp": true.to(DEV)}, 10.0)(src, x_s.to(DEV), x_r.to(DEV), 1e-3)
        direct = Propagator({"vp": torch.full((nz, ny), 1500.0, device=DEV)}, 10.0)(src, x_s.to(DEV), x_r.to(DEV), 1e-3)
        pred0 = Propagator({"vp": start.to(DEV)}, 10.0)(src, x_s.to(DEV), x_r.to(DEV), 1e-3)
    o = observed - direct
    obs_norm = o / (o.abs().amax(0, keepdim=True) + 1e-10)
    first = float(M.trace_misfit(pred0.cpu().double(), obs_norm.cpu().double(), direct.cpu().double(), kind="mse"))
    mask = torch.zeros(nz, ny, dtype=torch.bool, device=DEV)
    mask[:3] = True
    v, losses = inner_loop_fwi(start.to(DEV), observed, wavelet, x_s, x_r, 10.0, 1e-3, num_epochs=12, num_batches=2,
                               grad_mask=mask)
    assert len(losses) == 24 and tuple(v.shape) == (nz, ny)
    l_full = 0.5 * (losses[0] + losses[1])        # two half-batches of the first epoch bracket the full-survey misfit
    assert abs(l_full - first) < 0.15 * first
    assert max(losses[-2:]) < 0.95 * min(losses[:2])      # 12 Adam steps of 10 m/s each: a steady decrease
    assert torch.all(v[:3] == 2500.0)             # masked rows never move
    assert float(v.min()) >= 1500.0 and float(v.max()) <= 4400.0


def test_inner_loop_cuda_graph_replays_match_the_eager_loop(cuda_lib):
    """The loop of models/networks.py:15372-15415 in its own geometry (101 x 101, dx 10 m, 15 Hz, dt 1.5 ms, 10 shots x
    101 receivers at z = 0, Adam lr 10, clamp [1500, 4400]): captured as one CUDA graph per batch index and replayed
    (no host sync, losses read back once) against the same loop run eagerly."""
    from physicsbasedfwi2_b200.inversion import inner_loop_fwi
    from physicsbasedfwi2_b200.scalar import Propagator
    nz = ny = 101
    S, R, nt, dt, dx = 10, 101, 240, 1.5e-3, 10.0
    true = K.layered_model(nz, ny, seed=4, noise=0.0).clamp(1500.0, 4400.0)
    start = torch.full((nz, ny), 2400.0)
    x_s = torch.zeros(S, 1, 2)
    x_s[:, 0, 1] = torch.arange(S).float() * (101 * dx / S)
    x_r = torch.zeros(S, R, 2)
    x_r[:, :, 1] = (torch.arange(R).float() * dx)[None]
    wavelet = K.ricker(15.0, nt, dt, 1 / 15.0).reshape(-1, 1, 1)
    with torch.no_grad():
        observed = Propagator({"vp": true.to(DEV)}, dx, vpmax=4400.0)(wavelet.repeat(1, S, 1).contiguous().to(DEV),
                                                                      x_s.to(DEV), x_r.to(DEV), dt)
    outs = {}
    for graph, nb, shuffle in ((False, 1, False), (True, 1, False), (False, 2, False), (True, 2, False), (True, 2, True)):
        torch.manual_seed(0)
        v, losses = inner_loop_fwi(start.to(DEV), observed, wavelet, x_s, x_r, dx, dt, num_epochs=9, num_batches=nb,
                                   graph=graph, shuffle=shuffle)
        assert len(losses) == 9 * nb and all(l == l and l > 0 for l in losses)
        outs[(graph, nb, shuffle)] = (v, losses)
    for nb in (1, 2):
        (v0, l0), (v1, l1) = outs[(False, nb, False)], outs[(True, nb, False)]
        assert max(abs(a - b) / a for a, b in zip(l0, l1)) < 1e-5, (l0, l1)
        assert K.rel_l2(v1, v0) < 1e-6
        assert l0[-1] < l0[0]
    assert outs[(True, 2, True)][1][-1] < outs[(True, 2, True)][1][0]        # shuffled batches: still a descent


@pytest.mark.parametrize("small", ["1", "0"], ids=["cluster-resident", "step-kernel"])
def test_many_points_per_shot_and_a_very_shallow_grid(cuda_lib, small, monkeypatch):
    """More receivers than the 512 threads of a CTA of the cluster-resident kernel (its slow-path loops),
    several sources per shot with repeats, and a grid only 9 rows deep (bands shorter than a cluster)."""
    from oracle import cpu_ref as O
    from physicsbasedfwi2_b200.scalar import Propagator
    monkeypatch.setenv("FWI_B200_SMALL", small)
    for nz, ny, R, order, pml in ((9, 180, 700, 4, 6), (26, 90, 1100, 8, 3)):
        v = K.layered_model(nz, ny, seed=8, blob=False)
        g = torch.Generator().manual_seed(9)
        S, Ns, nt = 2, 3, 70
        x_s = torch.stack([torch.randint(0, nz, (S, Ns), generator=g), torch.randint(0, ny, (S, Ns), generator=g)], -1).float() * 10.0
        x_s[0, 2] = x_s[0, 0]                                   # a repeated source
        x_r = torch.stack([torch.randint(0, nz, (S, R), generator=g), torch.randint(0, ny, (S, R), generator=g)], -1).float() * 10.0
        src = torch.randn(nt, S, Ns, generator=g)
        gr = torch.randn(nt, S, R, generator=g)
        vg = v.to(DEV).requires_grad_(True)
        sg = src.to(DEV).requires_grad_(True)
        rec = Propagator({"vp": vg}, 10.0, pml_width=pml, accuracy=order)(sg, x_s.to(DEV), x_r.to(DEV), 1e-3)
        (rec * gr.to(DEV)).sum().backward()
        r0, gv0, gs0 = O.acoustic_gradient(v, 10.0, src, x_s, x_r, 1e-3, grad_rec=gr, order=order, pml_width=pml,
                                           dtype=torch.float32)
        assert K.rel_l2(rec, r0) <= 1e-5 and K.rel_l2(vg.grad, gv0) <= 1e-4 and K.rel_l2(sg.grad, gs0) <= 1e-4
