"""Host-side logic of the propagator (runs on CPU tensors): CFL sub-stepping, PML profile, FFT
resampling and its transpose, coordinate conversion, replicate-pad fold-back, tile lists, Ricker.
Checked against the oracle's independent statements of the same conventions."""
import math

import numpy as np
import pytest
import torch

from oracle import cpu_ref as O
from oracle import torch_ref as T
from physicsbasedfwi2_b200 import _host as H
from physicsbasedfwi2_b200 import wavelets


def test_ricker_matches_call_site_formula():
    # models/networks.py:5357  deepwave.wavelets.ricker(freq, nt, dt, 1/freq)
    w = wavelets.ricker(8.0, 4001, 1e-3, 1 / 8.0)
    assert w.shape == (4001,) and w.dtype == torch.float32
    assert abs(float(w[125]) - 1.0) < 1e-6            # peak at t = peak_time
    t = np.arange(4001) * 1e-3 - 0.125
    ref = (1 - 2 * (np.pi * 8 * t) ** 2) * np.exp(-(np.pi * 8 * t) ** 2)
    assert np.abs(w.numpy() - ref).max() < 1e-6
    assert torch.equal(w, T.ricker(8.0, 4001, 1e-3, 1 / 8.0))
    assert wavelets.ricker(8.0, 0, 1e-3, 0.1).numel() == 0


@pytest.mark.parametrize("dt,dx,vmax", [(1e-3, 10.0, 3550.0), (1e-3, 10.0, 4700.0), (1.5e-3, 10.0, 3000.0),
                                        (4e-4, 4.0, 4200.0), (1e-3, (15.0, 10.0), 2000.0)])
def test_inner_step_matches_oracle(dt, dx, vmax):
    dz, dy = H.spacing(dx)
    assert H.inner_step(dt, dz, dy, vmax) == T.cfl_substep(dt, dz, dy, vmax)
    ratio, dti = H.inner_step(dt, dz, dy, vmax)
    assert ratio >= 1 and abs(dti * ratio - dt) < 1e-15
    assert dti <= 0.6 / (math.sqrt(1 / dz ** 2 + 1 / dy ** 2) * vmax) + 1e-12


def test_spacing_forms():
    assert H.spacing(5) == (5.0, 5.0)
    assert H.spacing([4.0]) == (4.0, 4.0)
    assert H.spacing(torch.tensor([3.0, 2.0])) == (3.0, 2.0)
    with pytest.raises(RuntimeError):
        H.spacing([1.0, 2.0, 3.0])


@pytest.mark.parametrize("n,h,pw", [(124, 2, 10), (60, 4, 20), (30, 1, 0)])
def test_sigma_profile_matches_oracle(n, h, pw):
    a = H.sigma_profile(n, h, pw, 10.0, 3000.0)
    b = T.pml_profile(n, h, pw, 10.0, 3000.0)
    assert torch.equal(a, b)
    if pw:
        assert float(a[h + pw:n - h - pw].abs().max()) == 0.0
        assert float(a[0]) == float(a[h]) == float(a.max())


@pytest.mark.parametrize("n,ratio", [(40, 2), (41, 3), (16, 4), (10, 1)])
def test_resample_and_transpose(n, ratio):
    g = torch.Generator().manual_seed(n)
    x = torch.randn(n, 2, 3, generator=g, dtype=torch.float64)
    y = H.resample_up(x, ratio)
    assert torch.allclose(y, T.upsample(x, ratio), atol=1e-13)
    import scipy.signal
    assert np.allclose(y.numpy(), scipy.signal.resample(x.numpy(), n * ratio, axis=0), atol=1e-12)
    assert torch.allclose(y[::ratio], x, atol=1e-12)             # interpolating
    r = torch.randn(n * ratio, 2, 3, generator=g, dtype=torch.float64)
    xt = H.resample_up_adjoint(r, ratio, x)
    assert abs(float((y * r).sum()) - float((x * xt).sum())) < 1e-10


def test_to_cells_trunc_and_round():
    x = torch.tensor([[[0.0, 29.999998], [10.0, 117.05882]]])
    assert H.to_cells(x, 10.0, 10.0, "trunc").tolist() == [[[0, 3], [1, 11]]]
    assert H.to_cells(x, 10.0, 10.0, "round").tolist() == [[[0, 3], [1, 12]]]
    with pytest.raises(RuntimeError):
        H.to_cells(x, 10.0, 10.0, "nearest")


@pytest.mark.parametrize("pad", [0, 1, 5])
def test_replicate_pad_adjoint_is_the_transpose(pad):
    g = torch.Generator().manual_seed(pad)
    x = torch.randn(7, 9, generator=g, dtype=torch.float64)
    xp = torch.nn.functional.pad(x[None, None], (pad,) * 4, mode="replicate")[0, 0] if pad else x
    r = torch.randn(xp.shape, generator=g, dtype=torch.float64)
    assert abs(float((xp * r).sum()) - float((x * H.replicate_pad_adjoint(r, pad)).sum())) < 1e-12
    assert torch.allclose(H.replicate_pad_adjoint(r, pad), O.fold_replicate_pad(r, pad))


def test_injection_lists_sorted_by_cell_with_row_offsets():
    S, n, nzp, pitch = 3, 9, 20, 64
    g = torch.Generator().manual_seed(0)
    z = torch.randint(0, nzp, (S, n), generator=g)
    y = torch.randint(0, pitch, (S, n), generator=g)
    z[0, 5], y[0, 5] = z[0, 2], y[0, 2]                  # a duplicate point keeps its order
    cell = z * pitch + y
    start, c, col = H.injection_lists(cell, nzp, pitch)
    start = start.reshape(S, nzp + 1)
    assert int(start[0, 0]) == 0 and int(start[-1, -1]) == S * n
    for s in range(S):
        assert int(start[s, 0]) == s * n and int(start[s, nzp]) == (s + 1) * n
        seg = c[s * n:(s + 1) * n].tolist()
        assert seg == sorted(seg)
        for e in range(s * n, (s + 1) * n):
            assert int(cell[s, int(col[e])]) == int(c[e])
        for row in range(nzp):
            for e in range(int(start[s, row]), int(start[s, row + 1])):
                assert int(c[e]) // pitch == row
    dup = [e for e in range(n) if int(c[e]) == int(cell[0, 2])]
    assert [int(col[e]) for e in dup] == [2, 5]
    start0, c0, col0 = H.injection_lists(cell[:, :0], nzp, pitch)
    assert int(start0.abs().sum()) == 0 and start0.numel() == S * (nzp + 1)


def test_injection_lists_leave_out_masked_points():
    """x-slab windows (slab.py) drop the points outside the window: no row range may reach them."""
    S, n, nzp, pitch = 2, 12, 10, 32
    g = torch.Generator().manual_seed(1)
    cell = torch.randint(0, nzp, (S, n), generator=g) * pitch + torch.randint(0, pitch, (S, n), generator=g)
    keep = torch.rand(S, n, generator=g) < 0.6
    start, c, col = H.injection_lists(cell, nzp, pitch, keep)
    start = start.reshape(S, nzp + 1)
    seen = []
    for s in range(S):
        for row in range(nzp):
            for e in range(int(start[s, row]), int(start[s, row + 1])):
                assert int(c[e]) // pitch == row and bool(keep[s, int(col[e])])
                assert int(cell[s, int(col[e])]) == int(c[e])
                seen.append((s, int(col[e])))
    assert sorted(seen) == sorted((s, j) for s in range(S) for j in range(n) if bool(keep[s, j]))


def test_su_reader_round_trip(tmp_path):
    """SU seismograms (the files the reference's elastic prop shuffles at models/networks.py:9051-9071)."""
    import numpy as np
    from physicsbasedfwi2_b200.denise_glue import denise_geometry, read_su, write_su
    tr = np.random.default_rng(0).standard_normal((7, 301)).astype(np.float32)
    p = str(tmp_path / "seis_x.su.shot1")
    write_su(p, tr, 2e-3)
    back, dt = read_su(p)
    assert back.shape == (7, 301) and np.array_equal(back, tr) and abs(dt - 2e-3) < 1e-12
    x_s, x_r = denise_geometry([3, 0, 45, 7, 12, 20])
    assert tuple(x_s.shape) == (6, 1, 2) and tuple(x_r.shape) == (6, 370, 2)          # 370 receivers, :8994-9001
    assert float(x_s[2, 0, 1]) == 400.0 + 45 * 160.0 and float(x_r[0, -1, 1]) == 7780.0 and float(x_r[0, 0, 0]) == 460.0
