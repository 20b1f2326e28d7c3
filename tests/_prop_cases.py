"""Seeded inputs of the reference's `prop` methods used as drop-in fixtures, and the replay of their call sequence.

`auto22_inputs` / `auto_inputs` build what `AutoMarmousi22_Net.prop` (models/networks.py:5314) and `Auto_Net.prop`
(:2711) receive from their `forward`.  `replay_auto22` / `replay_auto` restate, line by line, what those methods do
around their deepwave calls, with `deepwave` a module argument: tests/test_dropin_gpu.py passes shim/deepwave (the
CUDA product) and compares with fixtures that tests/golden/make_prop_golden.py produced by running the reference's
own, unmodified methods (tests/_ref_harness.py).  The replays are test code, not product code.
"""
from __future__ import annotations

import torch


class CpuPinned(torch.Tensor):
    """`tensor.get_device()` is -1 on the CPU and `x.to(-1)` is an error; the reference's small-grid `prop` methods
    do exactly that (`devicek = net1out1.get_device()`, models/networks.py:2723).  This subclass answers with the
    tensor's device instead, so those methods run unmodified in the CPU harness."""

    def get_device(self):
        return self.device


def _layers(nz, ny, vmin, vmax, seed, water=0):
    g = torch.Generator().manual_seed(seed)
    z = torch.arange(nz, dtype=torch.float32)[:, None].expand(nz, ny)
    y = torch.arange(ny, dtype=torch.float32)[None, :].expand(nz, ny)
    v = vmin + (vmax - vmin) * (z / (nz - 1)) ** 0.8
    v = v + 0.06 * (vmax - vmin) * torch.sin(z / 9.0 + y / 40.0) + 0.02 * (vmax - vmin) * torch.randn(nz, ny, generator=g)
    v = v.clamp(vmin, vmax)
    if water:
        v[:water] = vmin
    return v.contiguous()


def _smooth(v, k):
    p = torch.nn.functional.pad(v[None, None], (k, k, k, k), mode="replicate")
    return torch.nn.functional.avg_pool2d(p, 2 * k + 1, stride=1)[0, 0]


def _gathers(shape, seed):
    """Band-limited seeded 'observed' gathers [1, shots, nt, receivers] (the misfit only normalises and compares them)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(shape, generator=g)
    k = torch.hann_window(31, periodic=False)
    k = (k / k.sum())[None, None, :, None]
    x = torch.nn.functional.conv2d(x.transpose(0, 1), k, padding=(15, 0)).transpose(0, 1)
    return x.contiguous()


def auto22_inputs():
    """AutoMarmousi22_Net: 151 x 200, 18 shots x 200 receivers x 4001 samples (models/networks.py:5339-5355)."""
    true = _layers(151, 200, 1500.0, 4500.0, seed=11, water=26)
    vel = _smooth(true, 6)
    vel[:26] = 1500.0
    vel = vel + 30.0 * torch.randn(151, 200, generator=torch.Generator().manual_seed(12))
    vel[:26] = 1500.0
    return dict(inputs=_gathers((1, 18, 4001, 200), 13), vel=vel[None, None].contiguous(), true=true[None, None].contiguous(),
                mintrue=true.min(), maxtrue=true.max(), seed=14)


def auto_inputs():
    """Auto_Net: 70 x 70 in km/s, 5 shots x 70 receivers x 1000 samples (models/networks.py:2726-2733)."""
    true = _layers(70, 70, 1.5, 4.0, seed=21)
    vel = _smooth(true, 4) + 0.03 * torch.randn(70, 70, generator=torch.Generator().manual_seed(22))
    return dict(inputs=_gathers((1, 5, 1000, 70), 23), vel=vel[None, None].contiguous(),
                mintrue=true.min(), maxtrue=true.max(), seed=24)


def replay_auto22(deepwave, c, device):
    """models/networks.py:5314-5510 around its deepwave calls.  Returns (grad [151, 200], loss, pred [4001, 9, 200])."""
    devicek = torch.device(device)
    inputs, vel, true, mintrue, maxtrue = c["inputs"], c["vel"], c["true"], c["mintrue"], c["maxtrue"]
    torch.manual_seed(c["seed"])
    net1out1 = torch.squeeze(vel.detach())                                                  # :5328-5329
    g1 = torch.arange(net1out1.size(dim=0)) ** 2.0                                          # :5330-5331
    ss = torch.transpose(g1.tile((200, 1)), 0, 1)                                           # :5332-5333
    net1out1 = net1out1.to(devicek)
    freq, dx, nt, dt, num_shots, R, ny = 8, 10, 4001, 0.001, 18, 200, 200                   # :5339-5349
    receiver_spacing = 200 * dx / R
    x_s = torch.zeros(num_shots, 1, 2)
    x_s[:, 0, 1] = torch.linspace(0, (ny - 1) * dx, num_shots)                              # :5352
    x_r = torch.zeros(num_shots, R, 2)
    x_r[0, :, 1] = torch.arange(R).float() * receiver_spacing
    x_r[:, :, 1] = x_r[0, :, 1].repeat(num_shots, 1)
    source_amplitudes_true = deepwave.wavelets.ricker(freq, nt, dt, 1 / freq).reshape(-1, 1, 1).to(devicek)   # :5357-5360
    num_batches = 2
    num_shots_per_batch = int(num_shots / num_batches)
    ss = ss.to(devicek)
    receiver_amplitudes_true = inputs[0, :, :, :].swapaxes(0, 1).to(devicek)                # :5379-5381
    mat2 = torch.ones(net1out1.size()[0], net1out1.size()[1]).to(devicek) * 1500            # :5398-5399
    src_amps = source_amplitudes_true.repeat(1, num_shots, 1)
    prop2 = deepwave.scalar.Propagator({'vp': mat2}, dx)                                    # :5408
    receiver_amplitudes_cte = prop2(src_amps, x_s.to(devicek), x_r.to(devicek), dt)         # :5409-5411
    rcv_amps_true_max, _ = torch.abs(receiver_amplitudes_true).max(dim=0, keepdim=True)     # :5418
    rcv_amps_true_norm = receiver_amplitudes_true / (rcv_amps_true_max.abs() + 1e-10)       # :5419
    criterion1 = torch.nn.L1Loss()
    net1out1.requires_grad = True                                                           # :5429
    idx = torch.randperm(num_shots)                                                         # :5434
    x_s = x_s.view(-1, 2)[idx].view(x_s.size())
    rcv_amps_true_norm = rcv_amps_true_norm[:, idx, :]
    receiver_amplitudes_cte = receiver_amplitudes_cte[:, idx, :]
    it = 0                                                                                  # :5442 `for it in range(1)`
    model2 = torch.clamp(net1out1, min=mintrue, max=maxtrue)                                # :5446
    prop = deepwave.scalar.Propagator({'vp': model2}, dx)                                   # :5449
    batch_src_amps = source_amplitudes_true.repeat(1, num_shots_per_batch, 1)
    batch_rcv_amps_true = rcv_amps_true_norm[:, it::num_batches]
    batch_rcv_amps_cte = receiver_amplitudes_cte[:, it::num_batches]
    batch_x_s = x_s[it::num_batches].to(devicek)
    batch_x_r = x_r[it::num_batches].to(devicek)
    pred0 = prop(batch_src_amps, batch_x_s, batch_x_r, dt)                                  # :5464
    pred = pred0 - batch_rcv_amps_cte                                                       # :5467
    pred_max, _ = torch.abs(pred).max(dim=0, keepdim=True)
    pred_norm = pred / (pred_max.abs() + 1e-10)                                             # :5470
    lossinner = criterion1(pred_norm, batch_rcv_amps_true)                                  # :5476
    lossinner.backward()                                                                    # :5491
    net1out1.grad = net1out1.grad * ss                                                      # :5492
    net1out1.grad[(true[0, 0, :, :] == 1500)] = 0                                           # :5493
    return net1out1.grad, lossinner.item(), pred0.detach()


def replay_auto(deepwave, c, device):
    """models/networks.py:2711-2850 (Auto_Net.prop, km/s model scaled by 1000).  Returns (grad [70, 70], loss, pred)."""
    devicek = torch.device(device)
    inputs, vel, mintrue, maxtrue = c["inputs"], c["vel"], c["mintrue"], c["maxtrue"]
    torch.manual_seed(c["seed"])
    net1out1 = torch.squeeze((vel * 1000).detach()).to(devicek)
    freq, dx, nt, dt, num_shots, R = 25, 15, 1000, 0.001, 5, 70
    source_spacing = 70 * dx / num_shots
    receiver_spacing = 70 * dx / R
    x_s = torch.zeros(num_shots, 1, 2)
    x_s[:, 0, 1] = torch.arange(num_shots).float() * source_spacing
    x_r = torch.zeros(num_shots, R, 2)
    x_r[0, :, 1] = torch.arange(R).float() * receiver_spacing
    x_r[:, :, 1] = x_r[0, :, 1].repeat(num_shots, 1)
    source_amplitudes_true = deepwave.wavelets.ricker(freq, nt, dt, 1 / freq).reshape(-1, 1, 1).to(devicek)
    receiver_amplitudes_true = inputs[0, :, :, :].swapaxes(0, 1).to(devicek)
    mat2 = torch.ones(net1out1.size()[0], net1out1.size()[1]).to(devicek) * mintrue * 1000
    src_amps = source_amplitudes_true.repeat(1, num_shots, 1)
    receiver_amplitudes_cte = deepwave.scalar.Propagator({'vp': mat2}, dx)(src_amps, x_s.to(devicek), x_r.to(devicek), dt)
    rcv_amps_true_max, _ = torch.abs(receiver_amplitudes_true).max(dim=0, keepdim=True)
    rcv_amps_true_norm = receiver_amplitudes_true / (rcv_amps_true_max.abs() + 1e-10)
    net1out1.requires_grad = True
    idx = torch.randperm(num_shots)
    x_s = x_s.view(-1, 2)[idx].view(x_s.size())
    rcv_amps_true_norm = rcv_amps_true_norm[:, idx, :]
    model2 = torch.clamp(net1out1, min=mintrue * 1000, max=maxtrue * 1000)
    prop = deepwave.scalar.Propagator({'vp': model2}, dx)
    pred = prop(source_amplitudes_true.repeat(1, num_shots, 1), x_s.to(devicek), x_r.to(devicek), dt)
    pred_max, _ = torch.abs(pred).max(dim=0, keepdim=True)
    pred_norm = pred / (pred_max.abs() + 1e-10)
    lossinner = torch.nn.L1Loss()(pred_norm, rcv_amps_true_norm)
    lossinner.backward()
    return net1out1.grad, lossinner.item(), pred.detach()
