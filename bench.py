#!/usr/bin/env python
"""bench.py - FWI gradient throughput (forward + adjoint) of the B200 acoustic propagator.

Metric (BASELINE.json): FWI gradient Gcell-updates/s (fwd+adjoint); % of HBM roofline.
Workload (BASELINE.json configs[1], SURVEY.md 8d "C2"): Marmousi-scale 2-D acoustic model
751 (z) x 2301 (y), dx = 4 m, 8th-order stencil, 20-cell PML, 3000 time steps, 15 Hz Ricker,
576 receivers per shot.  One "step" = one FWI gradient (forward, MSE misfit, adjoint with imaging)
of one BATCH of shots per GPU; the 256-shot survey is 256 / (batch x n_gpus) such steps, taken one
after the other exactly as the reference mini-batches its shots (models/networks.py:5454-5461).
The batch is what fits HBM with one snapshot per inner step (k = 1, the parity-graded mode).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--survey]

One step (default, c2) = the gradient of the 32 shots a GPU owns when the 256-shot survey is sharded over 8 GPUs
(six launches of 6 6 5 5 5 5 shots, gradients accumulated, ONE allreduce): per-GPU work is fixed ("weak") and the
8-GPU run is the north star's survey itself.  --survey: one step = all 256 shots sharded over the N ranks ("strong").
Under torchrun (N > 1) a small multi-GPU correctness check runs ahead of the timed region ("multi_gpu_check").
One JSON line on rank 0, with "roofline", "cpu_baseline", "parity" (GPU vs the CPU sample) and "e2e".

cell-update = one interior model cell advanced one inner time step in one pass for one shot;
work per step = nz*ny*nt*shots*2.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

WORKLOADS = {
    # BASELINE.json configs[1]
    "c2": dict(nz=751, ny=2301, dx=4.0, order=8, pml=20, nt=3000, dt=4e-4, freq=15.0, n_rec=576,
               rec_stride=4, z_acq=2, survey_shots=256, vmin=1500.0, vmax=4200.0),
    # BASELINE.json configs[0] (the reference's own CPU-runnable case) - parity-test sized
    # BASELINE.json configs[2]: elastic-Marmousi-sized vp / vs / rho (the AutoElFullRho path), explosive sources,
    # two-component receivers; vs = vp / sqrt(3) with a 24-row water layer, rho = 310 vp^0.25 (Gardner)
    "c3": dict(nz=376, ny=1151, dx=10.0, order=4, pml=10, nt=3000, dt=1e-3, freq=8.0, n_rec=370, rec_stride=3,
               z_acq=4, survey_shots=128, vmin=1500.0, vmax=4200.0, elastic=True),
    # BASELINE.json configs[3]: SEAM-sized 2-D acoustic slice with wavefield checkpointing (state kept every
    # 64 steps, snapshots recomputed per segment in the backward pass: one extra forward per gradient)
    "c4": dict(nz=1501, ny=3501, dx=10.0, order=8, pml=20, nt=4000, dt=1e-3, freq=8.0, n_rec=876, rec_stride=4,
               z_acq=2, survey_shots=64, vmin=1500.0, vmax=4500.0, ckpt=64),
    # BASELINE.json configs[3], x-slab variant: every shot is cut into one column slab per GPU, ghost columns
    # exchanged over NVLink every 8 steps (slab.py); snapshot every step; all ranks work on the SAME shots
    "c4dd": dict(nz=1501, ny=3501, dx=10.0, order=8, pml=20, nt=4000, dt=1e-3, freq=8.0, n_rec=876, rec_stride=4,
                 z_acq=2, survey_shots=64, vmin=1500.0, vmax=4500.0, slabs=8),
    # BASELINE.json configs[4]: end-to-end physics-guided inversion iteration in the Auto22 geometry
    # (models/networks.py:5339-5358: 151x200, dx 10 m, 8 Hz, 4001 steps of 1 ms, 18 shots x 200 receivers at z = 0,
    # default deepwave order 4 / pml 10): CNN velocity generator -> propagator -> direct-wave subtraction,
    # trace normalisation, L1 misfit (fused kernel) -> adjoint -> generator backward -> Adam step.
    "c5": dict(nz=151, ny=200, dx=10.0, order=4, pml=10, nt=4001, dt=1e-3, freq=8.0, n_rec=200, rec_stride=1,
               z_acq=0, survey_shots=18, vmin=1500.0, vmax=4500.0, e2e_iteration=True),
    # the reference's inner-loop FWI (models/networks.py:15278-15423, VaeLatentNoPhy_Net.prop):
    # 101 x 101, dx 10 m, 15 Hz, 800 steps of 1.5 ms, 10 shots x 101 receivers at z = 0, 100 Adam(lr=10) iterations
    "f3": dict(nz=101, ny=101, dx=10.0, order=4, pml=10, nt=800, dt=1.5e-3, freq=15.0, n_rec=101, rec_stride=1,
               z_acq=0, survey_shots=10, vmin=1500.0, vmax=4400.0, inner_loop=100),
    "c1": dict(nz=100, ny=300, dx=10.0, order=4, pml=10, nt=1000, dt=1e-3, freq=15.0, n_rec=101,
               rec_stride=3, z_acq=0, survey_shots=10, vmin=1500.0, vmax=3500.0),
}

B_FWD, B_ADJ = 20.0, 28.0   # algorithmic bytes per cell-update at k = 1: 16 + 4 / 16 + 4 + 8
B_FWD_EL, B_ADJ_EL = 80.0, 120.0   # elastic: 60 + 5*4 snapshot write / 60 + 5*4 read + 5*8 image RMW


def synth_model(w, seed=1):
    """Depth gradient vmin -> vmax plus band-limited dipping random layering (no dataset on disk)."""
    nz, ny = w["nz"], w["ny"]
    g = torch.Generator().manual_seed(seed)
    n = nz + ny // 8 + 64
    r = torch.randn(n, generator=g, dtype=torch.float64)
    k = torch.arange(-24, 25, dtype=torch.float64)
    ker = torch.exp(-0.5 * (k / 6.0) ** 2)
    lay = torch.nn.functional.conv1d(r[None, None], (ker / ker.norm())[None, None], padding=24)[0, 0]
    z = torch.arange(nz)[:, None]
    y = torch.arange(ny)[None, :]
    v = w["vmin"] + (w["vmax"] - w["vmin"]) * (z.double() / max(nz - 1, 1)) + 350.0 * lay[(z + y // 8)]
    return v.clamp(w["vmin"], w["vmax"]).float().contiguous()


def synth_survey(w, n_shots, first_shot=0):
    """Shots evenly spaced along y at depth z_acq; receivers every rec_stride-th cell."""
    ny, dx = w["ny"], w["dx"]
    total = w["survey_shots"]
    ys = torch.linspace(8, ny - 9, total).round()
    idx = (first_shot + torch.arange(n_shots)) % total
    x_s = torch.zeros(n_shots, 1, 2)
    x_s[:, 0, 0] = w["z_acq"] * dx
    x_s[:, 0, 1] = ys[idx] * dx
    yr = (torch.arange(w["n_rec"]) * w["rec_stride"]).clamp(max=ny - 1).float()
    x_r = torch.zeros(n_shots, w["n_rec"], 2)
    x_r[:, :, 0] = w["z_acq"] * dx
    x_r[:, :, 1] = yr[None, :] * dx
    return x_s, x_r


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, pw, reasons = [], None, [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1]); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port timed on the host cores.  ONE code path serves both the --impl reference arm
# and the GPU arm's cpu_baseline leg (which runs it in a fresh subprocess, so the number does not depend
# on what the CUDA context, pinned buffers and NCCL threads of the GPU process do to the host).
# ------------------------------------------------------------------------------------------------
def cpu_sample_inputs(w, n_steps):
    """The bounded sample both CPU legs and the bench line's parity block use: the middle shot of the survey,
    the first n_steps of the record, misfit 0.5 |rec - obs|^2 against seeded synthetic 'observed' data."""
    from physicsbasedfwi2_b200.wavelets import ricker
    x_s, x_r = synth_survey(w, 1, first_shot=w["survey_shots"] // 2)
    src = ricker(w["freq"], n_steps, w["dt"], 1.0 / w["freq"]).reshape(-1, 1, 1).contiguous()
    obs = torch.randn(n_steps, 1, w["n_rec"], generator=torch.Generator().manual_seed(2)).mul_(1e-3)
    return x_s, x_r, src, obs


def cpu_reference(w, steps, warmup, step_seconds=6.0, max_steps=400, dump=None):
    """Times `steps` fwd+adjoint gradients of the bounded sample on the C/OpenMP oracle with every host
    thread (set explicitly: torchrun exports OMP_NUM_THREADS=1).  Snapshot buffers are allocated and touched
    once (warm-up) and reused.  Returns the cpu_baseline dict; `dump`: path of an .npz that receives the
    sample's traces and gradient for the parity block."""
    from oracle import cpu_ref as O
    O.build()
    cores = O.set_threads(os.cpu_count() or 1)
    v = synth_model(w)
    kw = dict(order=w["order"], pml_width=w["pml"], vpmax=w["vmax"], dtype=torch.float32)
    ws = {}

    def gradient(n):
        x_s, x_r, src, obs = cpu_sample_inputs(w, n)
        t0 = time.perf_counter()
        out = O.acoustic_gradient(v, w["dx"], src, x_s, x_r, w["dt"], observed=obs, workspace=ws, **kw)
        return time.perf_counter() - t0, out

    probe = min(w["nt"], 24)
    gradient(probe)                                             # threads, page cache
    per_step = gradient(probe)[0] / probe
    n = int(max(10, min(w["nt"], step_seconds / max(per_step, 1e-9), max_steps)))
    t_warm = gradient(n)[0]                                     # first touch of the n-step snapshot buffer
    if t_warm > 1.3 * step_seconds:                             # the short probe ran from cache: resize the sample
        n = int(max(10, n * step_seconds / t_warm))
        gradient(n)
    for _ in range(max(0, warmup - 1)):
        gradient(n)
    t0 = time.perf_counter()
    for _ in range(steps):
        _, out = gradient(n)
    t = time.perf_counter() - t0
    val = w["nz"] * w["ny"] * n * 2 * steps / t / 1e9
    if dump:
        import numpy as np
        np.savez(dump, rec=out[0].numpy(), grad_v=out[1].numpy(), n=n)
    return {"value": round(val, 4), "unit": "Gcell-updates/s", "cores": cores, "kind": "port",
            "sample": f"1 shot x {n} of {w['nt']} steps per step, fwd+adjoint, same grid/order/PML, fp32, "
                      f"oracle/cpu_ref.c OpenMP, {steps} timed + {max(1, warmup)} warm-up repeats, buffers reused",
            "seconds": round(t, 2), "sample_steps": n, "ms_per_step": round(1e3 * t / steps, 3)}


def cpu_baseline(w, dump=None):
    """cpu_baseline leg of the GPU arm: the reference arm's own code in a fresh process (2 timed repeats)."""
    cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--workload", w["name"],
           "--steps", "2", "--warmup", "1", "--nt", str(w["nt"])]
    if dump:
        cmd += ["--dump", dump]
    env = dict(os.environ)
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "OMP_NUM_THREADS"):
        env.pop(k, None)
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=900)
    if r.returncode != 0:
        raise RuntimeError("cpu_baseline subprocess failed:\n" + r.stdout[-2000:] + r.stderr[-2000:])
    line = json.loads(r.stdout.strip().splitlines()[-1])
    return line["cpu_baseline"]


def cpu_baseline_elastic(w, n=5):
    """Elastic workloads: the torch restatement (oracle/elastic_ref.py; autograd adjoint) on the host cores,
    one shot, `n` time steps.  There is no C port of the elastic scheme; DENISE (the reference's CPU code
    for this path) is external and absent."""
    import math as _m
    from oracle import elastic_ref as E
    from physicsbasedfwi2_b200.wavelets import ricker
    v = synth_model(w).double()
    vs = v / _m.sqrt(3.0)
    vs[:24] = 0.0
    rho = 310.0 * v ** 0.25
    x_s, x_r = synth_survey(w, 1, first_shot=w["survey_shots"] // 2)
    src = ricker(w["freq"], n, w["dt"], 1.0 / w["freq"]).reshape(-1, 1, 1).double()
    g = torch.randn(2, n, 1, w["n_rec"], generator=torch.Generator().manual_seed(2), dtype=torch.float64)
    kw = dict(order=w["order"], pml_width=w["pml"], vpmax=w["vmax"])
    t0 = time.perf_counter()
    E.elastic_gradient(v.float(), vs.float(), rho.float(), w["dx"], src.float(), x_s, x_r, w["dt"], g.float(), **kw)
    t = time.perf_counter() - t0
    val = w["nz"] * w["ny"] * n * 2 / t / 1e9
    return {"value": round(val, 4), "unit": "Gcell-updates/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"1 shot x {n} of {w['nt']} steps, fwd+adjoint (torch autograd), same grid/order/C-PML, fp32, "
                      f"oracle/elastic_ref.py ({t:.1f} s)", "seconds": round(t, 2)}


def run_reference(args, w):
    """--impl reference: the reference's CPU path for this workload.  The reference's propagators
    (deepwave, DENISE) are external and absent, so this is the oracle port (kind 'port').  Under torchrun
    rank 0 alone runs it, with every host thread."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if w.get("elastic"):
        cb = cpu_baseline_elastic(w)
        ms = 1e3 * cb["seconds"]
    else:
        cb = cpu_reference(w, args.steps, args.warmup, dump=args.dump)
        ms = cb["ms_per_step"]
    val = cb["value"]
    line = {"impl": "reference", "metric": "FWI gradient Gcell-updates/s (fwd+adjoint)",
            "value": val, "unit": "Gcell-updates/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": bench_config(w, 1, 1, args.workload),
            "cpu_baseline": cb,
            "e2e": {"value": val, "unit": "Gcell-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def bench_config(w, batch, world, name):
    return {"workload": f"{name}: 2-D {'elastic P-SV' if w.get('elastic') else 'acoustic'} {w['nz']}x{w['ny']} dx={w['dx']}m order {w['order']} "
                        f"pml {w['pml']} nt {w['nt']} dt {w['dt']} {w['freq']}Hz Ricker, "
                        f"{w['n_rec']} receivers/shot, fwd+adjoint gradient (MSE), "
                        + (f"wavefield checkpoint every {w['ckpt']} steps + recompute" if w.get("ckpt") else
                           "snapshot every step" if w.get("save_every", 1) == 1 else f"snapshot every {w['save_every']} steps (decimated imaging)"),
            "shots_per_gpu_per_step": batch, "shots_per_step": batch * (1 if w.get("slabs") else world),
            "survey_shots": w["survey_shots"],
            "parallelism": (f"x-slab domain decomposition x{world}, ghost exchange every {w['slabs']} steps"
                            if w.get("slabs") else f"shot-parallel x{world}"),
            "l2": "working set per launch exceeds L2 (no flush needed)"}


# ------------------------------------------------------------------------------------------------
# C5: end-to-end physics-guided iteration (generator CNN + propagator gradient + optimiser step)
# ------------------------------------------------------------------------------------------------
def make_generator(n_shots, nz, ny, vmin, vmax):
    """Velocity generator with the layer shapes of the reference's Auto22 net (models/networks.py:5136-5294;
    blocks :2276-2314, :2482-2524): encoder on the gathers decimated 4x in time [1, shots, 1000, 200], four blocks of
    2 x (3x3 conv, batch norm, LeakyReLU 0.1) + 2x2 average pooling (ceil) with 16/32/64/128 channels ->
    [1, 128, 63, 13] -> Linear -> 8-number latent -> Linear -> [1, 128, 19, 25] -> three blocks of bilinear 2x
    upsampling + 2 x (conv, norm, LeakyReLU) with 64/32/16 channels -> crop to [nz, ny] -> 1x1 conv -> sigmoid ->
    vmin + f (vmax - vmin).  Written here from those shapes, random weights: bench.py may not read /root/reference
    on the GPU box, and the CNN is outside the propagator path - it has to have the reference's cost, not its code."""
    import torch.nn as nn

    def conv2(i, o):
        return [nn.Conv2d(i, o, 3, 1, 1), nn.BatchNorm2d(o), nn.LeakyReLU(0.1),
                nn.Conv2d(o, o, 3, 1, 1), nn.BatchNorm2d(o), nn.LeakyReLU(0.1)]

    def down(i, o):
        return nn.Sequential(*conv2(i, o), nn.AvgPool2d(2, 2, ceil_mode=True))

    def up(i, o):
        return nn.Sequential(nn.UpsamplingBilinear2d(scale_factor=2), *conv2(i, o))

    class Generator(nn.Module):
        def __init__(self):
            super().__init__()
            f = [16, 32, 64, 128]
            self.enc = nn.Sequential(down(n_shots, f[0]), down(f[0], f[1]), down(f[1], f[2]), down(f[2], f[3]))
            self.h0, self.w0 = (nz + 1 + 7) // 8, (ny + 7) // 8          # 19 x 25 for 151 x 200
            self.to_latent = nn.LazyLinear(8)
            self.from_latent = nn.Linear(8, f[3] * self.h0 * self.w0)
            self.dec = nn.Sequential(up(f[3], f[2]), up(f[2], f[1]), up(f[1], f[0]))
            self.head = nn.Conv2d(f[0], 1, 1)

        def forward(self, gathers):
            z = self.to_latent(torch.flatten(self.enc(gathers[:, :, 1:4001:4, :]), 1))
            x = self.from_latent(z).view(-1, 128, self.h0, self.w0)
            x = self.head(self.dec(x)[:, :, 1:1 + nz, :ny].contiguous())
            return vmin + torch.sigmoid(x) * (vmax - vmin)

    return Generator()


def run_c5(args, w, dev, rank, world, local):
    """One step = one training iteration of the Auto22 path (models/Auto22_model.py:357-362 +
    models/networks.py:5286-5287, 5445-5493) with the shots sharded over the ranks."""
    import torch.distributed as dist
    import physicsbasedfwi2_b200 as pkg
    from physicsbasedfwi2_b200 import _lib
    from physicsbasedfwi2_b200.distributed import allreduce_gradient, shard_bounds
    from physicsbasedfwi2_b200.misfit import normalize_traces, trace_misfit
    from physicsbasedfwi2_b200.scalar import Propagator
    from physicsbasedfwi2_b200.wavelets import ricker
    nz, ny, nt, S, R = w["nz"], w["ny"], w["nt"], w["survey_shots"], w["n_rec"]
    torch.manual_seed(0)
    gen = make_generator(S, nz, ny, w["vmin"], w["vmax"]).to(dev)
    opt = torch.optim.Adam(gen.parameters(), lr=5e-3)
    lo, hi = shard_bounds(S, rank, world)
    x_s = torch.zeros(S, 1, 2)
    x_s[:, 0, 1] = torch.linspace(0, (ny - 1) * w["dx"], S)
    x_r = torch.zeros(S, R, 2)
    x_r[:, :, 1] = (torch.arange(R).float() * w["dx"])[None]
    src = ricker(w["freq"], nt, w["dt"], 1.0 / w["freq"]).reshape(-1, 1, 1).repeat(1, hi - lo, 1).contiguous().to(dev)
    xs_d, xr_d = x_s[lo:hi].to(dev), x_r[lo:hi].to(dev)
    # --no-vpmax: the reference's own call, Propagator({'vp': v}, dx) without vpmax: the inner step and the PML then follow
    # the model's maximum, read back from the device once per call (as deepwave does)
    kw = dict(pml_width=w["pml"], accuracy=w["order"], vpmax=None if args.no_vpmax else w["vmax"])
    v_true = synth_model(w).to(dev)
    with torch.no_grad():
        observed_all = torch.zeros(nt, S, R, device=dev)      # the data set: all shots, on every rank (network input)
        observed_all[:, lo:hi] = Propagator({"vp": v_true}, w["dx"], **kw)(src, xs_d, xr_d, w["dt"])
        if world > 1:
            dist.all_reduce(observed_all)
        obs_norm = normalize_traces(observed_all[:, lo:hi]).contiguous()
        # direct wave of the constant 1500 m/s model: constant across iterations, computed once
        direct = Propagator({"vp": torch.full_like(v_true, 1500.0)}, w["dx"], **kw)(src, xs_d, xr_d, w["dt"]).contiguous()
    gathers_h = observed_all[None].permute(0, 2, 1, 3).contiguous().cpu().pin_memory()     # [1, S, nt, R]
    gathers_d = gathers_h.to(dev)
    loss_h = torch.empty(1).pin_memory()

    def iteration(gathers):
        opt.zero_grad(set_to_none=True)
        f1 = gen(gathers)                                         # [1, 1, nz, ny]
        v = f1.detach()[0, 0].clamp(w["vmin"], w["vmax"]).requires_grad_(True)
        pred = Propagator({"vp": v}, w["dx"], **kw)(src, xs_d, xr_d, w["dt"])
        loss = trace_misfit(pred, obs_norm, direct, kind="l1") * ((hi - lo) / S)   # mean over ALL traces of the survey
        loss.backward()
        g, l = allreduce_gradient(v.grad, loss)
        f1.backward(g[None, None])                                # models/Auto22_model.py:326
        opt.step()
        return l

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, sampler=None):
        barrier()
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = fn()
        e1.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), clocks, out

    def step_e2e():
        l = iteration(gathers_h.to(dev, non_blocking=True))
        loss_h.copy_(l.reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return l

    for _ in range(args.warmup):
        l = iteration(gathers_d)
    torch.cuda.synchronize()
    if not math.isfinite(float(l.detach())):
        raise RuntimeError("non-finite misfit in the C5 iteration")
    _lib.LOOP_TIMES = []
    n0 = pkg.kernel_launches()
    ms, clocks, _ = timed(lambda: iteration(gathers_d), args.steps, ClockSampler(local))
    launches = pkg.kernel_launches() - n0
    loops, _lib.LOOP_TIMES = _lib.LOOP_TIMES, None
    t_fwd = [a.elapsed_time(b) / n for tag, a, b, n, _ in loops if tag.endswith("_fwd")]
    t_adj = [a.elapsed_time(b) / n for tag, a, b, n, _ in loops if tag.endswith("_adj")]
    ms_fwd, ms_adj = sum(t_fwd) / len(t_fwd), sum(t_adj) / len(t_adj)
    from physicsbasedfwi2_b200 import _host
    nti = nt * _host.inner_step(w["dt"], w["dx"], w["dx"], w["vmax"])[0]
    work = nz * ny * nti * S * 2
    ms_e, _, _ = timed(step_e2e, args.steps)
    if rank == 0:
        peak, peak_src = hbm_peak()
        unit_launch = nz * ny * (hi - lo)
        ach = unit_launch * (B_FWD + B_ADJ) / ((ms_fwd + ms_adj) * 1e-3) / 1e9
        cfg = bench_config(w, hi - lo, world, "c5")
        cfg["workload"] += "; one step = generator forward + propagator gradient (fused L1 trace misfit, cached direct wave) + generator backward + Adam"
        cfg["shots_per_step"] = S
        cfg["l2"] = "wavefields of a launch fit L2 (latency-bound regime of the reference's own grids)"
        cfg["vpmax"] = "from the model, read back every call (the reference's call)" if args.no_vpmax else "passed by the caller"
        line = {"metric": "FWI gradient Gcell-updates/s (fwd+adjoint)", "value": round(work * args.steps / (ms * 1e-3) / 1e9, 3),
                "unit": "Gcell-updates/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
                "roofline": {"bound": "hbm", "kernel": "acoustic_step_kernel (forward MODE 1 + adjoint MODE 2 pair)",
                             "achieved": round(ach, 1), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4),
                             "traffic": None, "peak_source": peak_src,
                             "forward": {"ms_per_launch": round(ms_fwd, 5)}, "adjoint": {"ms_per_launch": round(ms_adj, 5)},
                             "kernel_share_of_step": round((ms_fwd + ms_adj) * nti * args.steps / ms, 4)},
                "clocks": clocks, "gpu_launches": int(launches),
                "e2e": {"value": round(work * args.steps / (ms_e * 1e-3) / 1e9, 3), "unit": "Gcell-updates/s",
                        "h2d_bytes_per_step": gathers_h.numel() * 4, "d2h_bytes_per_step": 4,
                        "ms_per_step": round(ms_e / args.steps, 3)}}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_f3(args, w, dev):
    """One step = the reference's whole inner loop (100 x: clamp, forward, misfit, adjoint, Adam) on one GPU:
    captured CUDA graph replays against the same loop driven from Python with a host sync per iteration."""
    import physicsbasedfwi2_b200 as pkg
    from physicsbasedfwi2_b200 import _host
    from physicsbasedfwi2_b200.inversion import inner_loop_fwi
    from physicsbasedfwi2_b200.scalar import Propagator
    from physicsbasedfwi2_b200.wavelets import ricker
    nz, ny, nt, S, R, E = w["nz"], w["ny"], w["nt"], w["survey_shots"], w["n_rec"], w["inner_loop"]
    x_s = torch.zeros(S, 1, 2)
    x_s[:, 0, 1] = torch.arange(S).float() * (101 * w["dx"] / S)
    x_r = torch.zeros(S, R, 2)
    x_r[:, :, 1] = (torch.arange(R).float() * w["dx"])[None]
    wavelet = ricker(w["freq"], nt, w["dt"], 1.0 / w["freq"]).reshape(-1, 1, 1)
    true = synth_model(w).to(dev)
    with torch.no_grad():
        observed = Propagator({"vp": true}, w["dx"], vpmax=w["vmax"])(wavelet.repeat(1, S, 1).contiguous().to(dev),
                                                                      x_s.to(dev), x_r.to(dev), w["dt"])
    start = torch.full((nz, ny), 2500.0, device=dev)

    def loop(graph):
        return inner_loop_fwi(start, observed, wavelet, x_s, x_r, w["dx"], w["dt"], num_epochs=E, lr=10.0,
                              vmin=w["vmin"], vmax=w["vmax"], kind="mse", graph=graph)

    out = {}
    for graph in (True, False):
        for _ in range(max(1, args.warmup // 3)):
            loop(graph)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            v, losses = loop(graph)
        torch.cuda.synchronize()
        out[graph] = (time.perf_counter() - t0) / args.steps
    nti = nt * _host.inner_step(w["dt"], w["dx"], w["dx"], w["vmax"])[0]
    work = nz * ny * nti * S * 2 * E
    line = {"metric": "FWI gradient Gcell-updates/s (fwd+adjoint)", "value": round(work / out[True] / 1e9, 3),
            "unit": "Gcell-updates/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(1e3 * out[True], 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"f3: inner-loop FWI of models/networks.py:15278-15423: {nz}x{ny}, {S} shots x {R} receivers, "
                                   f"{nt} steps (x{nti // nt} inner), {E} Adam iterations per step; one CUDA graph replayed"},
            "ms_per_iteration": {"cuda_graph": round(1e3 * out[True] / E, 4), "python_loop": round(1e3 * out[False] / E, 4)},
            "first_last_misfit": [losses[0], losses[-1]], "gpu_launches": int(pkg.kernel_launches())}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def split_batches(n_shots, cap):
    """[lo, hi) ranges of the launches one rank makes for n_shots shots, at most `cap` shots per launch: full launches,
    then the rest (32 shots, cap 6 -> 6 6 6 6 6 2: measured on C2, an adjoint launch costs 9.4 us per shot at 6 shots
    and 10.7 at 5, so full launches plus a short one beat an even 6 6 5 5 5 5 split by 3 %)."""
    out, lo = [], 0
    while lo < n_shots:
        hi = min(n_shots, lo + cap)
        out.append((lo, hi))
        lo = hi
    return out


def multi_gpu_check(w, dev, rank, world):
    """Driver-visible multi-GPU correctness, run under torchrun ahead of the timed region (the 2-GPU pytest cases
    skip on a 1-GPU test box).  Shot-parallel: the N-rank gradient (shards + one allreduce) against rank 0
    computing every shot alone.  Slab workloads: the x-slab result against the undecomposed run."""
    import torch.distributed as dist
    from physicsbasedfwi2_b200.distributed import allreduce_gradient, shard_bounds
    from physicsbasedfwi2_b200.scalar import Propagator
    from physicsbasedfwi2_b200.wavelets import ricker
    small = dict(w, nz=160, ny=520 if not w.get("slabs") else 96 * world + 40, nt=240, n_rec=64, rec_stride=4,
                 survey_shots=2 * world)
    if w.get("slabs"):
        small["survey_shots"] = 2
    v = synth_model(small).to(dev)
    S = small["survey_shots"]
    x_s, x_r = synth_survey(small, S)
    src = ricker(small["freq"], small["nt"], small["dt"], 1.0 / small["freq"]).reshape(-1, 1, 1).repeat(1, S, 1).contiguous()
    obs = torch.randn(small["nt"], S, small["n_rec"], generator=torch.Generator().manual_seed(5)).mul_(1e-3)
    x_s, x_r, src, obs = (t.to(dev) for t in (x_s, x_r, src, obs))
    kw = dict(pml_width=small["pml"], accuracy=small["order"], vpmax=small["vmax"])

    def grad(lo, hi, **extra):
        vg = v.clone().requires_grad_(True)
        rec = Propagator({"vp": vg}, small["dx"], **kw, **extra)(src[:, lo:hi], x_s[lo:hi], x_r[lo:hi], small["dt"])
        loss = 0.5 * ((rec - obs[:, lo:hi]) ** 2).sum()
        loss.backward()
        return vg.grad, loss.detach(), rec.detach()

    def rel(a, b):
        return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-300))

    out = {"grid": f"{small['nz']}x{small['ny']}", "shots": S, "steps": small["nt"], "ranks": world}
    if w.get("slabs"):
        g_dd, l_dd, rec_dd = grad(0, S, slabs=True, halo_steps=4)
        g_1, l_1, rec_1 = grad(0, S)
        out.update(mode="x-slab vs undecomposed", traces=rel(rec_dd, rec_1), grad_v=rel(g_dd, g_1),
                   loss=abs(float(l_dd) - float(l_1)) / abs(float(l_1)))
        ok = out["traces"] <= 1e-5 and out["grad_v"] <= 1e-4
    else:
        lo, hi = shard_bounds(S, rank, world)
        g_loc, l_loc, _ = grad(lo, hi)
        g_n, l_n = allreduce_gradient(g_loc, l_loc)
        g_1, l_1, _ = grad(0, S)                 # every rank computes the reference; rank 0's comparison is reported
        out.update(mode="shot-parallel N ranks + one allreduce vs one GPU", grad_v=rel(g_n, g_1),
                   loss=abs(float(l_n) - float(l_1)) / abs(float(l_1)))
        ok = out["grad_v"] <= 1e-6 and out["loss"] <= 1e-6
    flag = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    out["ok"] = bool(flag.item() == 1.0)
    for k in ("traces", "grad_v", "loss"):
        if k in out:
            out[k] = float(f"{out[k]:.3e}")
    return out


def run_gpu(args, w, dev, rank, world, local):
    import torch.distributed as dist
    import physicsbasedfwi2_b200 as pkg
    from physicsbasedfwi2_b200 import _host, _lib
    from physicsbasedfwi2_b200.distributed import allreduce_gradient, shard_bounds
    from physicsbasedfwi2_b200.scalar import Propagator
    from physicsbasedfwi2_b200.wavelets import ricker

    nz, ny, nt = w["nz"], w["ny"], w["nt"]
    elastic, slabs = bool(w.get("elastic")), bool(w.get("slabs"))
    P = w["pml"] + w["order"] // 2
    nti = nt * (1 if elastic else _host.inner_step(w["dt"], w["dx"], w["dx"], w["vmax"])[0])
    # shots per launch: one q snapshot per inner step per shot must fit HBM next to the wavefields
    kq = w.get("save_every", 1)
    per_shot = (nz + 2 * P) * (ny + 2 * P + 64) * 4 * (nti // kq + 8) * (5 if elastic else 1)
    if w.get("ckpt"):
        per_shot = (nz + 2 * P) * (ny + 2 * P + 64) * 4 * (6 * (nti // w["ckpt"] + 2) + w["ckpt"] + 24)
    if slabs:
        cols = (ny + 2 * P) // world + 2 * w["slabs"] * (w["order"] // 2) + 64
        per_shot = (nz + 2 * P) * cols * 4 * (nti + 16)
    free, _ = torch.cuda.mem_get_info(dev)
    # 6 shots per launch is where the C2 kernel pair was tuned (7 fit, but leave the allocator no slack)
    full_nt = WORKLOADS[args.workload]["nt"]              # the cap must not change with a --nt override (ncu runs)
    cap = 16 if (w.get("ckpt") or per_shot * (full_nt / nt) < (2 << 30)) else (6 if kq == 1 else 12)
    cap = args.shots or max(1, min(cap, w["survey_shots"], int((0.6 if w.get("ckpt") else 0.88) * free // per_shot)))

    # shots of one step on this rank.  Default for the shot-parallel acoustic survey (C2): the 1/8 slice of the
    # survey a GPU owns when the 256 shots are sharded over 8 GPUs (32 shots = six launches of 6 6 5 5 5 5), so
    # that the 8-GPU point IS the north star's survey and the smaller N are the same per-GPU work ("weak").
    # --survey: the whole survey sharded over the N ranks ("strong").  Other workloads: one launch per step.
    total = w["survey_shots"]
    if args.survey and not slabs:
        lo, hi = shard_bounds(total, rank, world)
        first, n_rank, scaling = lo, hi - lo, "strong"
    elif args.workload == "c2" and not args.shots and not args.one_batch:
        n_rank = max(1, total // 8)
        first, scaling = (rank * n_rank) % total, "weak"
    else:
        n_rank, first, scaling = cap, (0 if slabs else rank * cap), ("strong" if slabs else "weak")
    batches = split_batches(n_rank, cap)
    shots_step = n_rank if slabs else (total if scaling == "strong" else n_rank * world)

    v_host = synth_model(w).pin_memory()
    x_s_h, x_r_h = synth_survey(w, n_rank, first_shot=first)
    src_h = ricker(w["freq"], nt, w["dt"], 1.0 / w["freq"]).reshape(-1, 1, 1).repeat(1, n_rank, 1).contiguous().pin_memory()
    gen = torch.Generator().manual_seed(2 + (0 if slabs else rank))
    obs_shape = (2, nt, n_rank, w["n_rec"]) if elastic else (nt, n_rank, w["n_rec"])
    obs_h = torch.randn(*obs_shape, generator=gen).mul_(1e-3).pin_memory()   # "observed" data
    x_s_h, x_r_h = x_s_h.pin_memory(), x_r_h.pin_memory()
    grad_h = torch.empty(nz, ny).pin_memory()
    loss_h = torch.empty(1).pin_memory()

    def gradient(v_d, src_d, xs_d, xr_d, obs_d):
        """The call sequence of the reference (models/networks.py:5449,5464,5476,5491) for one launch of shots."""
        v_d = v_d.detach().requires_grad_(True)
        if elastic:
            vs_d = v_d.detach() / math.sqrt(3.0)
            vs_d[:24] = 0.0
            rho_d = 310.0 * v_d.detach() ** 0.25
            prop = pkg.Propagator({"vp": v_d, "vs": vs_d, "rho": rho_d}, w["dx"], pml_width=w["pml"],
                                  accuracy=w["order"], vpmax=w["vmax"], checkpoint_every=w.get("ckpt", 0))
        else:
            prop = Propagator({"vp": v_d}, w["dx"], pml_width=w["pml"], accuracy=w["order"],
                              vpmax=None if args.no_vpmax else w["vmax"], checkpoint_every=w.get("ckpt", 0), save_every=kq, slabs=slabs and world > 1,
                              halo_steps=w.get("slabs") or 8)
        rec = prop(src_d, xs_d, xr_d, w["dt"])
        loss = 0.5 * ((rec - obs_d) ** 2).sum()
        loss.backward()
        return v_d.grad, loss.detach()

    sdim = 2 if elastic else 1          # shot dimension of the observed data
    # per-launch inputs: pinned host copies (the e2e leg uploads them every step) and device-resident ones
    host_b = [tuple(t.narrow(d, lo, hi - lo).contiguous().pin_memory()
                    for t, d in ((src_h, 1), (x_s_h, 0), (x_r_h, 0), (obs_h, sdim))) for lo, hi in batches]
    dev_b = [tuple(t.to(dev) for t in hb) for hb in host_b]
    v_d = v_host.to(dev)

    def step(v_in, per_launch, to_dev):
        """All launches of this rank's shots, gradients accumulated, then ONE allreduce."""
        v_cur = to_dev(v_in)
        g_acc, l_acc = None, None
        for tens in per_launch:
            g, l = gradient(v_cur, *(to_dev(t) for t in tens))
            g_acc = g if g_acc is None else g_acc.add_(g)
            l_acc = l if l_acc is None else l_acc + l
        if slabs:
            return g_acc, l_acc           # slab mode: traces and gradient are already whole on every rank
        return allreduce_gradient(g_acc, l_acc)

    def step_resident():
        return step(v_d, dev_b, lambda t: t)

    def step_e2e():
        g, l = step(v_host, host_b, lambda t: t.to(dev, non_blocking=True))
        grad_h.copy_(g, non_blocking=True)
        loss_h.copy_(l.reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return g, l

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, sampler=None):
        barrier()
        if sampler:
            sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = fn()
        e1.record()
        barrier()
        clocks = sampler.stop() if sampler else None
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), clocks, out

    mgc = None
    if world > 1 and not elastic and not args.no_check:
        mgc = multi_gpu_check(w, dev, rank, world)

    for _ in range(args.warmup):
        g, l = step_resident()
    torch.cuda.synchronize()
    if not bool(torch.isfinite(g).all()) or not math.isfinite(float(l.detach())):
        raise RuntimeError("non-finite gradient / misfit in the benchmark step")

    _lib.LOOP_TIMES = []
    n0 = pkg.kernel_launches()
    ms, clocks, _ = timed(step_resident, args.steps, ClockSampler(local))
    launches = pkg.kernel_launches() - n0
    loops = _lib.LOOP_TIMES
    _lib.LOOP_TIMES = None

    cells = nz * ny
    work = cells * nti * shots_step * 2                 # cell-updates (inner steps) per step, whole job
    value = work * args.steps / (ms * 1e-3) / 1e9

    e2e = None
    if not args.no_e2e:
        e_steps = args.steps if ms / args.steps < 2500 else max(1, min(args.steps, 3))   # long steps: 3 e2e repeats
        step_e2e()
        ms_e, _, _ = timed(step_e2e, e_steps)
        h2d = v_host.numel() * 4 + sum(t.numel() * t.element_size() for hb in host_b for t in hb)
        d2h = grad_h.numel() * 4 + 4
        e2e = {"value": round(work * e_steps / (ms_e * 1e-3) / 1e9, 3), "unit": "Gcell-updates/s",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e_steps,
               "ms_per_step": round(ms_e / e_steps, 3)}

    # ---- roofline from the native time loops of the timed region (CUDA events on the launching stream) ----
    # per pass: algorithmic bytes per cell-update: forward + snapshot 20 B, adjoint + imaging 28 B,
    # forward only (checkpoint pass, MODE 0) 16 B; elastic 80 / 120 B
    bytes_of = {"acoustic_fwd": 16.0 + 4.0 / kq, "acoustic_fwd0": 16.0, "acoustic_adj": 16.0 + 12.0 / kq,
                "elastic_fwd": B_FWD_EL, "elastic_adj": B_ADJ_EL}
    div = world if slabs else 1                          # a slab launch advances 1/world of each shot
    agg = {}
    for tag, a, b, n, shots in loops:
        t = agg.setdefault(tag, {"ms": 0.0, "launches": 0, "bytes": 0.0, "cell_updates": 0.0})
        t["ms"] += a.elapsed_time(b)
        t["launches"] += n
        t["cell_updates"] += float(cells) * shots * n / div
        t["bytes"] += float(cells) * shots * n / div * bytes_of[tag]

    if rank == 0:
        peak, peak_src = hbm_peak()
        tot_ms = sum(t["ms"] for t in agg.values())
        tot_bytes = sum(t["bytes"] for t in agg.values())
        ach = tot_bytes / (tot_ms * 1e-3) / 1e9
        kname = ("elastic step (f1 + f2 forward, b2 + b1 fused adjoint; one time step = one 'launch')"
                 if elastic else "acoustic_step_kernel (forward: MODE 1 per L2-sized shot group; adjoint: MODE 0 advance-only + MODE 3 "
                 "paired-imaging launches" + (" + forward-only MODE 0 of the checkpoint pass)" if "acoustic_fwd0" in agg else ")"))
        roof = {"bound": "hbm", "kernel": kname, "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
                "frac": round(ach / peak, 4), "algorithmic_frac": round(ach / peak, 4), "dram_frac": None,
                "traffic": None, "peak_source": peak_src,
                "algorithmic_bytes_per_cell_update": {k: bytes_of[k] for k in agg},
                "kernel_share_of_step": round(tot_ms / ms, 4)}
        names = {"acoustic_fwd": "forward", "acoustic_fwd0": "forward_only", "acoustic_adj": "adjoint",
                 "elastic_fwd": "forward", "elastic_adj": "adjoint"}
        for tag, t in agg.items():
            a_k = t["bytes"] / (t["ms"] * 1e-3) / 1e9
            roof[names[tag]] = {"ms_per_launch": round(t["ms"] / t["launches"], 5), "launches": t["launches"],
                                "achieved": round(a_k, 1), "frac": round(a_k / peak, 4)}
        # measured DRAM traffic (ncu --set full, one launch of each kernel at the 6-shot launch size; profiles/)
        prof = os.path.join(ROOT, "profiles", "traffic.json")
        tr = None
        if os.path.exists(prof):
            try:
                tr = json.load(open(prof)).get(args.workload)
            except Exception:
                tr = None
        if tr:
            roof["traffic"] = tr
            fwd_k, adj_k = ("elastic_fwd", "elastic_adj") if elastic else ("acoustic_fwd", "acoustic_adj")
            if fwd_k in agg and adj_k in agg and len(agg) == 2:
                # measured DRAM bytes per cell-update of each pass (one ncu capture per kernel) x the cell-updates timed here
                if "forward_bytes_per_cell_update" in tr:
                    bf_m, ba_m = tr["forward_bytes_per_cell_update"], tr["adjoint_bytes_per_cell_update"]
                else:
                    s_ref = float(tr.get("shots_per_launch", 5 if elastic else 6))
                    bf_m = tr["forward_bytes_per_launch"] / (cells * s_ref)
                    ba_m = tr["adjoint_bytes_per_launch"] / (cells * s_ref)
                dram = bf_m * agg[fwd_k]["cell_updates"] + ba_m * agg[adj_k]["cell_updates"]
                roof["dram_frac"] = round(dram / (tot_ms * 1e-3) / 1e9 / peak, 4)
        cfg = bench_config(w, cap, world, args.workload)
        cfg.update(shots_per_gpu_per_step=n_rank, launches_per_gpu_per_step=[hi - lo for lo, hi in batches],
                   shots_per_step=shots_step, first_shot_of_rank0=first,
                   survey=("every rank works on the same shots, one column slab each" if slabs else
                           "whole survey sharded over the ranks, one allreduce" if scaling == "strong" else
                           f"{n_rank} shots per GPU per step, one allreduce per step"
                           + (" (the per-GPU share of the 256-shot survey on 8 GPUs)" if args.workload == "c2" and n_rank == total // 8 else "")))
        line = {"metric": "FWI gradient Gcell-updates/s (fwd+adjoint)", "value": round(value, 3),
                "unit": "Gcell-updates/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True, "scaling": scaling,
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg, "roofline": roof,
                "clocks": clocks, "gpu_launches": int(launches), "e2e": e2e}
        if mgc is not None:
            line["multi_gpu_check"] = mgc
        if world == 1 and not args.no_cpu:
            if elastic:
                line["cpu_baseline"] = cpu_baseline_elastic(w)
            else:
                import tempfile
                dump = os.path.join(tempfile.gettempdir(), f"fwi_bench_cpu_{os.getpid()}.npz")
                line["cpu_baseline"] = cpu_baseline(w, dump=dump)
                line["parity"] = parity_block(w, dump, dev)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def parity_block(w, dump, dev):
    """The CPU sample the cpu_baseline leg just ran (one shot, n steps; traces and model gradient kept in `dump`)
    recomputed on the GPU through the public API: rel-L2 of traces and gradient (north star: 1e-5 / 1e-4)."""
    import numpy as np
    from physicsbasedfwi2_b200.scalar import Propagator
    z = np.load(dump)
    n = int(z["n"])
    x_s, x_r, src, obs = (t.to(dev) for t in cpu_sample_inputs(w, n))
    vg = synth_model(w).to(dev).requires_grad_(True)
    rec = Propagator({"vp": vg}, w["dx"], pml_width=w["pml"], accuracy=w["order"], vpmax=w["vmax"])(src, x_s, x_r, w["dt"])
    (0.5 * ((rec - obs) ** 2).sum()).backward()

    def rel(a, b):
        a, b = a.detach().double().cpu(), torch.from_numpy(b).double()
        return float(f"{float((a - b).norm() / b.norm()):.3e}")
    out = {"traces": rel(rec, z["rec"]), "grad_v": rel(vg.grad, z["grad_v"]),
           "sample": f"1 shot x {n} steps of this workload, GPU (public API) vs oracle/cpu_ref.c fp32",
           "tolerance": {"traces": 1e-5, "grad_v": 1e-4}}
    out["ok"] = out["traces"] <= 1e-5 and out["grad_v"] <= 1e-4
    try:
        os.remove(dump)
    except OSError:
        pass
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--shots", type=int, default=0, help="shots per GPU per step (0 = fit HBM, max 6)")
    ap.add_argument("--nt", type=int, default=0, help="override the number of time steps (tuning only)")
    ap.add_argument("--halo-steps", type=int, default=0, help="slab workloads: steps between ghost exchanges (tuning only)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer end-to-end leg")
    ap.add_argument("--pml", type=int, default=-1, help="override the PML width (tuning only: 0 = PML-free grid)")
    ap.add_argument("--no-vpmax", action="store_true", help="call the propagator as the reference does, without vpmax (one device-to-host read per call)")
    ap.add_argument("--save-every", type=int, default=1, help="snapshot decimation k (performance mode; k = 1 is the parity-graded default)")
    ap.add_argument("--survey", action="store_true", help="one step = the WHOLE survey sharded over the ranks (strong scaling)")
    ap.add_argument("--one-batch", action="store_true", help="c2: one launch of shots per step instead of the 32-shot slice")
    ap.add_argument("--no-check", action="store_true", help="skip the multi-GPU correctness check ahead of the timed region")
    ap.add_argument("--dump", default=None, help="--impl reference: .npz receiving the CPU sample's traces / gradient")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl != "reference":
        args.warmup = 3
    w = dict(WORKLOADS[args.workload])
    w["name"] = args.workload
    if args.nt:
        w["nt"] = args.nt
    if args.pml >= 0:
        w["pml"] = args.pml
    w["save_every"] = max(1, args.save_every)

    if args.impl == "reference":
        run_reference(args, w)
        return

    import torch.distributed as dist
    import physicsbasedfwi2_b200 as pkg
    from physicsbasedfwi2_b200 import _lib
    from physicsbasedfwi2_b200.distributed import allreduce_gradient
    if w.get("slabs"):
        # ghost-exchange interval: --halo-steps, else the package's own choice for this slab width
        from physicsbasedfwi2_b200.slab import auto_halo_steps
        world_env = int(os.environ.get("WORLD_SIZE", "1"))
        w["slabs"] = args.halo_steps or auto_halo_steps(w["ny"] + 2 * (w["pml"] + w["order"] // 2), w["order"], world_env)
    from physicsbasedfwi2_b200.scalar import Propagator
    from physicsbasedfwi2_b200.wavelets import ricker

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    L = pkg.lib()
    if L.fwi_device_ok(local) != 0:
        raise RuntimeError(L.fwi_last_error().decode())

    if w.get("e2e_iteration"):
        return run_c5(args, w, dev, rank, world, local)
    if w.get("inner_loop"):
        return run_f3(args, w, dev)

    return run_gpu(args, w, dev, rank, world, local)


if __name__ == "__main__":
    main()
