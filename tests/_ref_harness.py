"""Runs the reference's OWN `prop` glue (models/networks.py, unmodified, imported from /root/reference) on the CPU.

TEST INFRASTRUCTURE.  The reference's `import deepwave` (models/networks.py:10) is answered here by a stand-in
with deepwave's v0.0.x class API whose time loops are the C oracle (oracle/cpu_ref.c) behind a
torch.autograd.Function - the reference cannot travel to the GPU box, and this container has no GPU, so the
reference code and the CUDA product never meet in one process.  What this harness is for:

* tests/golden/make_prop_golden.py runs `AutoMarmousi22_Net.prop` / `Auto_Net.prop` through it and commits what the
  reference code computed (loss, masked gradient, subsampled traces) as fixtures;
* tests/test_reference_prop_cpu.py re-runs the small case here (when /root/reference exists) against the fixture;
* tests/test_dropin_gpu.py replays the same inputs through shim/deepwave (the product) on the B200 and must
  reproduce the fixtures.

Only the things the reference hard-codes for its author's machine are redirected: `torch.cuda.set_device(4)` /
`torch.device('cuda:4')` (models/networks.py:5316-5318) become no-ops / the CPU, and the third-party imports that
are absent here and unrelated to the path (pyapi_denise, FrEIA) are empty modules.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

REFERENCE = "/root/reference"


def available() -> bool:
    return os.path.exists(os.path.join(REFERENCE, "models", "networks.py"))


# ---- deepwave v0.0.x stand-in on the C oracle ------------------------------------------------------------------
class _OracleScalar(torch.autograd.Function):
    @staticmethod
    def forward(ctx, v, source_amplitudes, x_s, x_r, dt, dx, log):
        from oracle import cpu_ref as O
        rec, state = O.acoustic_forward(v.detach().cpu(), dx, source_amplitudes.detach().cpu(), x_s.cpu(), x_r.cpu(),
                                        float(dt), order=4, pml_width=10, dtype=torch.float32, save=True)
        ctx.state, ctx.src = state, source_amplitudes.detach().cpu()
        if log is not None:
            log.append(dict(v=v.detach().clone(), src_shape=tuple(source_amplitudes.shape), x_s=x_s.detach().clone(),
                            x_r=x_r.detach().clone(), dt=float(dt), dx=dx, rec=rec.clone(),
                            contiguous=(x_s.is_contiguous(), x_r.is_contiguous())))
        return rec

    @staticmethod
    def backward(ctx, grad_rec):
        from oracle import cpu_ref as O
        gv, gs = O.acoustic_adjoint(ctx.state, grad_rec.detach().cpu().contiguous(), ctx.src, torch.float32)
        ctx.state = None
        return gv, gs, None, None, None, None, None


def oracle_deepwave(log=None) -> types.ModuleType:
    """A module object that quacks like deepwave v0.0.x for the calls the reference makes:
    deepwave.scalar.Propagator({'vp': v}, dx)(src, x_s, x_r, dt) and deepwave.wavelets.ricker(f, n, dt, t0).
    Every forward call is appended to `log` (arguments as the reference passed them + the traces)."""
    from oracle import torch_ref as T

    class Propagator(torch.nn.Module):
        def __init__(self, model, dx, pml_width=None, survey_pad=None, vpmax=None):
            super().__init__()
            self.model, self.dx = model, dx

        def forward(self, source_amplitudes, source_locations, receiver_locations, dt):
            return _OracleScalar.apply(self.model["vp"], source_amplitudes, source_locations, receiver_locations, dt,
                                       self.dx, log)

    dw = types.ModuleType("deepwave")
    dw.scalar = types.ModuleType("deepwave.scalar")
    dw.scalar.Propagator = Propagator
    dw.wavelets = types.ModuleType("deepwave.wavelets")
    dw.wavelets.ricker = lambda freq, length, dt, peak_time, dtype=None: T.ricker(freq, length, dt, peak_time).float()
    return dw


# ---- importing the reference unmodified ------------------------------------------------------------------------
class _CudaProxy:
    """torch.cuda for the reference module: device pinning is a no-op, everything else is the real thing."""

    def __init__(self, real):
        self._real = real

    def set_device(self, *_a, **_k):
        return None

    def __getattr__(self, name):
        return getattr(self._real, name)


class _TorchProxy:
    """The name `torch` inside models/networks.py: torch.device('cuda:N') -> the harness device,
    torch.cuda.set_device -> no-op (models/networks.py:5316-5318 pin the author's GPU 4)."""

    def __init__(self, device):
        self._device = torch.device(device)
        self.cuda = _CudaProxy(torch.cuda)

    def device(self, *_a, **_k):
        return self._device

    def __getattr__(self, name):
        return getattr(torch, name)


def load_networks(deepwave_module, device="cpu"):
    """Imports /root/reference/models/networks.py unmodified with `deepwave_module` answering `import deepwave`."""
    if not available():
        raise RuntimeError("/root/reference is not present")
    import scipy
    if not hasattr(scipy, "arange"):
        scipy.arange = np.arange                      # `from scipy import arange` (models/networks.py:38): removed upstream
    for name in ("pyapi_denise", "FrEIA", "FrEIA.framework", "FrEIA.modules"):
        sys.modules.setdefault(name, types.ModuleType(name))
    saved = {k: sys.modules.get(k) for k in ("deepwave", "deepwave.scalar", "deepwave.wavelets")}
    sys.modules["deepwave"] = deepwave_module
    sys.modules["deepwave.scalar"] = deepwave_module.scalar
    sys.modules["deepwave.wavelets"] = deepwave_module.wavelets
    if REFERENCE not in sys.path:
        sys.path.insert(0, REFERENCE)
    for k in [k for k in sys.modules if k == "models" or k.startswith("models.")]:
        del sys.modules[k]
    try:
        import importlib
        N = importlib.import_module("models.networks")
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    N.deepwave = deepwave_module
    N.torch = _TorchProxy(device)
    return N


def load_lbfgs():
    """functions/LBFGS.py of the reference (needs only torch), imported unmodified."""
    import importlib.util
    sys.modules.setdefault("matplotlib", types.ModuleType("matplotlib"))          # imported for plots only; absent here
    sys.modules.setdefault("matplotlib.pyplot", types.ModuleType("matplotlib.pyplot"))
    spec = importlib.util.spec_from_file_location("ref_functions_LBFGS", os.path.join(REFERENCE, "functions", "LBFGS.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
