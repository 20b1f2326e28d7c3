"""physicsbasedfwi2_b200 — B200-native finite-difference wave propagator for FWI.

Drop-in for the propagator path of ADharaUTEXAS123007/PhysicsBasedFWI2 (the deepwave-style
call ``Propagator({'vp'[, 'vs', 'rho']}, dx)(source_amplitudes, x_s, x_r, dt)``; reference
call sites models/networks.py:5408-5411, :5449, :5464, :5491).  The hot path is hand-written
sm_100a CUDA behind the C ABI in include/fwi_b200.h; there is no CPU fallback.
"""
from . import wavelets  # noqa: F401
from . import scalar  # noqa: F401
from . import elastic  # noqa: F401
from . import misfit  # noqa: F401
from . import slab  # noqa: F401
from . import fwiloss  # noqa: F401
from . import inversion  # noqa: F401
from . import denise_glue  # noqa: F401
from ._lib import build, kernel_launches, lib  # noqa: F401


def Propagator(model: dict, dx, *args, **kwargs):
    """``Propagator({'vp'[, 'vs', 'rho']}, dx)``: the scalar (acoustic) propagator for a ``{'vp'}``
    model, the elastic one when ``'vs'`` and ``'rho'`` are given too."""
    if isinstance(model, dict) and ("vs" in model or "rho" in model):
        return elastic.Propagator(model, dx, *args, **kwargs)
    return scalar.Propagator(model, dx, *args, **kwargs)


__all__ = ["scalar", "elastic", "misfit", "slab", "fwiloss", "inversion", "denise_glue", "wavelets", "Propagator", "build", "lib", "kernel_launches"]
