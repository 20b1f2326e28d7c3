"""`deepwave` (v0.0.x class API) served by the B200 propagator.

Put this directory on PYTHONPATH and the reference's `import deepwave` (models/networks.py:10)
resolves here: `deepwave.scalar.Propagator` (models/networks.py:5408,5449) and
`deepwave.wavelets.ricker` (models/networks.py:5357) are the B200 implementations.
"""
from physicsbasedfwi2_b200 import scalar, wavelets  # noqa: F401

__version__ = "0.0.9+b200"
__all__ = ["scalar", "wavelets"]
