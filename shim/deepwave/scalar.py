"""deepwave.scalar — re-export (reference call sites models/networks.py:5408,5449,5464)."""
from physicsbasedfwi2_b200.scalar import Propagator  # noqa: F401
