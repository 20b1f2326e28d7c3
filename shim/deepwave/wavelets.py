"""deepwave.wavelets — re-export (reference call site models/networks.py:5357)."""
from physicsbasedfwi2_b200.wavelets import ricker  # noqa: F401
