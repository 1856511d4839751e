"""fleetx_b200 — B200-native GPT hybrid-parallel decoder layer (PaddleFleetX hot path).

Python here is the host-side mirror of the reference's layer API; all arithmetic runs in
libfleetx_b200.so (hand-written sm_100a CUDA behind the C ABI in include/fleetx_b200.h).
"""
from ._lib import FleetXError, lib, launch_count  # noqa: F401

__version__ = "0.1.0"
