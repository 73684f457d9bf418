"""TEST INFRASTRUCTURE -- nmigen.sim subset (see __init__.py)."""
from ._sim import Simulator, Delay, Settle, Tick, Passive, Active  # noqa: F401

__all__ = ["Simulator", "Delay", "Settle", "Tick", "Passive", "Active"]
