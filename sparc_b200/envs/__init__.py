from .batched import BatchedWireEDMEnv  # noqa: F401
from .grouped import GroupedWireEDMEnv  # noqa: F401
from .vector import WireEDMVectorEnv  # noqa: F401
from .wire_edm import WireEDMEnv  # noqa: F401

__all__ = ["WireEDMEnv", "BatchedWireEDMEnv", "GroupedWireEDMEnv", "WireEDMVectorEnv"]
