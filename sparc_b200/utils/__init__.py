"""Host-side helpers around the batched engine (the reference's `wedm.utils`)."""
from .logger import BatchedSimulationLogger, LoggerConfig, SimulationLogger  # noqa: F401
