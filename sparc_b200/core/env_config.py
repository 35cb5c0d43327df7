"""`wedm.core.env_config` compatibility."""
from ..params import EnvironmentConfig  # noqa: F401
