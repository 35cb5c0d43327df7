"""`wedm.core.state` compatibility."""
from ..state import EDMState, crater_dtype  # noqa: F401
