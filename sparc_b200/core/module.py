"""`wedm.core.module` compatibility."""
from ..modules import EDMModule  # noqa: F401
