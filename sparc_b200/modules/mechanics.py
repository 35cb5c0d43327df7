"""Import-path compatibility with the reference (`wedm.modules.mechanics`)."""
from . import MechanicsModule, MechanicsModuleParameters  # noqa: F401
