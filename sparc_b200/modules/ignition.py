"""Import-path compatibility with the reference (`wedm.modules.ignition`)."""
from . import IgnitionModule, IgnitionModuleParameters  # noqa: F401
