"""Import-path compatibility with the reference (`wedm.modules.wire`)."""
from . import WireModule, WireModuleParameters  # noqa: F401
