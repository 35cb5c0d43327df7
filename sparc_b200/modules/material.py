"""Import-path compatibility with the reference (`wedm.modules.material`)."""
from . import MaterialRemovalModule, MaterialModuleParameters  # noqa: F401
