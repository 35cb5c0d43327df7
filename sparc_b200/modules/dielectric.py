"""Import-path compatibility with the reference (`wedm.modules.dielectric`)."""
from . import DielectricModule, DielectricModuleParameters  # noqa: F401
