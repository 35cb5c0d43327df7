"""`wedm.core.material_db` compatibility."""
from ..params import MaterialDatabase, WireMaterial, get_material_db  # noqa: F401
