"""Import-path compatibility with the reference's `wedm.core` package."""
from ..modules import EDMModule  # noqa: F401
from ..params import EnvironmentConfig, MaterialDatabase, WireMaterial, get_material_db  # noqa: F401
from ..state import EDMState, crater_dtype  # noqa: F401

__all__ = ["EDMState", "crater_dtype", "EnvironmentConfig", "MaterialDatabase", "WireMaterial",
           "get_material_db", "EDMModule"]
