"""sparc_b200 — B200-native batched implementation of geduardo/SPARC's `WireEDMEnv.step`.

Public surface mirrors the reference package (`wedm/__init__.py`): the same names import from
here, plus the vectorised `BatchedWireEDMEnv` and the low-level `Engine`.  Importing the package
does not need a GPU; constructing an environment does (there is no CPU fallback).
"""
from .params import (DielectricModuleParameters, EnvironmentConfig, IgnitionModuleParameters,
                     MaterialDatabase, MaterialModuleParameters, MechanicsModuleParameters, WireMaterial,
                     WireModuleParameters, derive_params, get_material_db)
from .state import DeviceStateView, EDMState, crater_dtype
from .modules import (DielectricModule, EDMModule, IgnitionModule, MaterialRemovalModule, MechanicsModule,
                      WireModule)

__version__ = "0.1.0"


def __getattr__(name):
    # envs/engine import torch; keep `import sparc_b200` light
    if name in ("WireEDMEnv", "BatchedWireEDMEnv", "GroupedWireEDMEnv", "WireEDMVectorEnv"):
        from . import envs

        return getattr(envs, name)
    if name in ("GapController", "VoltageController"):
        from . import controllers

        return getattr(controllers, name)
    if name == "Engine":
        from .engine import Engine

        return Engine
    raise AttributeError(name)


__all__ = [
    "EDMState", "DeviceStateView", "crater_dtype", "EnvironmentConfig", "MaterialDatabase", "WireMaterial",
    "get_material_db", "WireEDMEnv", "BatchedWireEDMEnv", "GroupedWireEDMEnv", "WireEDMVectorEnv", "GapController",
    "VoltageController", "Engine", "EDMModule",
    "IgnitionModule", "IgnitionModuleParameters", "WireModule", "WireModuleParameters",
    "MaterialRemovalModule", "MaterialModuleParameters", "DielectricModule", "DielectricModuleParameters",
    "MechanicsModule", "MechanicsModuleParameters", "derive_params",
]
