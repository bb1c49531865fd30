"""pcoll_b200 -- B200-native collision-probability hot path behind the mcsimplugin command interface.

Host side mirrors the reference's plugin interface (mcsimplugin/mcsimplugin.cpp:13-44): `MCModule`
with `SendCommand("<name> <args>")`.  Compute lives in csrc/ (sm_100a kernels behind a C ABI).
The directory name carries hyphens (it is the one the build contract names); import it as
`pcoll_b200` through the shim module at the repository root.
"""
from . import cuda_abi  # noqa: F401
from . import plugin  # noqa: F401
from . import inputs  # noqa: F401
from . import sharding  # noqa: F401
from . import planner  # noqa: F401
from .inputs import default_params, load_plan_json, plan_from_pickles, py2_str  # noqa: F401
from .plugin import MCModule, McsimError, push_parameters  # noqa: F401

__all__ = ["cuda_abi", "plugin", "MCModule", "McsimError", "push_parameters", "default_params", "load_plan_json",
           "plan_from_pickles", "py2_str"]
