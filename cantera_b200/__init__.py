"""cantera_b200 — B200-native (sm_100a, FP64) batched ideal-gas chemistry engine: a drop-in for the
reference's Kinetics / ReactorNet hot path over many independent states. See DESIGN.md."""
from .batched import BatchedGasKinetics, BatchReactorNet, mechanism_path  # noqa: F401
from .clib import CtbError  # noqa: F401
from .yaml_mech import MechanismError, tables_from_yaml  # noqa: F401

__version__ = "0.1.0"
from .solution_array import BatchedSolutionArray, MultiGpuKinetics, MultiGpuReactorNet  # noqa: F401,E402
from .transport import BatchedMixTransport  # noqa: F401,E402
