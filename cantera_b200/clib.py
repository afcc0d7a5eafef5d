"""ctypes binding of the C ABI in include/ctb_clib.h (cantera_b200/libctb.so).

This is the same stub a maintainer of the reference would add beside interfaces/clib (see
INTEGRATION.md). There is no CPU fallback: if the library is missing it is built with nvcc, and
if that fails, or no CUDA device is usable, constructors raise.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# CTB_LIB: an alternative build of the same library (kernel experiments); the default is in-tree
LIB_PATH = os.environ.get("CTB_LIB") or os.path.join(_HERE, "libctb.so")
_LIB = None

ERR = -999
DERR = -999.999
HOST, DEVICE = 0, 1
NSTATS = 8
NOALLOC_MAX_STATES = 1 << 22      # CTB_NOALLOC_MAX_STATES
NOALLOC_SCRATCH_BYTES = 32 << 20  # CTB_NOALLOC_SCRATCH_BYTES

_dp = C.c_void_p  # host or device pointer, by memspace
_i32, _i64, _f64 = C.c_int32, C.c_int64, C.c_double

_STATE = [_i32, _i64, _dp, _dp, _dp, _i64]  # handle, nStates, T, P, Y, ldY

#: every symbol include/ctb_clib.h declares -> (restype, argtypes)
SIGNATURES = {
    "ctb_getError": (_i32, [_i32, C.c_char_p]),
    "ctb_setError": (_i32, [C.c_char_p]),
    "ctb_nDevices": (_i32, []),
    "ctb_version": (_i32, [_i32, C.c_char_p]),
    "ctb_fp64PeakTflops": (_f64, [_i32]),
    "ctb_fp64DmmaPeakTflops": (_f64, [_i32]),
    "ctb_kin_newFromTables": (_i32, [C.c_void_p, _i32]),
    "ctb_kin_newFromFile": (_i32, [C.c_char_p, _i32]),
    "ctb_kin_del": (_i32, [_i32]),
    "ctb_kin_setStream": (_i32, [_i32, C.c_void_p]),
    "ctb_kin_synchronize": (_i32, [_i32]),
    "ctb_kin_nSpecies": (_i32, [_i32]),
    "ctb_kin_nReactions": (_i32, [_i32]),
    "ctb_kin_setMultiplier": (_i32, [_i32, _i32, _f64]),
    "ctb_kin_multiplier": (_f64, [_i32, _i32]),
    "ctb_kin_setDerivativeSettings": (_i32, [_i32, _i32, _i32, _f64]),
    "ctb_kin_getNetProductionRates": (_i32, _STATE + [_dp, _i64, _i32]),
    "ctb_kin_getCreationRates": (_i32, _STATE + [_dp, _i64, _i32]),
    "ctb_kin_getDestructionRates": (_i32, _STATE + [_dp, _i64, _i32]),
    "ctb_kin_getFwdRatesOfProgress": (_i32, _STATE + [_dp, _i64, _i32]),
    "ctb_kin_getRevRatesOfProgress": (_i32, _STATE + [_dp, _i64, _i32]),
    "ctb_kin_getNetRatesOfProgress": (_i32, _STATE + [_dp, _i64, _i32]),
    "ctb_kin_getFwdRateConstants": (_i32, _STATE + [_dp, _i64, _i32]),
    "ctb_kin_getEquilibriumConstants": (_i32, _STATE + [_dp, _i64, _i32]),
    "ctb_kin_getQuantity": (_i32, [_i32, _i32, _i64, _dp, _dp, _dp, _i64, _dp, _i64, _i32]),
    "ctb_kin_getNetProductionRates_ddX": (_i32, _STATE + [_dp, _i64, _i32]),
    "ctb_kin_getNetProductionRates_ddCi": (_i32, _STATE + [_dp, _i64, _i32]),
    "ctb_kin_getNetProductionRates_ddX_noAlloc": (_i32, _STATE + [_dp, _i64]),
    "ctb_trans_newFromArrays": (_i32, [_i32, _dp, _dp, _dp, _dp, _i32]),
    "ctb_trans_del": (_i32, [_i32]),
    "ctb_trans_nSpecies": (_i32, [_i32]),
    "ctb_trans_setStream": (_i32, [_i32, C.c_void_p]),
    "ctb_trans_synchronize": (_i32, [_i32]),
    "ctb_trans_getMixtureTransport": (_i32, _STATE + [_dp, _dp, _dp, _i64, _i32]),
    "ctb_reactornet_new": (_i32, [_i32, _i64]),
    "ctb_reactornet_newTyped": (_i32, [_i32, _i64, _i32]),
    "ctb_reactornet_nEquations": (_i32, [_i32]),
    "ctb_reactornet_del": (_i32, [_i32]),
    "ctb_reactornet_setState": (_i32, [_i32, _dp, _dp, _dp, _i64, _i32]),
    "ctb_reactornet_setRelativeTolerance": (_i32, [_i32, _f64]),
    "ctb_reactornet_setAbsoluteTolerance": (_i32, [_i32, _f64]),
    "ctb_reactornet_setMaxSteps": (_i32, [_i32, _i32]),
    "ctb_reactornet_getTolerances": (_i32, [_i32, _i32, _dp]),
    "ctb_reactornet_setInitialTime": (_i32, [_i32, _f64]),
    "ctb_reactornet_advance": (_i32, [_i32, _f64]),
    "ctb_reactornet_time": (_f64, [_i32]),
    "ctb_reactornet_getState": (_i32, [_i32, _dp, _dp, _i64, _i32]),
    "ctb_reactornet_getStateVector": (_i32, [_i32, _dp, _i64, _i32]),
    "ctb_reactornet_eval": (_i32, [_i32, _dp, _i64, _i32]),
    "ctb_reactornet_getJacobian": (_i32, [_i32, _dp, _i64, _i32]),
    "ctb_reactornet_lastRounds": (_i32, [_i32]),
    "ctb_reactornet_status": (_i32, [_i32, _i64, C.c_void_p]),
    "ctb_reactornet_solverStats": (_i32, [_i32, _i64, C.c_void_p]),
    "ctb_reactornet_setIgnitionMonitor": (_i32, [_i32, _f64]),
    "ctb_reactornet_ignitionDelay": (_i32, [_i32, _i64, _dp]),
}


class CtbError(RuntimeError):
    """The reference's CanteraError as seen through the C ABI (return code -1 / ERR / DERR)."""


def lib():
    """Load (building first if necessary) libctb.so and declare every prototype."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            from . import build as _build
            _build.build()
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


#: include/ctb_cantera.h — the optional reference bridge (libctb_cantera.so, built where the reference is)
BRIDGE_PATH = os.path.join(_HERE, "libctb_cantera.so")
BRIDGE_SIGNATURES = {
    "ctb_kin_new": (_i32, [_i32, C.c_char_p, C.c_char_p, _i32]),
    "ctb_kin_newFromSolution": (_i32, [C.c_void_p, _i32]),
    "ctb_addDataDirectory": (_i32, [C.c_char_p]),
    "ctb_trans_new": (_i32, [C.c_char_p, C.c_char_p, _i32]),
    "ctb_trans_newFromSolution": (_i32, [C.c_void_p, _i32]),
}
_BRIDGE = None


def bridge():
    """The reference bridge, or None if it was not built (it needs the reference at build time)."""
    global _BRIDGE
    if _BRIDGE is None and os.path.exists(BRIDGE_PATH):
        lib()  # libctb.so first: the bridge resolves ctb_kin_newFromTables / ctb_setError against it
        B = C.CDLL(BRIDGE_PATH, mode=C.RTLD_GLOBAL)
        for name, (res, args) in BRIDGE_SIGNATURES.items():
            fn = getattr(B, name)
            fn.restype = res
            fn.argtypes = args
        _BRIDGE = B
    return _BRIDGE


def last_error():
    L = lib()
    n = L.ctb_getError(0, None)
    buf = C.create_string_buffer(max(n, 1))
    L.ctb_getError(n, buf)
    return buf.value.decode(errors="replace")


def check(rc):
    """Raise CtbError for the error codes of clib_defs.h:27-33, otherwise pass the value through."""
    if isinstance(rc, float):
        if rc == DERR:
            raise CtbError(last_error())
        return rc
    if rc == -1 or rc == ERR:
        raise CtbError(last_error())
    return rc
