"""TEST INFRASTRUCTURE — ctypes front end of oracle/libctb_oracle.so (the plain-C restatement of
the reference's algorithm, oracle/ctb_oracle.c). Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module; nothing under
cantera_b200/ does."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_dp = C.POINTER(C.c_double)


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libctb_oracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        L.ctbo_load.restype = C.c_void_p
        L.ctbo_load.argtypes = [C.c_char_p]
        L.ctbo_free.argtypes = [C.c_void_p]
        L.ctbo_nSpecies.argtypes = [C.c_void_p]
        L.ctbo_nReactions.argtypes = [C.c_void_p]
        L.ctbo_setMultiplier.argtypes = [C.c_void_p, C.c_int32, C.c_double]
        L.ctbo_rates_batch.argtypes = [C.c_void_p, C.c_int64, _dp, _dp, _dp, C.c_int64] + [_dp] * 8
        L.ctbo_ddX.argtypes = [C.c_void_p, C.c_double, C.c_double, _dp, C.c_int32, C.c_int32,
                               C.c_double, C.c_int32, C.c_int32, _dp]
        L.ctbo_reactor_rhs.argtypes = [C.c_void_p, C.c_double, _dp, _dp]
        L.ctbo_reactor_initial_state.argtypes = [C.c_void_p, C.c_double, C.c_double, _dp, _dp]
        L.ctbo_reactor_contents.argtypes = [C.c_void_p, _dp, _dp, _dp]
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(_dp) if a is not None else None


class OracleMech:
    """One mechanism loaded into the CPU oracle."""

    def __init__(self, ctbm_path):
        self._h = lib().ctbo_load(str(ctbm_path).encode())
        if not self._h:
            raise RuntimeError(f"oracle: cannot load {ctbm_path}")
        self.K = lib().ctbo_nSpecies(self._h)
        self.R = lib().ctbo_nReactions(self._h)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ctbo_free(self._h)
            self._h = None

    def set_multiplier(self, i, f):
        lib().ctbo_setMultiplier(self._h, i, f)

    def rates(self, T, P, Y, want=("kf", "kc", "ropf", "ropr", "ropnet", "wdot", "cdot", "ddot")):
        T = np.ascontiguousarray(T, dtype=np.float64)
        P = np.ascontiguousarray(P, dtype=np.float64)
        Y = np.ascontiguousarray(Y, dtype=np.float64).reshape(len(T), self.K)
        n = len(T)
        names = ("kf", "kc", "ropf", "ropr", "ropnet", "wdot", "cdot", "ddot")
        out = {}
        for nm in names:
            if nm in want:
                out[nm] = np.empty((n, self.K if nm in ("wdot", "cdot", "ddot") else self.R))
        rc = lib().ctbo_rates_batch(self._h, n, _p(T), _p(P), _p(Y), self.K,
                                    *[_p(out.get(nm)) for nm in names])
        out["rc"] = rc
        return out

    def ddX(self, T, P, Y, skip_third_bodies=False, skip_falloff=False, rtol_delta=1e-8,
            analytic_falloff=False, ddCi=False):
        T = np.atleast_1d(np.asarray(T, dtype=np.float64))
        P = np.atleast_1d(np.asarray(P, dtype=np.float64))
        Y = np.ascontiguousarray(Y, dtype=np.float64).reshape(len(T), self.K)
        jac = np.empty((len(T), self.K * self.K))
        for n in range(len(T)):
            lib().ctbo_ddX(self._h, T[n], P[n], _p(Y[n]), int(skip_third_bodies), int(skip_falloff),
                           rtol_delta, int(analytic_falloff), int(ddCi), _p(jac[n]))
        return jac

    def reactor_initial_state(self, T, P, Y):
        y = np.empty(self.K + 1)
        Y = np.ascontiguousarray(Y, dtype=np.float64)
        lib().ctbo_reactor_initial_state(self._h, float(T), float(P), _p(Y), _p(y))
        return y

    def reactor_rhs(self, P, y):
        y = np.ascontiguousarray(y, dtype=np.float64)
        ydot = np.empty_like(y)
        lib().ctbo_reactor_rhs(self._h, float(P), _p(y), _p(ydot))
        return ydot

    def reactor_contents(self, y):
        y = np.ascontiguousarray(y, dtype=np.float64)
        Y = np.empty(self.K)
        T = C.c_double()
        lib().ctbo_reactor_contents(self._h, _p(y), C.byref(T), _p(Y))
        return T.value, Y
