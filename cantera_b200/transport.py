"""Python mirror of the reference's mixture-averaged transport model for batches of states
(``Transport.viscosity``, ``thermal_conductivity``, ``mix_diff_coeffs`` of interfaces/cython/cantera/transport.py
after ``gas.TPY = ...``), through the C ABI (include/ctb_clib.h, ctb_trans_*).

    tr = BatchedMixTransport("gri30")            # cantera_b200/data/gri30_transport.ctbm
    mu, lam, D = tr.mixture_transport(T, P, Y)   # [N], [N], [N, K]

The temperature fits of the species properties are tables written by the reference-side tool
(oracle/ref_build/gen_golden.cpp ``trans``: GasTransport's public polynomial getters), like the mechanism
tables; numpy arrays = host memory, torch CUDA tensors = device memory (queued on torch's current stream)."""
import os

import numpy as np

from . import clib
from .batched import _Batch, _is_torch, _ptr
from .ctbm import read_arrays

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


def transport_path(name):
    if os.path.exists(name) and name.endswith((".ctbm", ".ctbm.gz")):
        return name
    base = os.path.basename(name)
    for suffix in (".yaml", ".yml", ".ctbm"):
        if base.endswith(suffix):
            base = base[: -len(suffix)]
    p = os.path.join(_DATA, base + "_transport.ctbm")
    if not os.path.exists(p):
        raise clib.CtbError("no transport table for '%s' (looked for %s); write one with oracle/_ref/gen_golden trans" % (name, p))
    return p


class BatchedMixTransport:
    def __init__(self, tables, device=0):
        """tables: a mechanism name / table file, or a dict with mol_weights, visc_poly, cond_poly, diff_poly."""
        self._lib = clib.lib()
        t = tables if isinstance(tables, dict) else read_arrays(transport_path(tables))
        mw, vc, cc, dc = (np.ascontiguousarray(t[k], dtype=np.float64) for k in ("mol_weights", "visc_poly", "cond_poly", "diff_poly"))
        K = mw.size
        if vc.size != 5 * K or cc.size != 5 * K or dc.size != 5 * K * (K + 1) // 2:
            raise clib.CtbError("transport tables of inconsistent size")
        self._h = clib.check(self._lib.ctb_trans_newFromArrays(K, mw.ctypes.data, vc.ctypes.data, cc.ctypes.data, dc.ctypes.data, device))
        self.n_species = K
        self.device = device

    @classmethod
    def from_yaml(cls, yaml, phase="", device=0, data_dirs=()):
        """newSolution(yaml, phase, "mixture-averaged") by the reference itself, in this process, through the optional
        bridge libctb_cantera.so (include/ctb_cantera.h: ctb_trans_new)."""
        b = clib.bridge()
        if b is None:
            raise clib.CtbError("libctb_cantera.so (the reference bridge) is not built; use a transport table file")
        for d in data_dirs:
            clib.check(b.ctb_addDataDirectory(str(d).encode()))
        self = cls.__new__(cls)
        self._lib = clib.lib()
        self._h = clib.check(b.ctb_trans_new(str(yaml).encode(), phase.encode(), device))
        self.n_species = clib.check(self._lib.ctb_trans_nSpecies(self._h))
        self.device = device
        return self

    def close(self):
        if getattr(self, "_h", None) is not None:
            self._lib.ctb_trans_del(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def mixture_transport(self, T, P, Y, want=("viscosity", "thermal_conductivity", "mix_diff_coeffs")):
        """-> tuple of the wanted quantities, in the order of `want`."""
        b = _Batch(T, P, Y, self.n_species)
        if b.space == clib.DEVICE:
            clib.check(self._lib.ctb_trans_setStream(self._h, b.torch.cuda.current_stream().cuda_stream))
        K = self.n_species
        out = {"viscosity": b.empty(b.N) if "viscosity" in want else None,
               "thermal_conductivity": b.empty(b.N) if "thermal_conductivity" in want else None,
               "mix_diff_coeffs": b.empty(b.N, K) if "mix_diff_coeffs" in want else None}
        if b.N:
            p = [(_ptr(v) if v is not None else None) for v in out.values()]
            clib.check(self._lib.ctb_trans_getMixtureTransport(self._h, b.N, _ptr(b.T), _ptr(b.P), _ptr(b.Y), b.ldY,
                                                               p[0], p[1], p[2], K, b.space))
        return tuple(out[w] for w in want)

    def viscosity(self, T, P, Y):
        return self.mixture_transport(T, P, Y, ("viscosity",))[0]

    def thermal_conductivity(self, T, P, Y):
        return self.mixture_transport(T, P, Y, ("thermal_conductivity",))[0]

    def mix_diff_coeffs(self, T, P, Y):
        return self.mixture_transport(T, P, Y, ("mix_diff_coeffs",))[0]
