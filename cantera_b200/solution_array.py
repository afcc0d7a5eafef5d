"""SolutionArray-shaped front end and the one-box multi-GPU split.

The reference's "many states" entry point is ``SolutionArray`` (interfaces/cython/cantera/
composite.py): ``states.TPY = T, P, Y`` followed by ``states.net_production_rates`` runs a Python loop
over the states (composite.py:2164-2173), one ``setState`` + getter each. ``BatchedSolutionArray``
keeps that surface — same property names, same [shape..., K] / [shape..., R] result shapes — and
evaluates the whole array in one call of the batched engine.

``MultiGpuKinetics`` / ``MultiGpuReactorNet`` are the library-level split over the GPUs of one box:
states are independent, so every device holds the (<= 100 KB) mechanism and owns a slice of the batch
(contiguous for the rate getters; shares of a seeded permutation for reactors, whose cost varies ~10x
with the initial temperature), one host thread per device drives its own handle and stream, and there
is NO collective on the data path.
"""
import threading

import numpy as np

from . import clib
from .batched import BatchedGasKinetics, BatchReactorNet
from .sharding import permuted_indices, slice_bounds


def _run_threads(jobs):
    """One host thread per device (ctypes releases the GIL inside the C ABI call)."""
    errs = []

    def wrap(f):
        def g():
            try:
                f()
            except Exception as e:  # noqa: BLE001
                errs.append(e)
        return g
    ts = [threading.Thread(target=wrap(f)) for f in jobs]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    if errs:
        raise errs[0]


class MultiGpuKinetics:
    """BatchedGasKinetics on several devices: host arrays in, host arrays out, contiguous slices."""

    def __init__(self, mechanism, devices=None):
        n = clib.check(clib.lib().ctb_nDevices())
        self.devices = list(range(n)) if devices is None else list(devices)
        if not self.devices:
            raise clib.CtbError("no usable CUDA device; this engine has no CPU path")
        self.parts = [BatchedGasKinetics(mechanism, device=d) for d in self.devices]
        self.n_species = self.parts[0].n_species
        self.n_reactions = self.parts[0].n_reactions

    def close(self):
        for p in self.parts:
            p.close()

    def _split(self, name, T, P, Y, width):
        T = np.ascontiguousarray(np.atleast_1d(T), dtype=np.float64)
        P = np.ascontiguousarray(np.atleast_1d(P), dtype=np.float64)
        Y = np.ascontiguousarray(Y, dtype=np.float64).reshape(T.size, -1)
        out = np.empty((T.size,) + width)
        w = len(self.parts)

        def job(r):
            b, e = slice_bounds(T.size, r, w)
            if e > b:
                getattr(self.parts[r], name)(T[b:e], P[b:e], Y[b:e], out=out[b:e])
        _run_threads([lambda r=r: job(r) for r in range(w)])
        return out

    def set_multiplier(self, i, f):
        for p in self.parts:
            p.set_multiplier(i, f)

    def set_derivative_settings(self, **kw):
        for p in self.parts:
            p.set_derivative_settings(**kw)


for _name, _kind in (("net_production_rates", "K"), ("creation_rates", "K"), ("destruction_rates", "K"),
                     ("forward_rates_of_progress", "R"), ("reverse_rates_of_progress", "R"),
                     ("net_rates_of_progress", "R"), ("forward_rate_constants", "R"),
                     ("equilibrium_constants", "R"), ("net_production_rates_ddX", "KK"),
                     ("net_production_rates_ddCi", "KK")):
    def _make(name, kind):
        def f(self, T, P, Y):
            K, R = self.n_species, self.n_reactions
            return self._split(name, T, P, Y, {"K": (K,), "R": (R,), "KK": (K, K)}[kind])
        f.__name__ = name
        f.__doc__ = "BatchedGasKinetics.%s over all devices (contiguous slices, one thread per device)." % name
        return f
    setattr(MultiGpuKinetics, _name, _make(_name, _kind))


class MultiGpuReactorNet:
    """N x (IdealGasConstPressureMoleReactor + ReactorNet) split over several devices: every device
    integrates its share of a seeded permutation of the reactors (static balance; inside a device the
    integrator's slot queue balances dynamically)."""

    def __init__(self, kinetics, n_reactors, seed=20261104):
        self.kinetics = kinetics
        self.n_reactors = int(n_reactors)
        w = len(kinetics.parts)
        self.shares = [permuted_indices(self.n_reactors, r, w, seed) for r in range(w)]
        self.nets = [BatchReactorNet(k, len(s)) if len(s) else None for k, s in zip(kinetics.parts, self.shares)]

    def close(self):
        for n in self.nets:
            if n is not None:
                n.close()

    def _each(self, f):
        _run_threads([lambda n=n, s=s: f(n, s) for n, s in zip(self.nets, self.shares) if n is not None])

    def set_state(self, T, P, Y):
        T, P, Y = np.asarray(T, dtype=np.float64), np.asarray(P, dtype=np.float64), np.asarray(Y, dtype=np.float64)
        self._each(lambda n, s: n.set_state(T[s], P[s], Y[s]))

    def set_tolerances(self, rtol=None, atol=None):
        for n in self.nets:
            if n is not None:
                if rtol is not None:
                    n.rtol = rtol
                if atol is not None:
                    n.atol = atol

    def set_ignition_monitor(self, dT):
        for n in self.nets:
            if n is not None:
                n.set_ignition_monitor(dT)

    def advance(self, t):
        self._each(lambda n, s: n.advance(t))

    def get_state(self):
        K = self.kinetics.n_species
        T, Y = np.empty(self.n_reactors), np.empty((self.n_reactors, K))

        def job(n, s):
            Ts, Ys = n.get_state()
            T[s], Y[s] = Ts, Ys
        self._each(job)
        return T, Y

    def _gather(self, name, dtype, width=()):
        out = np.empty((self.n_reactors,) + width, dtype=dtype)

        def job(n, s):
            out[s] = getattr(n, name)()
        self._each(job)
        return out

    def status(self):
        return self._gather("status", np.int32)

    def ignition_delay(self):
        return self._gather("ignition_delay", np.float64)

    def solver_stats(self):
        return self._gather("solver_stats", np.int64, (clib.NSTATS,))


class BatchedSolutionArray:
    """The SolutionArray surface (composite.py) for the quantities on the batched path.

        states = BatchedSolutionArray("gri30", shape=(nx, ny))
        states.TPY = T, P, Y                 # broadcast like SolutionArray's setters
        wdot = states.net_production_rates   # [nx, ny, K], one batched call
    """

    def __init__(self, mechanism, shape=(0,), devices=None):
        self.kinetics = MultiGpuKinetics(mechanism, devices) if (devices is None or len(devices) != 1) \
            else _Single(mechanism, devices[0])
        self.shape = (shape,) if np.isscalar(shape) else tuple(shape)
        self.size = int(np.prod(self.shape)) if self.shape else 1
        K = self.kinetics.n_species
        self._T = np.full(self.size, 300.0)
        self._P = np.full(self.size, 101325.0)
        self._Y = np.zeros((self.size, K))
        self._Y[:, 0] = 1.0

    n_species = property(lambda self: self.kinetics.n_species)
    n_reactions = property(lambda self: self.kinetics.n_reactions)

    def close(self):
        self.kinetics.close()

    # ---- state (SolutionArray.TPY etc.: values broadcast against the array shape)
    def _bc(self, v, trailing=()):
        return np.ascontiguousarray(np.broadcast_to(np.asarray(v, dtype=np.float64), self.shape + trailing)).reshape((self.size,) + trailing)

    @property
    def TPY(self):
        return self.T, self.P, self.Y

    @TPY.setter
    def TPY(self, tpy):
        T, P, Y = tpy
        if T is not None:
            self._T = self._bc(T)
        if P is not None:
            self._P = self._bc(P)
        if Y is not None:
            self._Y = self._bc(Y, (self.n_species,))

    T = property(lambda self: self._T.reshape(self.shape))
    P = property(lambda self: self._P.reshape(self.shape))
    Y = property(lambda self: self._Y.reshape(self.shape + (self.n_species,)))

    def _get(self, name, trailing):
        return getattr(self.kinetics, name)(self._T, self._P, self._Y).reshape(self.shape + trailing)


class _Single(BatchedGasKinetics):
    """One device behind the same calls as MultiGpuKinetics (no threads)."""

    def __init__(self, mechanism, device):
        super().__init__(mechanism, device=device)
        self.parts = [self]


for _name, _kind in (("net_production_rates", "K"), ("creation_rates", "K"), ("destruction_rates", "K"),
                     ("forward_rates_of_progress", "R"), ("reverse_rates_of_progress", "R"),
                     ("net_rates_of_progress", "R"), ("forward_rate_constants", "R"),
                     ("equilibrium_constants", "R"), ("net_production_rates_ddX", "KK"),
                     ("net_production_rates_ddCi", "KK")):
    def _prop(name, kind):
        def f(self):
            K, R = self.n_species, self.n_reactions
            return self._get(name, {"K": (K,), "R": (R,), "KK": (K, K)}[kind])
        f.__doc__ = "SolutionArray.%s for every state of the array." % name
        return property(f)
    setattr(BatchedSolutionArray, _name, _prop(_name, _kind))
