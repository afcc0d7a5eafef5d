"""Host-side mirror of the reference's Kinetics / ReactorNet interface for a BATCH of states.

``BatchedGasKinetics`` follows ``Kinetics`` (include/cantera/kinetics/Kinetics.h:348,572,1117;
Python names as in interfaces/cython/cantera/kinetics.pyx: ``net_production_rates``,
``forward_rates_of_progress``, ``net_production_rates_ddX`` ...), ``BatchReactorNet`` follows
``ReactorNet`` + ``IdealGasConstPressureMoleReactor`` (include/cantera/zeroD/ReactorNet.h:57,159).
Every call goes through the C ABI (include/ctb_clib.h). Arrays are either numpy (host memory,
staged by the library) or torch CUDA tensors (device memory, no copies; every such call first points
the handle at torch's current stream, so it is ordered against the producers and consumers of the
tensors).
"""
import os

import numpy as np

from . import clib

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")


def mechanism_path(name):
    """Resolve 'gri30', 'gri30.yaml' or a path to the flat-table file (.ctbm) of a mechanism."""
    if os.path.exists(name) and name.endswith(".ctbm"):
        return name
    base = os.path.basename(name)
    for suffix in (".yaml", ".yml", ".ctbm"):
        if base.endswith(suffix):
            base = base[: -len(suffix)]
    p = os.path.join(_DATA, base + ".ctbm")
    if not os.path.exists(p):
        raise clib.CtbError(
            f"no flat mechanism table for '{name}' (looked for {p}); produce one from the YAML file with the "
            "reference-side extractor (include/cantera_b200/MechanismExtract.h, oracle/ref_build/ctb_extract)")
    return p


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _ptr(x):
    return x.data_ptr() if _is_torch(x) else x.ctypes.data


class _Batch:
    """Validated (T, P, Y) in one memory space."""

    def __init__(self, T, P, Y, K):
        if _is_torch(T):
            import torch
            if not (T.is_cuda and P.is_cuda and Y.is_cuda):
                raise clib.CtbError("torch inputs must be CUDA tensors (use numpy arrays for host memory)")
            self.space = clib.DEVICE
            self.T = T.contiguous().to(torch.float64)
            self.P = P.contiguous().to(torch.float64)
            self.Y = Y.to(torch.float64)
            if self.Y.dim() != 2 or self.Y.stride(1) != 1:
                self.Y = self.Y.reshape(self.T.numel(), -1).contiguous()
            self.ldY = self.Y.stride(0) if self.Y.shape[0] > 1 else self.Y.shape[1]
            self.N = self.T.numel()
            self.torch = torch
        else:
            self.space = clib.HOST
            self.T = np.ascontiguousarray(np.atleast_1d(T), dtype=np.float64)
            self.P = np.ascontiguousarray(np.atleast_1d(P), dtype=np.float64)
            self.N = self.T.size
            Y = np.ascontiguousarray(Y, dtype=np.float64)
            self.Y = Y.reshape(self.N, -1) if self.N else Y.reshape(0, K)
            self.ldY = self.Y.shape[1]
            self.torch = None
        if self.P.shape[0] != self.N or self.Y.shape[0] != self.N:
            raise clib.CtbError("T, P and Y must describe the same number of states")
        if self.Y.shape[1] < K:
            raise clib.CtbError(f"Y has {self.Y.shape[1]} columns; the mechanism has {K} species")

    def empty(self, *shape):
        if self.torch is not None:
            return self.torch.empty(shape, dtype=self.torch.float64, device=self.T.device)
        return np.empty(shape, dtype=np.float64)


#: python name -> (CTB_Q_* id, "R" | "K"): the reference's Python property names (kinetics.pyx)
QUANTITIES = {}
for _i, _fam in enumerate(("forward_rate_constants", "forward_rates_of_progress", "reverse_rates_of_progress",
                           "net_rates_of_progress")):
    for _j, _d in enumerate(("ddT", "ddP", "ddC")):
        QUANTITIES["%s_%s" % (_fam, _d)] = (3 * _i + _j, "R")
for _i, _fam in enumerate(("creation_rates", "destruction_rates", "net_production_rates")):
    for _j, _d in enumerate(("ddT", "ddP", "ddC")):
        QUANTITIES["%s_%s" % (_fam, _d)] = (12 + 3 * _i + _j, "K")
for _i, _nm in enumerate(("delta_gibbs", "delta_enthalpy", "delta_entropy", "delta_standard_gibbs",
                          "delta_standard_enthalpy", "delta_standard_entropy")):
    QUANTITIES[_nm] = (21 + _i, "R")


class BatchedGasKinetics:
    """Kinetics/BulkKinetics of one ideal-gas mechanism evaluated for N states per call."""

    def __init__(self, mechanism, device=0, _handle=None):
        self._lib = clib.lib()
        self._h = _handle if _handle is not None else \
            clib.check(self._lib.ctb_kin_newFromFile(mechanism_path(mechanism).encode(), device))
        self.device = device
        self.n_species = clib.check(self._lib.ctb_kin_nSpecies(self._h))
        self.n_reactions = clib.check(self._lib.ctb_kin_nReactions(self._h))

    @classmethod
    def from_yaml(cls, yaml, phase="", device=0, data_dirs=()):
        """newSolution(yaml, phase) by the reference itself, in this process, through the optional bridge
        libctb_cantera.so (include/ctb_cantera.h: ctb_kin_new)."""
        b = clib.bridge()
        if b is None:
            raise clib.CtbError("libctb_cantera.so (the reference bridge) is not built; use a .ctbm table file")
        for d in data_dirs:
            clib.check(b.ctb_addDataDirectory(str(d).encode()))
        return cls(yaml, device, _handle=clib.check(b.ctb_kin_new(-1, str(yaml).encode(), phase.encode(), device)))

    # ---- lifetime
    @classmethod
    def from_yaml_standalone(cls, yaml_path, phase="", device=0, data_dirs=()):
        """The mechanism read from the reference's YAML format WITHOUT the reference (cantera_b200/yaml_mech.py:
        bit-identical tables for the config mechanisms and the reference's kinetics test inputs)."""
        import os
        import tempfile
        from .ctbm import write_arrays
        from .yaml_mech import tables_from_yaml
        tables = tables_from_yaml(yaml_path, phase, data_dirs)
        fd, path = tempfile.mkstemp(suffix=".ctbm")
        os.close(fd)
        try:
            write_arrays(path, tables)
            return cls(path, device=device)
        finally:
            os.unlink(path)

    def close(self):
        if getattr(self, "_h", None) is not None:
            self._lib.ctb_kin_del(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    def use_torch_stream(self):
        """Queue subsequent device-memory calls on torch's current CUDA stream. (Every call that is
        handed torch tensors does this itself, see _on_torch_stream; kept for explicit use.)"""
        import torch
        clib.check(self._lib.ctb_kin_setStream(self._h, torch.cuda.current_stream().cuda_stream))

    def _on_torch_stream(self, batch):
        """Device-memory (torch) arguments: the kernels are queued on torch's CURRENT stream, so they are
        ordered after the kernels that produced the inputs and before whatever consumes the outputs,
        and the caching allocator's stream bookkeeping stays valid."""
        if batch.space == clib.DEVICE:
            clib.check(self._lib.ctb_kin_setStream(self._h, batch.torch.cuda.current_stream().cuda_stream))

    def synchronize(self):
        clib.check(self._lib.ctb_kin_synchronize(self._h))

    # ---- Kinetics settings
    def set_multiplier(self, i, f):
        clib.check(self._lib.ctb_kin_setMultiplier(self._h, i, float(f)))

    def multiplier(self, i):
        return clib.check(self._lib.ctb_kin_multiplier(self._h, i))

    def set_derivative_settings(self, skip_third_bodies=False, skip_falloff=False, rtol_delta=1e-8):
        clib.check(self._lib.ctb_kin_setDerivativeSettings(self._h, int(skip_third_bodies), int(skip_falloff),
                                                           float(rtol_delta)))

    # ---- getters
    def _species(self, fn, T, P, Y, out=None):
        b = _Batch(T, P, Y, self.n_species)
        self._on_torch_stream(b)
        if out is None:
            out = b.empty(b.N, self.n_species)
        if b.N:
            clib.check(fn(self._h, b.N, _ptr(b.T), _ptr(b.P), _ptr(b.Y), b.ldY, _ptr(out), out.shape[1], b.space))
        return out

    def _reaction(self, fn, T, P, Y, out=None):
        b = _Batch(T, P, Y, self.n_species)
        self._on_torch_stream(b)
        if out is None:
            out = b.empty(b.N, self.n_reactions)
        if b.N:
            clib.check(fn(self._h, b.N, _ptr(b.T), _ptr(b.P), _ptr(b.Y), b.ldY, _ptr(out), out.shape[1], b.space))
        return out

    def net_production_rates(self, T, P, Y, out=None):
        """Kinetics::getNetProductionRates per state, [N, K] kmol/m^3/s."""
        return self._species(self._lib.ctb_kin_getNetProductionRates, T, P, Y, out)

    def creation_rates(self, T, P, Y, out=None):
        return self._species(self._lib.ctb_kin_getCreationRates, T, P, Y, out)

    def destruction_rates(self, T, P, Y, out=None):
        return self._species(self._lib.ctb_kin_getDestructionRates, T, P, Y, out)

    def forward_rates_of_progress(self, T, P, Y, out=None):
        return self._reaction(self._lib.ctb_kin_getFwdRatesOfProgress, T, P, Y, out)

    def reverse_rates_of_progress(self, T, P, Y, out=None):
        return self._reaction(self._lib.ctb_kin_getRevRatesOfProgress, T, P, Y, out)

    def net_rates_of_progress(self, T, P, Y, out=None):
        return self._reaction(self._lib.ctb_kin_getNetRatesOfProgress, T, P, Y, out)

    def forward_rate_constants(self, T, P, Y, out=None):
        return self._reaction(self._lib.ctb_kin_getFwdRateConstants, T, P, Y, out)

    def equilibrium_constants(self, T, P, Y, out=None):
        return self._reaction(self._lib.ctb_kin_getEquilibriumConstants, T, P, Y, out)

    def quantity(self, name, T, P, Y, out=None):
        """Any of QUANTITIES (forward_rates_of_progress_ddT, net_production_rates_ddC, delta_gibbs, ...):
        [N, R] or [N, K] per the reference's getter of that name."""
        qid, kind = QUANTITIES[name]
        b = _Batch(T, P, Y, self.n_species)
        self._on_torch_stream(b)
        if out is None:
            out = b.empty(b.N, self.n_reactions if kind == "R" else self.n_species)
        if b.N:
            clib.check(self._lib.ctb_kin_getQuantity(self._h, qid, b.N, _ptr(b.T), _ptr(b.P), _ptr(b.Y), b.ldY, _ptr(out),
                                                     out.shape[1], b.space))
        return out

    def __getattr__(self, name):
        if name in QUANTITIES:
            return lambda T, P, Y, out=None: self.quantity(name, T, P, Y, out)
        raise AttributeError(name)

    def _jac(self, fn, T, P, Y, out):
        b = _Batch(T, P, Y, self.n_species)
        self._on_torch_stream(b)
        K = self.n_species
        if out is None:
            out = b.empty(b.N, K, K)
        if b.N:
            buf_len = out.numel() if _is_torch(out) else out.size
            clib.check(fn(self._h, b.N, _ptr(b.T), _ptr(b.P), _ptr(b.Y), b.ldY, _ptr(out), buf_len, b.space))
        return out

    def net_production_rates_ddX(self, T, P, Y, out=None):
        """Kinetics::netProductionRates_ddX per state as [N, K, K] with out[n, j, k] = d wdot_k / d X_j
        (the column-major K x K buffer of kin_getNetProductionRates_ddX)."""
        return self._jac(self._lib.ctb_kin_getNetProductionRates_ddX, T, P, Y, out)

    def net_production_rates_ddCi(self, T, P, Y, out=None):
        return self._jac(self._lib.ctb_kin_getNetProductionRates_ddCi, T, P, Y, out)

    def net_production_rates_ddX_no_alloc(self, T, P, Y, out):
        """net_production_rates_ddX for callers that may not allocate device memory
        (ctb_kin_getNetProductionRates_ddX_noAlloc: static scratch array, at most clib.NOALLOC_MAX_STATES
        states per call, calls must not overlap on the GPU). CUDA float64 tensors only, contiguous, and
        ``out`` [N, K, K] is the caller's: nothing is allocated here either."""
        if not (_is_torch(T) and _is_torch(P) and _is_torch(Y) and _is_torch(out)):
            raise clib.CtbError("net_production_rates_ddX_no_alloc takes CUDA tensors (device memory only)")
        import torch
        K = self.n_species
        for x in (T, P, Y, out):
            if not (x.is_cuda and x.dtype == torch.float64 and x.is_contiguous()):
                raise clib.CtbError("net_production_rates_ddX_no_alloc: contiguous CUDA float64 tensors only")
        clib.check(self._lib.ctb_kin_setStream(self._h, torch.cuda.current_stream().cuda_stream))
        clib.check(self._lib.ctb_kin_getNetProductionRates_ddX_noAlloc(
            self._h, T.numel(), T.data_ptr(), P.data_ptr(), Y.data_ptr(), Y.shape[-1], out.data_ptr(), out.numel()))
        return out


REACTOR_TYPES = ("IdealGasConstPressureMoleReactor", "IdealGasMoleReactor", "IdealGasReactor",
                 "IdealGasConstPressureReactor")


class BatchReactorNet:
    """N x (reactor + ReactorNet) advanced together; ``reactor_type`` is one of REACTOR_TYPES (the reference's
    class names), default IdealGasConstPressureMoleReactor."""

    def __init__(self, kinetics, n_reactors, reactor_type="IdealGasConstPressureMoleReactor"):
        self._lib = clib.lib()
        self.kinetics = kinetics
        self.n_reactors = int(n_reactors)
        if reactor_type not in REACTOR_TYPES:
            raise clib.CtbError("unknown reactor type '%s'; one of %s" % (reactor_type, ", ".join(REACTOR_TYPES)))
        self.reactor_type = reactor_type
        self._h = clib.check(self._lib.ctb_reactornet_newTyped(kinetics.handle, self.n_reactors, REACTOR_TYPES.index(reactor_type)))
        self.n_equations = clib.check(self._lib.ctb_reactornet_nEquations(self._h))
        self._rtol, self._atol = 1e-9, 1e-15  # ReactorNet.h:483-485

    def close(self):
        if getattr(self, "_h", None) is not None:
            self._lib.ctb_reactornet_del(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_state(self, T, P, Y):
        b = _Batch(T, P, Y, self.kinetics.n_species)
        if b.N != self.n_reactors:
            raise clib.CtbError(f"expected {self.n_reactors} states, got {b.N}")
        self.kinetics._on_torch_stream(b)
        clib.check(self._lib.ctb_reactornet_setState(self._h, _ptr(b.T), _ptr(b.P), _ptr(b.Y), b.ldY, b.space))

    @property
    def rtol(self):
        return self._rtol

    @rtol.setter
    def rtol(self, v):
        clib.check(self._lib.ctb_reactornet_setRelativeTolerance(self._h, float(v)))
        self._rtol = float(v)

    @property
    def atol(self):
        return self._atol

    @atol.setter
    def atol(self, v):
        clib.check(self._lib.ctb_reactornet_setAbsoluteTolerance(self._h, float(v)))
        self._atol = float(v)

    def effective_tolerances(self):
        """(rtol, atol of the T row, atol of the species rows, user_set) as the integrator applies them
        (ctb_reactornet_getTolerances); without a user-set atol the species value is per kmol of contents."""
        out = np.empty(4)
        clib.check(self._lib.ctb_reactornet_getTolerances(self._h, 4, out.ctypes.data))
        return float(out[0]), float(out[1]), float(out[2]), bool(out[3])

    def set_max_steps(self, n):
        clib.check(self._lib.ctb_reactornet_setMaxSteps(self._h, int(n)))

    def set_initial_time(self, t0):
        clib.check(self._lib.ctb_reactornet_setInitialTime(self._h, float(t0)))

    def set_ignition_monitor(self, dT):
        clib.check(self._lib.ctb_reactornet_setIgnitionMonitor(self._h, float(dT)))

    def advance(self, t):
        clib.check(self._lib.ctb_reactornet_advance(self._h, float(t)))

    @property
    def time(self):
        return clib.check(self._lib.ctb_reactornet_time(self._h))

    def get_state(self, device_like=None):
        """(T[N], Y[N, K]) of the reactor contents; torch CUDA tensors if ``device_like`` is one."""
        K = self.kinetics.n_species
        if device_like is not None and _is_torch(device_like):
            import torch
            T = torch.empty(self.n_reactors, dtype=torch.float64, device=device_like.device)
            Y = torch.empty(self.n_reactors, K, dtype=torch.float64, device=device_like.device)
            space = clib.DEVICE
            clib.check(self._lib.ctb_kin_setStream(self.kinetics.handle, torch.cuda.current_stream().cuda_stream))
        else:
            T = np.empty(self.n_reactors)
            Y = np.empty((self.n_reactors, K))
            space = clib.HOST
        clib.check(self._lib.ctb_reactornet_getState(self._h, _ptr(T), _ptr(Y), K, space))
        return T, Y

    def get_state_vector(self):
        """ReactorNet::getState: [N, n_equations] rows, layout by reactor model (host)."""
        y = np.empty((self.n_reactors, self.n_equations))
        clib.check(self._lib.ctb_reactornet_getStateVector(self._h, y.ctypes.data, y.size, clib.HOST))
        return y

    def eval(self):
        """ReactorNet::eval at the current state: ydot [N, n_equations] (host)."""
        yd = np.empty((self.n_reactors, self.n_equations))
        clib.check(self._lib.ctb_reactornet_eval(self._h, yd.ctypes.data, yd.size, clib.HOST))
        return yd

    def jacobian(self):
        """d(ydot)/dy at the current state: [N, K+1, K+1] with out[n, j, i] = d ydot_i / d y_j (host): the
        reference's ``Reactor.jacobian`` layout transposed per reactor (column-major buffer of
        ctb_reactornet_getJacobian); includes the exact volume-dilution terms, see include/ctb_clib.h."""
        n = self.kinetics.n_species + 1
        J = np.empty((self.n_reactors, n, n))
        clib.check(self._lib.ctb_reactornet_getJacobian(self._h, J.ctypes.data, J.size, clib.HOST))
        return J

    def last_rounds(self):
        """Kernel rounds of the batch-synchronous integrator in the last advance."""
        return int(clib.check(self._lib.ctb_reactornet_lastRounds(self._h)))

    def status(self):
        s = np.empty(self.n_reactors, dtype=np.int32)
        clib.check(self._lib.ctb_reactornet_status(self._h, s.size, s.ctypes.data))
        return s

    def solver_stats(self):
        """[N, 8] counters: steps, rhs_evals, jac_evals, lin_solve_setups, nonlinear_iters,
        nonlinear_conv_fails, err_test_fails, last_order (ReactorNet::solverStats)."""
        s = np.empty((self.n_reactors, clib.NSTATS), dtype=np.int64)
        clib.check(self._lib.ctb_reactornet_solverStats(self._h, s.size, s.ctypes.data))
        return s

    def ignition_delay(self):
        t = np.empty(self.n_reactors)
        clib.check(self._lib.ctb_reactornet_ignitionDelay(self._h, t.size, t.ctypes.data))
        return t
