"""Parity of the CUDA kinetics path (through the C ABI) against the reference's golden vectors and
the CPU oracle. Tolerance (BASELINE.json north_star): 1e-12 relative, FP64, summation-order
differences only. For quantities formed by cancellation (ropnet, wdot) the error is measured
against the magnitude of the terms that were subtracted (SURVEY.md §7 hard part 1)."""
import numpy as np
import pytest

from conftest import MECHS, golden

pytestmark = pytest.mark.gpu
TOL = 1e-12
CASES = sorted(MECHS)


def _rel(a, b, scale=None):
    s = np.abs(b) if scale is None else scale
    return float(np.max(np.abs(a - b) / np.maximum(s, 1e-300)))


@pytest.fixture(scope="module")
def kin_cache():
    from cantera_b200 import BatchedGasKinetics
    cache = {}

    def get(case):
        if case not in cache:
            cache[case] = BatchedGasKinetics(MECHS[case])
        return cache[case]
    yield get
    for k in cache.values():
        k.close()


@pytest.mark.parametrize("case", CASES)
def test_rates_vs_reference_golden(case, kin_cache):
    kin = kin_cache(case)
    g = golden("rates", case)
    N, K, R = g["dims"]
    T, P, Y = g["T"], g["P"], g["Y"].reshape(N, K)
    ref = {k: g[k].reshape(N, -1) for k in ("kf", "kc", "ropf", "ropr", "ropnet", "wdot", "cdot", "ddot")}
    assert _rel(kin.forward_rate_constants(T, P, Y), ref["kf"]) <= TOL
    assert _rel(kin.equilibrium_constants(T, P, Y), ref["kc"]) <= TOL
    assert _rel(kin.forward_rates_of_progress(T, P, Y), ref["ropf"]) <= TOL
    assert _rel(kin.reverse_rates_of_progress(T, P, Y), ref["ropr"]) <= TOL
    assert _rel(kin.net_rates_of_progress(T, P, Y), ref["ropnet"], np.maximum(ref["ropf"], ref["ropr"])) <= TOL
    assert _rel(kin.creation_rates(T, P, Y), ref["cdot"]) <= TOL
    assert _rel(kin.destruction_rates(T, P, Y), ref["ddot"]) <= TOL
    assert _rel(kin.net_production_rates(T, P, Y), ref["wdot"], ref["cdot"] + ref["ddot"]) <= TOL


@pytest.mark.parametrize("case", CASES)
def test_ddx_vs_reference_golden(case, kin_cache):
    kin = kin_cache(case)
    g = golden("ddx", case)
    N, K, R = g["dims"]
    T, P, Y = g["T"], g["P"], g["Y"].reshape(N, K)

    def scale(ref):
        return np.abs(ref).max(axis=(1, 2), keepdims=True) * np.ones_like(ref)

    # analytic terms: exact parity
    for key, kw in (("ddX_skip_falloff", dict(skip_falloff=True)),
                    ("ddX_skip_third_bodies", dict(skip_third_bodies=True))):
        kin.set_derivative_settings(**kw)
        ref = g[key].reshape(N, K, K)
        assert _rel(kin.net_production_rates_ddX(T, P, Y), ref, scale(ref)) <= TOL, key
    # default settings: the reference's falloff term is a finite difference (MultiRate.h:135-165)
    # whose own noise is ~1e-8; ours is the exact derivative
    kin.set_derivative_settings()
    for key, fn in (("ddX", kin.net_production_rates_ddX), ("ddCi", kin.net_production_rates_ddCi)):
        ref = g[key].reshape(N, K, K)
        assert _rel(fn(T, P, Y), ref, scale(ref)) <= 1e-6, key


@pytest.mark.parametrize("case", CASES)
def test_ddx_vs_oracle_analytic(case, kin_cache, oracle_lib):
    """Against the oracle with the exact falloff derivative the default-settings Jacobian is 1e-12."""
    kin = kin_cache(case)
    kin.set_derivative_settings()
    m = oracle_lib.OracleMech(MECHS[case])
    g = golden("ddx", case)
    N, K, R = g["dims"]
    T, P, Y = g["T"], g["P"], g["Y"].reshape(N, K)
    for ddCi, fn in ((False, kin.net_production_rates_ddX), (True, kin.net_production_rates_ddCi)):
        ref = m.ddX(T, P, Y, analytic_falloff=True, ddCi=ddCi).reshape(N, K, K)
        sc = np.abs(ref).max(axis=(1, 2), keepdims=True) * np.ones_like(ref)
        assert _rel(fn(T, P, Y), ref, sc) <= TOL


@pytest.mark.parametrize("case,n", [("h2o2", 20000), ("gri30", 6000), ("nDodecane", 3000)])
def test_rates_vs_oracle_random(case, n, kin_cache, oracle_lib):
    """Seeded random states (T 800-2500 K, P 1-50 atm, random Y) beyond the golden files, host and
    device memory paths, with a ragged leading dimension."""
    import torch
    kin = kin_cache(case)
    m = oracle_lib.OracleMech(MECHS[case])
    K = m.K
    rng = np.random.default_rng(20261104)
    T = rng.uniform(800, 2500, n)
    P = 101325.0 * rng.uniform(1, 50, n)
    Y = rng.uniform(0, 1, (n, K))
    Y /= Y.sum(axis=1, keepdims=True)
    o = m.rates(T, P, Y, want=("ropf", "wdot", "cdot", "ddot"))
    sc = o["cdot"] + o["ddot"]
    assert _rel(kin.net_production_rates(T, P, Y), o["wdot"], sc) <= TOL
    assert _rel(kin.forward_rates_of_progress(T, P, Y), o["ropf"]) <= TOL
    # device memory, ldY > K
    Yp = torch.zeros(n, K + 3, dtype=torch.float64, device="cuda")
    Yp[:, :K] = torch.from_numpy(Y).cuda()
    w = kin.net_production_rates(torch.from_numpy(T).cuda(), torch.from_numpy(P).cuda(), Yp[:, :K])
    kin.synchronize()
    assert _rel(w.cpu().numpy(), o["wdot"], sc) <= TOL


@pytest.mark.parametrize("case,Tlo,Thi", [("gri30", 120.0, 400.0), ("nDodecane", 150.0, 600.0), ("nDodecane", 250.0, 3500.0)])
def test_rates_extreme_temperatures(case, Tlo, Thi, kin_cache, oracle_lib):
    """Very low temperatures make |g'_k| = |mu0_k/RT - ln(P/RT)| exceed the range in which the tile
    kernel's per-species factors exp(+-g') can be multiplied safely: such tiles switch to the safe
    mode (rates_tile.cuh: one exp of the summed g' per reaction, the reference's own formula). The
    batch mixes tiles of both kinds; parity with the oracle as everywhere else."""
    kin = kin_cache(case)
    m = oracle_lib.OracleMech(MECHS[case])
    K = m.K
    n = 2000
    rng = np.random.default_rng(5)
    T = rng.uniform(Tlo, Thi, n)
    P = 101325.0 * rng.uniform(0.1, 50, n)
    Y = rng.uniform(0, 1, (n, K))
    Y /= Y.sum(axis=1, keepdims=True)
    o = m.rates(T, P, Y, want=("wdot", "cdot", "ddot"))
    sc = o["cdot"] + o["ddot"]
    w = kin.net_production_rates(T, P, Y)
    assert np.all(np.isfinite(w))
    live = sc > 1e-250  # below that the terms themselves are denormal
    assert float(np.max(np.abs(w - o["wdot"])[live] / sc[live])) <= TOL


@pytest.mark.parametrize("yaml,phase,case", [("data/gri30.yaml", "gri30", "gri30"), ("testdata/pdep-test.yaml", "gas", "pdep-test")])
def test_kin_new_in_process(yaml, phase, case, kin_cache):
    """ctb_kin_new (include/ctb_cantera.h): newSolution() by the reference in this process -> tables ->
    ctb_kin_newFromTables, against the handle loaded from the .ctbm file: bit-identical outputs."""
    import os
    from conftest import ROOT
    from cantera_b200 import BatchedGasKinetics, clib
    if clib.bridge() is None:
        pytest.skip("libctb_cantera.so not built")
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    kin2 = BatchedGasKinetics.from_yaml(os.path.join(ref_dir, yaml), phase, data_dirs=[os.path.join(ref_dir, "data")])
    kin = kin_cache(case)
    assert (kin2.n_species, kin2.n_reactions) == (kin.n_species, kin.n_reactions)
    rng = np.random.default_rng(11)
    n, K = 777, kin.n_species
    T = rng.uniform(800, 2500, n)
    P = 101325.0 * rng.uniform(1, 50, n)
    Y = rng.uniform(0, 1, (n, K))
    Y /= Y.sum(axis=1, keepdims=True)
    assert np.array_equal(kin2.net_production_rates(T, P, Y), kin.net_production_rates(T, P, Y))
    assert np.array_equal(kin2.forward_rates_of_progress(T, P, Y), kin.forward_rates_of_progress(T, P, Y))
    assert np.array_equal(kin2.net_production_rates_ddX(T[:50], P[:50], Y[:50]), kin.net_production_rates_ddX(T[:50], P[:50], Y[:50]))
    kin2.close()


def test_kin_from_standalone_yaml_loader(kin_cache):
    """f4: the mechanism read by cantera_b200/yaml_mech.py (no reference at run time) against the handle loaded
    from the reference-extracted .ctbm file: the tables are identical (tests/test_yaml_mech.py), so are the
    outputs. The toy input of tests/golden runs end to end as well."""
    import os
    from conftest import GOLDEN, ROOT
    from cantera_b200 import BatchedGasKinetics
    toy = BatchedGasKinetics.from_yaml_standalone(os.path.join(GOLDEN, "toy_mech.yaml"), "toy")
    assert (toy.n_species, toy.n_reactions) == (5, 6)
    w = toy.net_production_rates(np.array([1500.0]), np.array([2e5]), np.array([[0.2, 0.1, 0.3, 0.2, 0.2]]))
    assert np.all(np.isfinite(w)) and np.any(w != 0.0)
    toy.close()
    path = os.path.join(ROOT, "oracle", "_ref", "data", "gri30.yaml")
    if not os.path.exists(path):
        pytest.skip("the reference's YAML data files are not on this machine")
    kin2 = BatchedGasKinetics.from_yaml_standalone(path, "gri30")
    kin = kin_cache("gri30")
    rng = np.random.default_rng(12)
    n, K = 513, kin.n_species
    T = rng.uniform(800, 2500, n)
    P = 101325.0 * rng.uniform(1, 50, n)
    Y = rng.uniform(0, 1, (n, K))
    Y /= Y.sum(axis=1, keepdims=True)
    assert np.array_equal(kin2.net_production_rates(T, P, Y), kin.net_production_rates(T, P, Y))
    assert np.array_equal(kin2.net_production_rates_ddX(T[:40], P[:40], Y[:40]), kin.net_production_rates_ddX(T[:40], P[:40], Y[:40]))
    kin2.close()


DERIV_CASES = ["h2o2", "gri30", "nDodecane", "pdep-test", "sri-falloff", "tsang-falloff", "chemically-activated-reaction",
               "explicit-third-bodies"]


@pytest.mark.parametrize("case", DERIV_CASES)
def test_derivative_families_and_deltas(case, kin_cache):
    """f2: the _ddT / _ddP / _ddC families and the reaction deltas against the reference (golden derivs_*).
    The deltas and every Arrhenius-rate derivative are analytic on both sides: 1e-12 of the row's
    largest magnitude. Where the reference itself takes a forward difference with rtol-delta = 1e-8
    (falloff / Plog / Chebyshev rate constants, d(1/Kc)/dT) two implementations agree only to the noise
    of that difference, ~1e-16 / 1e-8 amplified by the cancellation in the quotient: 2e-5 of the row's
    largest magnitude (observed up to 6e-6), 2e-4 for the pure difference quotient kf_ddC (observed 5e-5)."""
    from cantera_b200.batched import QUANTITIES
    kin = kin_cache(case)
    kin.set_derivative_settings()
    g = golden("derivs", case)
    N, K, R = g["dims"]
    T, P, Y = g["T"], g["P"], g["Y"].reshape(N, K)
    names = {"forward_rate_constants": "kf", "forward_rates_of_progress": "ropf", "reverse_rates_of_progress": "ropr",
             "net_rates_of_progress": "ropnet", "creation_rates": "cdot", "destruction_rates": "ddot",
             "net_production_rates": "wdot"}
    gold = {"delta_gibbs": "deltaGibbs", "delta_enthalpy": "deltaEnthalpy", "delta_entropy": "deltaEntropy",
            "delta_standard_gibbs": "deltaSSGibbs", "delta_standard_enthalpy": "deltaSSEnthalpy",
            "delta_standard_entropy": "deltaSSEntropy"}
    for fam, short in names.items():
        for d in ("ddT", "ddP", "ddC"):
            gold["%s_%s" % (fam, d)] = "%s_%s" % (short, d)
    assert sorted(gold) == sorted(QUANTITIES)
    worst = {}
    for name, key in gold.items():
        ref = g[key].reshape(N, -1)
        out = kin.quantity(name, T, P, Y)
        assert out.shape == ref.shape, name
        scale = np.maximum(np.abs(ref).max(axis=1, keepdims=True), 1e-300)
        err = float(np.max(np.abs(out - ref) / scale))
        worst[name] = err
        # forward_rate_constants_ddC is nothing but the [M]-difference quotient of the falloff rates: near
        # the high-pressure limit the quotient itself is ~1e-3, so its noise is relatively larger
        tol = 1e-12 if name.startswith("delta") else 2e-4 if name == "forward_rate_constants_ddC" else 2e-5
        assert err <= tol, (name, err)
    # reactions with plain Arrhenius rates are analytic on both sides: exact parity of their rows
    from cantera_b200.ctbm import read_arrays
    arr = read_arrays(MECHS[case])["rxn_type"] == 0
    if arr.any():
        for name in ("forward_rate_constants_ddT", "forward_rates_of_progress_ddT", "forward_rates_of_progress_ddC"):
            ref = g[gold[name]].reshape(N, R)[:, arr]
            out = kin.quantity(name, T, P, Y)[:, arr]
            sc = np.maximum(np.abs(ref).max(axis=1, keepdims=True), 1e-300)
            assert float(np.max(np.abs(out - ref) / sc)) <= 1e-12, name
    # the getters are also attributes named like the reference's Python properties
    assert np.array_equal(kin.net_production_rates_ddT(T, P, Y), kin.quantity("net_production_rates_ddT", T, P, Y))


def test_multiplier_and_errors(kin_cache, oracle_lib):
    from cantera_b200 import CtbError
    kin = kin_cache("h2o2")
    m = oracle_lib.OracleMech(MECHS["h2o2"])
    g = golden("rates", "h2o2")
    N, K, R = g["dims"]
    T, P, Y = g["T"], g["P"], g["Y"].reshape(N, K)
    kin.set_multiplier(3, 2.5)
    m.set_multiplier(3, 2.5)
    assert kin.multiplier(3) == 2.5
    o = m.rates(T, P, Y, want=("ropf",))
    assert _rel(kin.forward_rates_of_progress(T, P, Y), o["ropf"]) <= TOL
    kin.set_multiplier(3, 1.0)
    with pytest.raises(CtbError):
        kin.set_multiplier(R, 1.0)
    with pytest.raises(CtbError):
        kin.net_production_rates(T, P, Y[:, : K - 1])
    assert kin.net_production_rates(T[:0], P[:0], Y[:0]).shape == (0, K)


def test_full_size_properties(kin_cache):
    """BASELINE config sizes are too large for the oracle: check size-independent properties on 2^20
    gri30 states — element conservation (sum_k wdot_k W_k = 0 to rounding) and invariance of the
    result under batch splitting (same state => same bits wherever it sits in the batch)."""
    import torch
    from cantera_b200.ctbm import read_arrays
    kin = kin_cache("gri30")
    K = kin.n_species
    n = 1 << 20
    gen = torch.Generator(device="cuda").manual_seed(7)
    T = 800 + 1700 * torch.rand(n, dtype=torch.float64, device="cuda", generator=gen)
    P = 101325.0 * (1 + 49 * torch.rand(n, dtype=torch.float64, device="cuda", generator=gen))
    Y = torch.rand(n, K, dtype=torch.float64, device="cuda", generator=gen)
    Y /= Y.sum(dim=1, keepdim=True)
    w = kin.net_production_rates(T, P, Y)
    c = kin.creation_rates(T, P, Y)
    d = kin.destruction_rates(T, P, Y)
    kin.synchronize()
    W = torch.from_numpy(read_arrays(MECHS["gri30"])["mol_weights"]).cuda()
    assert torch.isfinite(w).all()
    mass = (w * W).sum(dim=1).abs()
    scale = ((c + d) * W).sum(dim=1)
    assert float((mass / scale).max()) < 1e-13
    w2 = kin.net_production_rates(T[12345:12345 + 4097], P[12345:12345 + 4097], Y[12345:12345 + 4097])
    kin.synchronize()
    assert torch.equal(w2, w[12345:12345 + 4097])
