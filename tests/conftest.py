import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
DATA = os.path.join(ROOT, "cantera_b200", "data")

#: case name -> flat mechanism table
MECHS = {
    "h2o2": os.path.join(DATA, "h2o2.ctbm"),
    "gri30": os.path.join(DATA, "gri30.ctbm"),
    "nDodecane": os.path.join(DATA, "nDodecane_Reitz.ctbm"),
}
for _t in ("pdep-test", "sri-falloff", "tsang-falloff", "jacobian-tests", "chemically-activated-reaction",
           "explicit-third-bodies", "reaction-orders"):
    MECHS[_t] = os.path.join(GOLDEN, "mech_%s.ctbm" % _t)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden(kind, case):
    from cantera_b200.ctbm import read_arrays
    a = read_arrays(os.path.join(GOLDEN, "%s_%s.ctbm.gz" % (kind, case)))
    if len(a["dims"]) > 3:  # rhs fixtures carry the number of equations of the reactor type
        a["neq"] = int(a["dims"][3])
        a["dims"] = a["dims"][:3]
    return a


def pdep_state(T, P):
    """The state of the reference's PdepTest fixture (test/kinetics/pdep.cpp:23-36): equal mole
    fractions of H, R1A, R1B, R2..R6 at (T, P), as mass fractions."""
    import numpy as np
    from cantera_b200.ctbm import read_arrays
    tab = read_arrays(MECHS["pdep-test"])
    names = tab["species_names"].split("\n")
    X = np.array([1.0 if n in ("H", "R1A", "R1B", "R2", "R3", "R4", "R5", "R6") else 0.0 for n in names])
    Y = X * tab["mol_weights"]
    return np.array([float(T)]), np.array([float(P)]), (Y / Y.sum())[None, :]


#: (T, P, quantity, {reaction index: (expected, abs tolerance)}) of test/kinetics/pdep.cpp:119-222
#: (Chemkin / RMG-py numbers quoted by the reference)
PDEP_KATS = [
    (1100.0, 20 * 101325.0, "ropf", {0: (3.100682e+05, 1e2), 1: (2.006871e+05, 1e2), 2: (4.468658e+06, 1e2), 3: (1.774796e+06, 1e2)}),
    (1100.0, 0.5 * 101325.0, "ropf", {0: (5.244649e+02, 5e-2), 1: (2.252537e+02, 2e-2), 2: (2.985338e+03, 3e-1), 3: (1.109248e+03, 1e-1)}),
    (800.0, 70 * 101325.0, "ropf", {0: (2.274501e+04, 1e+1), 1: (2.307191e+05, 1e+2), 2: (2.224601e+07, 1e+3), 3: (1.007440e+07, 1e+3)}),
    (1100.0, 20 * 101325.0, "kf", {4: (3.130698657e+06, 1e-1), 5: (1.187949573e+00, 1e-7)}),
    (400.0, 0.1 * 101325.0, "kf", {4: (1.713599902e+05, 1e-3), 5: (9.581780687e-24, 1e-31)}),
    (1100.0, 30 * 101325.0, "ropf", {4: (4.552930e+03, 1e-1), 5: (4.877390e-02, 1e-5)}),
    (500.0, 1000.0, "kf", {4: (1.225785655e+06, 1e-2)}),
    (500.0, 1.0e7, "kf", {4: (1.580981157e+03, 1e-5)}),
    (300.0, 101325.0, "kf", {4: (5.405987017e+03, 1e-5)}),
    (2000.0, 101325.0, "kf", {4: (3.354054351e+07, 1e-1)}),
]


def check_pdep_kats(evaluate):
    """evaluate(T, P, Y, quantity) -> [1, R] array; asserts the reference's known answers, the
    Plog corner cases (pdep.cpp:100-117) and the molec/mol Chebyshev twin (:203-205)."""
    import numpy as np
    for T, P, what, exp in PDEP_KATS:
        got = evaluate(*pdep_state(T, P), what)[0]
        for i, (val, tol) in exp.items():
            assert abs(got[i] - val) <= tol, (T, P, what, i, got[i], val)
    T, P = 500.0, 101325.0
    kf = evaluate(*pdep_state(T, P), "kf")[0]
    RT = 1.98720425864083 * T  # GasConst_cal_mol_K

    def k(A, n, Ea):
        return A * T ** n * np.exp(-Ea / RT)
    kf0 = k(4.910800e+28, -4.8507, 24772.8)
    kf1 = k(1.2600e+17, -1.83, 15003.0) + k(1.2300e+01, 2.68, 6335.0)
    kf2 = k(3.4600e+9, 0.442, 5463.0)
    for got, ref in ((kf[0], kf0), (kf[1], kf1), (kf[2], kf2)):
        assert abs(got - ref) <= 1e-9 * ref
    for T, P, tol in ((1100.0, 20 * 101325.0, 5e2), (400.0, 0.1 * 101325.0, 1e2)):
        kf = evaluate(*pdep_state(T, P), "kf")[0]
        assert abs(kf[4] - kf[6] / (6.02214076e26 * 1e-3)) <= tol


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import oracle_py
    oracle_py.build()
    return oracle_py
