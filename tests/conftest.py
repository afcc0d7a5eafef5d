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
    return read_arrays(os.path.join(GOLDEN, "%s_%s.ctbm.gz" % (kind, case)))


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import oracle_py
    oracle_py.build()
    return oracle_py
