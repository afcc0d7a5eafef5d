"""GPU ReactorNet::eval vs the CPU oracle at states ALONG ignition trajectories (tiny / zero /
slightly negative mole numbers, low temperature) — complements the random-state RHS parity test."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import MECHS, golden  # noqa: E402
from cantera_b200 import BatchedGasKinetics, BatchReactorNet  # noqa: E402
from oracle.oracle_py import OracleMech  # noqa: E402

for case, times in (("nDodecane", (0, 1e-6, 1e-5, 1e-4, 3e-4, 1e-3, 3e-3, 1e-2)), ("gri30", (0, 1e-6, 1e-4, 1e-3))):
    g = golden("ignite", case)
    N, K, R = g["dims"]
    kin = BatchedGasKinetics(MECHS[case])
    ora = OracleMech(MECHS[case])
    net = BatchReactorNet(kin, N)
    net.set_state(g["T"], g["P"], g["Y"].reshape(N, K))
    for t in times:
        if t > 0:
            net.advance(t)
        y = net.get_state_vector()
        yd = net.eval()
        worst = (0, 0, 0)
        for n in range(N):
            r = ora.reactor_rhs(g["P"][n], y[n])
            es = np.max(np.abs(r[1:] - yd[n, 1:])) / max(np.abs(r[1:]).max(), 1e-300)
            eT = abs(r[0] - yd[n, 0]) / max(abs(r[0]), 1e-300)
            if max(es, eT) > worst[0]:
                k = int(np.argmax(np.abs(r[1:] - yd[n, 1:])))
                worst = (max(es, eT), n, k, es, eT, y[n, 0], y[n, 1:].min(), r[1 + k], yd[n, 1 + k])
        print(case, "t=%g worst" % t, worst)
