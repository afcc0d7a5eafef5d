"""Accuracy vs cost of the batched BDF solver against the converged reference (rtol 1e-11)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import MECHS, golden  # noqa: E402
from cantera_b200 import BatchedGasKinetics, BatchReactorNet  # noqa: E402

settings = [("", ""), ("A0.1", ""), ("A0.01", ""), ("A0.001", ""), ("A0.01", "0.3")]
for case in ("h2o2", "gri30", "nDodecane"):
    g = golden("ignite", case)
    t = golden("ignite_tight", case)
    N, K, R = g["dims"]
    kin = BatchedGasKinetics(MECHS[case])
    tEnd = float(g["tEnd"][0])
    Yt, Yc = t["Yend"].reshape(N, K), g["Yend"].reshape(N, K)
    tol = 1e-6 * np.abs(Yt) + 1e-12
    front = np.isfinite(t["tign"]) & (np.abs(t["tign"] - tEnd) <= 0.05 * tEnd)
    ign = np.isfinite(t["tign"])
    cs = g["stats"].reshape(N, 8).sum(axis=0)
    print("%s cvodes(1e-9): steps %d rhs %d | Yerr/tol %.2e tign %.2e T %.2e" % (
        case, cs[0], cs[1], (np.abs(Yc - Yt) / tol)[~front].max(),
        np.max(np.abs(g["tign"][ign] - t["tign"][ign]) / t["tign"][ign]), np.max(np.abs(g["Tend"] - t["Tend"]) / t["Tend"])))
    for floor, scale in settings:
        os.environ.pop("CTB_BDF_ATOL_SCALE", None)
        if floor.startswith("A"):
            os.environ["CTB_BDF_ATOL_SCALE"] = floor[1:]
        for k, v in (("CTB_BDF_CRATE_FLOOR", "" if floor.startswith("A") else floor), ("CTB_BDF_RTOL_SCALE", scale)):
            if v:
                os.environ[k] = v
            else:
                os.environ.pop(k, None)
        net = BatchReactorNet(kin, N)
        net.set_state(g["T"], g["P"], g["Y"].reshape(N, K))
        net.set_ignition_monitor(float(g["dTign"][0]))
        net.advance(tEnd)
        T, Y = net.get_state()
        st = net.solver_stats().sum(axis=0)
        tig = net.ignition_delay()
        print("  floor=%-4s rtolscale=%-4s steps %d rhs %d setups %d nni %d ncfn %d netf %d | Yerr/tol %.2e tign %.2e T %.2e | vs cvodes Y %.2e" % (
            floor or "0", scale or "1", st[0], st[1], st[3], st[4], st[5], st[6], (np.abs(Y - Yt) / tol)[~front].max(),
            np.max(np.abs(tig[ign] - t["tign"][ign]) / t["tign"][ign]), np.max(np.abs(T - t["Tend"]) / t["Tend"]),
            (np.abs(Y - Yc) / (1e-6 * np.abs(Yc) + 1e-12))[~front].max()))
        net.close()
    kin.close()
