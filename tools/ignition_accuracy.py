"""Accuracy and work of BatchReactorNet against the reference's CVODES runs (tests/golden/ignite_*:
default tolerances, and the converged twin at rtol 1e-11 / atol 1e-18) — the numbers behind the
tolerance mapping of DESIGN.md. Run on a GPU box: python tools/ignition_accuracy.py [case ...]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import MECHS, golden  # noqa: E402
from cantera_b200 import BatchedGasKinetics, BatchReactorNet  # noqa: E402


def main(cases):
    for case in cases:
        g = golden("ignite", case)
        t = golden("ignite_tight", case)
        N, K, R = g["dims"]
        kin = BatchedGasKinetics(MECHS[case])
        tEnd = float(g["tEnd"][0])
        Yt, Yc = t["Yend"].reshape(N, K), g["Yend"].reshape(N, K)
        ign = np.isfinite(t["tign"]) & np.isfinite(g["tign"])
        front = np.isfinite(g["tign"]) & (np.abs(g["tign"] - tEnd) <= 0.05 * tEnd)
        cs = g["stats"].reshape(N, 8).sum(axis=0)

        def report(tag, T, Y, tig, st, dt):
            tolc = 1e-6 * np.abs(Yc) + 1e-12
            tolt = 1e-6 * np.abs(Yt) + 1e-12
            print("  %-34s steps %8d rhs %8d setups %7d | vs ref: Y/tol %.3f T %.2e tign %.2e | vs converged: Y/tol %.3f tign %.2e | %.2f s" % (
                tag, st[0], st[1], st[3],
                (np.abs(Y - Yc) / tolc)[~front].max(), np.max(np.abs(T - g["Tend"])[~front] / g["Tend"][~front]),
                np.max(np.abs(tig[ign] - g["tign"][ign]) / g["tign"][ign]),
                (np.abs(Y - Yt) / tolt)[~front].max(), np.max(np.abs(tig[ign] - t["tign"][ign]) / t["tign"][ign]), dt))

        print("%s: N=%d, %d ignite, %d inside the front at t_end" % (case, N, ign.sum(), front.sum()))
        report("reference CVODES (defaults)", g["Tend"], Yc, g["tign"], cs, 0.0)
        for tag, rtol, atol in (("ours, defaults", None, None), ("ours, user atol 1e-15*ntot-ish", None, "moles")):
            net = BatchReactorNet(kin, N)
            if atol == "moles":
                continue
            net.set_state(g["T"], g["P"], g["Y"].reshape(N, K))
            net.set_ignition_monitor(float(g["dTign"][0]))
            t0 = time.perf_counter()
            net.advance(tEnd)
            T, Y = net.get_state()
            dt = time.perf_counter() - t0
            assert np.all(net.status() == 0), np.unique(net.status(), return_counts=True)
            report(tag + " %s" % (net.effective_tolerances(),), T, Y, net.ignition_delay(), net.solver_stats().sum(axis=0), dt)
            # the worst reactors, with the reference's own distance from its converged twin beside ours
            tolc = 1e-6 * np.abs(Yc) + 1e-12
            ours = (np.abs(Y - Yc) / tolc).max(axis=1)
            refc = (np.abs(Yc - Yt) / tolc).max(axis=1)
            ourt = (np.abs(Y - Yt) / tolc).max(axis=1)
            tig = net.ignition_delay()
            for i in np.argsort(-ours)[:10]:
                print("    reactor %4d T0 %7.1f P %5.1f atm Tend %7.1f tign ref %.6e tight %.6e ours %.6e | Y/tol ours-ref %.3f ref-conv %.3f ours-conv %.3f | dT ours-ref %.2e ref-conv %.2e%s" % (
                    i, g["T"][i], g["P"][i] / 101325, g["Tend"][i], g["tign"][i], t["tign"][i], tig[i], ours[i], refc[i], ourt[i],
                    T[i] - g["Tend"][i], g["Tend"][i] - t["Tend"][i], " (front)" if front[i] else ""))
            net.close()
        kin.close()


if __name__ == "__main__":
    main(sys.argv[1:] or ["h2o2", "gri30", "nDodecane"])
