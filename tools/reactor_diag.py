"""Diagnostics for the batched BDF solver on a GPU box: accuracy and work vs the reference golden
ignition cases, and throughput for each group width (CTB_BDF_NT)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import MECHS, golden  # noqa: E402
from cantera_b200 import BatchedGasKinetics, BatchReactorNet  # noqa: E402

names = ["steps", "rhs", "jac", "setups", "nni", "ncfn", "netf", "q"]
for case, kind in (("h2o2", "ignite"), ("gri30", "ignite"), ("gri30", "substep"), ("nDodecane", "ignite")):
    g = golden(kind, case)
    N, K, R = g["dims"]
    kin = BatchedGasKinetics(MECHS[case])
    net = BatchReactorNet(kin, N)
    net.set_state(g["T"], g["P"], g["Y"].reshape(N, K))
    dT = float(g["dTign"][0])
    if dT > 0:
        net.set_ignition_monitor(dT)
    t0 = time.perf_counter()
    net.advance(float(g["tEnd"][0]))
    T, Y = net.get_state()
    dt = time.perf_counter() - t0
    st = net.solver_stats()
    ref = g["stats"].reshape(N, 8)
    tig, tref = net.ignition_delay(), g["tign"]
    ign = np.isfinite(tref) & np.isfinite(tig)
    eT = np.abs(T - g["Tend"]) / g["Tend"]
    eY = np.abs(Y - g["Yend"].reshape(N, K)) / (1e-6 * np.abs(g["Yend"].reshape(N, K)) + 1e-12)
    print("== %s %s N=%d  %.3fs  status=%s" % (case, kind, N, dt, np.bincount(net.status()).tolist()))
    print("   ours  " + " ".join("%s=%d" % (n, v) for n, v in zip(names, st.sum(axis=0))))
    print("   cvodes" + " ".join(" %s=%d" % (n, v) for n, v in zip(names, ref.sum(axis=0))))
    print("   max relT %.2e  max Yerr/tol %.2e  tign rel %s  (nan mismatch %d)" % (
        eT.max(), eY.max(), ("%.2e" % np.max(np.abs(tig[ign] - tref[ign]) / tref[ign])) if ign.any() else "-",
        int(np.sum(np.isfinite(tref) != np.isfinite(tig)))))
    if kind == "ignite":
        t = golden("ignite_tight", case)
        Yt = t["Yend"].reshape(N, K)
        tol = 1e-6 * np.abs(Yt) + 1e-12
        front = np.isfinite(tref) & (np.abs(tref - float(g["tEnd"][0])) <= 0.05 * float(g["tEnd"][0]))
        eo = (np.abs(Y - Yt) / tol)[~front]
        er = (np.abs(g["Yend"].reshape(N, K) - Yt) / tol)[~front]
        print("   vs converged reference (rtol 1e-11): ours Yerr/tol %.2e  cvodes(1e-9) Yerr/tol %.2e; T ours %.2e cvodes %.2e" % (
            eo.max(), er.max(), np.max(np.abs(T - t["Tend"]) / t["Tend"]), np.max(np.abs(g["Tend"] - t["Tend"]) / t["Tend"])))
        rr, kk = np.unravel_index(np.argmax(np.abs(Y - Yt) / tol), Y.shape)
        print("   worst: reactor %d species %d Y ours %.6e conv %.6e cvodes %.6e  T0 %.0f tign %.3g" % (
            rr, kk, Y[rr, kk], Yt[rr, kk], g["Yend"].reshape(N, K)[rr, kk], g["T"][rr], tref[rr]))
    worst = np.argsort(-eT)[:4]
    for n in worst:
        print("     r%d T0=%.0f P=%.1fatm Tend ref %.2f ours %.2f tign ref %.4g ours %.4g steps %d/%d" % (
            n, g["T"][n], g["P"][n] / 101325, g["Tend"][n], T[n], tref[n], tig[n], st[n, 0], ref[n, 0]))
    net.close()
    kin.close()

# throughput per group width
sys.path.insert(0, ROOT)
from bench import make_ignition_states  # noqa: E402
kin = BatchedGasKinetics("gri30")
for n in (16384,):
    Tn, Pn, Yn = make_ignition_states(None, n, 5, None)
    for nt in ("64", "128"):
        os.environ["CTB_BDF_NT"] = nt
        net = BatchReactorNet(kin, n)
        net.set_state(Tn, Pn, Yn)
        kin.synchronize()
        t0 = time.perf_counter()
        net.advance(1e-3)
        kin.synchronize()
        dt = time.perf_counter() - t0
        st = net.solver_stats().sum(axis=0)
        print("gri30 ignition n=%d NT=%s: %.3f s  %.0f reactors/s  steps/reactor %.0f rhs %.0f setups %.0f fails %s" % (
            n, nt, dt, n / dt, st[0] / n, st[1] / n, st[3] / n, np.bincount(net.status()).tolist()))
        net.close()
