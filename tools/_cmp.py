import os, sys, numpy as np
sys.path.insert(0, os.getcwd()); sys.path.insert(0, "tests")
from conftest import MECHS, golden
from cantera_b200 import BatchedGasKinetics, BatchReactorNet
for case in ("gri30", "nDodecane"):
    kin = BatchedGasKinetics(MECHS[case])
    g = golden("ignite", case)
    N, K, R = g["dims"]
    net = BatchReactorNet(kin, N)
    os.environ["CTB_RHS_KERNEL"] = "group"
    net.set_state(g["T"], g["P"], g["Y"].reshape(N, K))
    for t in (0.0, 1e-8, 1e-6, 0.3, 0.6, 1.0):
        if t > 0: net.advance(t * float(g["tEnd"][0]))
        os.environ["CTB_RHS_KERNEL"] = "group"
        fg = net.eval()
        os.environ.pop("CTB_RHS_KERNEL")
        ft = net.eval()
        os.environ["CTB_RHS_KERNEL"] = "group"
        y = net.get_state_vector()
        sc = np.abs(fg[:, 1:]).max(axis=1, keepdims=True) + 1e-300
        d = np.abs(ft - fg)
        rel = d[:, 1:] / sc
        i, k = np.unravel_index(np.argmax(rel), rel.shape)
        print(case, "t", t, "status", (net.status() != 0).sum(), "max species err/scale %.3e at reactor %d species %d (f tile %.6e group %.6e, y %.3e, T %.1f) ; T row rel %.3e ; nan tile %d group %d" % (
            rel.max(), i, k, ft[i, 1 + k], fg[i, 1 + k], y[i, 1 + k], y[i, 0], np.nanmax(d[:, 0] / (np.abs(fg[:, 0]) + 1e-300)), np.isnan(ft).sum(), np.isnan(fg).sum()))
