"""TEST INFRASTRUCTURE — numpy restatement of the reference's mixture-averaged transport
(src/transport/MixTransport.cpp:33-50,175-202, GasTransport.cpp:28-58,60-121,123-146,170-201,288-294), one state
at a time, pinned by tests/test_transport.py against tests/golden/trans_*.ctbm.gz (outputs of the reference
itself). Only tests may import this module."""
import numpy as np

GAS_CONSTANT = 6.02214076e26 * 1.380649e-23
TINY = 1e-20


def mixture_transport(tab, T, P, Y):
    """tab: mol_weights [K], visc_poly / cond_poly [K*5], diff_poly [K(K+1)/2*5] (GasTransport's non-CK fits).
    -> viscosity [N], thermal conductivity [N], mixture-averaged diffusion coefficients [N, K]."""
    mw = np.asarray(tab["mol_weights"])
    K = mw.size
    vc = np.asarray(tab["visc_poly"]).reshape(K, 5)
    cc = np.asarray(tab["cond_poly"]).reshape(K, 5)
    dc = np.asarray(tab["diff_poly"]).reshape(-1, 5)
    iu = np.triu_indices(K)
    # GasTransport.cpp:288-294
    wratjk = np.zeros((K, K)); wratkj1 = np.zeros((K, K))
    for j in range(K):
        for k in range(j, K):
            wratjk[j, k] = np.sqrt(mw[j] / mw[k])
            wratjk[k, j] = np.sqrt(wratjk[j, k])
            wratkj1[j, k] = np.sqrt(1.0 + mw[k] / mw[j])
    N = len(T)
    mu, lam, dmix = np.empty(N), np.empty(N), np.empty((N, K))
    for n in range(N):
        y = np.maximum(np.asarray(Y[n], dtype=float), 0.0)
        y = y / y.sum()
        ym = y * (1.0 / mw)                                         # Phase: m_ym = y / W through 1 / W (Phase.cpp:342-360)
        mmw = 1.0 / ym.sum()
        x = ym * mmw
        x = np.maximum(x, TINY)                                     # MixTransport::update_C
        logt = np.log(T[n]); sqt = np.sqrt(T[n]); t14 = np.sqrt(sqt)
        poly = np.array([1.0, logt, logt * logt, logt * logt * logt, logt * logt * logt * logt])
        sqv = t14 * (vc @ poly)                                     # the fit is for sqrt(visc / sqrt(T))
        visc = sqv * sqv
        phi = np.zeros((K, K))
        for j in range(K):                                          # updateViscosity_T (Poling et al. eq. 9-5.14)
            for k in range(j, K):
                f1 = 1.0 + (sqv[k] / sqv[j]) * wratjk[k, j]
                phi[k, j] = f1 * f1 / (np.sqrt(8.0) * wratkj1[j, k])
                phi[j, k] = phi[k, j] / ((visc[k] / visc[j]) * (mw[j] / mw[k]))
        mu[n] = np.sum(x * visc / (phi @ x))
        cond = sqt * (cc @ poly)
        lam[n] = 0.5 * (np.sum(x * cond) + 1.0 / np.sum(x / cond))
        bd = np.zeros((K, K))
        bd[iu] = T[n] * sqt * (dc @ poly)
        bd = np.triu(bd) + np.triu(bd, 1).T
        p = P[n]
        for k in range(K):
            others = np.arange(K) != k
            s2 = np.sum(x[others] / bd[others, k])
            dmix[n, k] = bd[k, k] / p if s2 <= 0.0 else (mmw - x[k] * mw[k]) / (p * mmw * s2)
    return mu, lam, dmix
