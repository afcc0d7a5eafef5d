// reactor_thread.cuh — the IdealGasConstPressureMoleReactor right-hand side for ONE THREAD per
// reactor, on the [species][thread] shared-memory layout of
// rates_thread.cuh. Same expressions as reactorRhs of bdf.cuh (the cooperative-group version);
// only the mixture sums run sequentially over the species, like the reference's own loops.
#pragma once

#include "rates_thread.cuh"

namespace ctb {

//! IdealGasConstPressureMoleReactor::updateState + eval, ReactorNet::eval
//! (IdealGasConstPressureMoleReactor.cpp:38-106, ReactorNet.cpp:617-635).
//! On entry a.c[k] holds the moles n_k; on exit a.w[k] = dn_k/dt. Returns dT/dt.
template <int U>
CTB_HD double reactorRhsT(const MechView& v, const unsigned char* sm, const double T, const double P,
                          const TArr& a)
{
    const int K = v.K, st = a.st;
    const double* molwt = CTB_TAB(double, sm, v.o_molwt);
    double totalMoles = 0.0, totalMass = 0.0;
    for (int k = 0; k < K; k++) {
        const double nk = a.c[k * st];
        totalMoles += nk;
        totalMass += nk * molwt[k];
    }
    const double mmw = totalMass / totalMoles;
    const double f2 = 1.0 / (mmw * totalMoles);
    const double dens = P * mmw / (T * CTB_GAS_CONSTANT);
    GasScalars s;
    s.T = T;
    s.ctot = dens / mmw;
    s.P = CTB_GAS_CONSTANT * s.ctot * T;
    fillScalarsT(v, s);
    const double tmp = s.logPref * s.RT;
    double sumq = 0.0;
    for (int k = 0; k < K; k++) {
        double cp, h, sr;
        nasaEval(v, sm, k, s.T, s.logT, s.recipT, cp, h, sr);
        a.g[k * st] = (h - sr) * s.RT + tmp;
        const double ym = a.c[k * st] * f2;
        sumq += ym * cp;
        a.c[k * st] = dens * ym;
        a.w[k * st] = 0.0;
    }
    reactionsT<true, U>(v, sm, s, a);
    const double vol = totalMass / dens;
    const double cp_mole = CTB_GAS_CONSTANT * (mmw * sumq);
    const double totalCp = totalMass * (cp_mole / mmw);
    double mcpdTdt = 0.0;
    for (int k = 0; k < K; k++) {
        double cp, h, sr;
        nasaEval(v, sm, k, s.T, s.logT, s.recipT, cp, h, sr);
        const double wdot = a.w[k * st];
        const double hk = h * s.RT;
        mcpdTdt -= wdot * hk * vol;
        a.w[k * st] = wdot * vol;
    }
    return mcpdTdt / totalCp;
}

} // namespace ctb
