// derivs_kernels.cu — the remaining kinetics getters on the intermediates of updateROP (SURVEY.md §8(f)
// rank 2): the _ddT / _ddP / _ddC families of rate constants, rates of progress and production rates
// (BulkKinetics.cpp:284-385,600-655, Kinetics.cpp:459-602) and getDelta{,SS}{Gibbs,Enthalpy,Entropy}
// (BulkKinetics.cpp:200-257). A warp per state; not a throughput kernel (these getters feed
// user-side Jacobians and diagnostics), so it stays on the lane-group device code of rates.cuh.
//
// What the reference does, and what is reproduced here term by term:
//   X_ddT = X * s_T    s_T = (b + Ea/RT)/T for Arrhenius rates (Arrhenius.h:205-207), else the forward
//                      difference (k(T (1 + d)) / k(T) - 1) / (T d) at fixed [M] (MultiRate.h:100-118);
//           ropr and ropnet also carry ropr * (d ln(1/Kc)/dT) with DeltaG0 differenced at T (1 + d)
//           (BulkKinetics.cpp:571-604);
//   X_ddP = X * s_P    forward difference at P (1 + d) for Plog / Chebyshev, 0 otherwise (MultiRate.h:120-143);
//   X_ddC              d/d c_tot at fixed T and X: X * (sum of orders [+ 1 for a mass-action third body]) / c_tot
//                      plus, for falloff rates, X / c_tot * (k([M] (1 + d)) / k - 1) / d (BulkKinetics.cpp:622-655);
//   d = rtol-delta of setDerivativeSettings (default 1e-8).
#define CTB_EXP_LIB
#include "derivs_kernels.h"
#include "rates.cuh"

#include <stdexcept>

namespace ctb {

namespace {

//! Rate constant of a non-Arrhenius reaction at (possibly perturbed) scalars and third-body concentration.
__device__ __forceinline__ double evalSpecialRate(const MechView& v, const unsigned char* sm, const RxnRec& r, int e,
                                                  const GasScalars& s, double concm)
{
    const int a = CTB_TAB(int16_t, sm, v.o_auxIdx)[e];
    if (r.cls <= CLS_TSANG) {
        return falloffRate(r.cls, r.flags & CTB_RXN_CHEM_ACTIVATED, CTB_TAB(double, sm, v.o_fallLow) + 3 * a,
                           CTB_TAB(double, sm, v.o_fallCoef) + 5 * a, CTB_TAB(double, sm, v.o_A)[e], CTB_TAB(double, sm, v.o_b)[e],
                           CTB_TAB(double, sm, v.o_EaR)[e], s, concm, nullptr);
    }
    return r.cls == CLS_PLOG ? plogRate(v, sm, a, s) : chebyshevRate(v, sm, a, s);
}

__device__ __forceinline__ double sideOrders(const MechView& v, const unsigned char* sm, int n, const uint16_t* slot, bool useOrder)
{
    if (n <= 3) return (double) n;
    const int32_t* ptr = CTB_TAB(int32_t, sm, v.o_genPtr);
    const double* ex = CTB_TAB(double, sm, useOrder ? v.o_genOrder : v.o_genStoich);
    double acc = 0.0;
    for (int i = ptr[slot[0]]; i < ptr[slot[0] + 1]; i++) acc += ex[i];
    return acc;
}

__global__ void __launch_bounds__(128) derivsKernel(const MechView v, const unsigned char* __restrict__ gmech, const DerivArgs a,
                                                    const int workDoubles)
{
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + v.hotBytes);
    tmaStage(smem, gmech, v.hotBytes, bar);
    const unsigned char* sm = smem;
    using G = SubWarp<32>;
    const G g;
    const int K = v.K, R = v.R;
    const int warp = threadIdx.x >> 5, warps = blockDim.x >> 5;
    double* work = reinterpret_cast<double*>(smem + v.hotBytes + 16) + (size_t) warp * workDoubles;
    double* c = work;            // [K] concentrations
    double* mu0 = c + K;         // [K] standard chemical potentials at (T, P)
    double* hk = mu0 + K;        // [K] h_k / RT
    double* sk = hk + K;         // [K] s_k / R (reference pressure)
    double* mu1 = sk + K;        // [K] standard chemical potentials at (T (1 + d), P)
    double* ropf = mu1 + K;      // [R] execution order
    double* ropr = ropf + R;
    double* rfn = ropr + R;
    double* rkc = rfn + R;
    double* concm = rkc + R;
    double* dF[3] = {concm + R, concm + 2 * R, concm + 3 * R};          // ropf_ddT, _ddP, _ddC
    double* dR[3] = {concm + 4 * R, concm + 5 * R, concm + 6 * R};      // ropr_ddT, _ddP, _ddC
    const int32_t* perm = CTB_TAB(int32_t, gmech, v.o_perm);
    const double delta = a.rtolDelta;
    for (long long n = (long long) blockIdx.x * warps + warp; n < a.N; n += (long long) gridDim.x * warps) {
        GasScalars s;
        stateFromTPY(g, v, sm, a.T[n], a.P[n], a.Y + n * a.ldY, c, s);
        // perturbed temperature at the same pressure (ThermoPhase::setState_TP(T (1 + d), P))
        GasScalars s1 = s;
        s1.T = s.T * (1.0 + delta);
        s1.logT = log(s1.T); s1.recipT = 1.0 / s1.T; s1.RT = s1.T * CTB_GAS_CONSTANT; s1.rrt = 1.0 / s1.RT;
        GasScalars sP = s;
        sP.P = s.P * (1.0 + delta);
        sP.logP = log(sP.P);
        const double tmp = s.logPref * s.RT, tmp1 = s.logPref * s1.RT;
        for (int k = g.lane(); k < K; k += 32) {
            double cp, h, sr;
            nasaEval(v, sm, k, s.T, s.logT, s.recipT, cp, h, sr);
            mu0[k] = (h - sr) * s.RT + tmp;
            hk[k] = h;
            sk[k] = sr;
            nasaEval(v, sm, k, s1.T, s1.logT, s1.recipT, cp, h, sr);
            mu1[k] = (h - sr) * s1.RT + tmp1;
        }
        g.sync();
        RopOut o = {};
        o.ropf = ropf; o.ropr = ropr; o.rfn = rfn; o.rkc = rkc; o.concm = concm;
        ropPhase(g, v, sm, s, c, mu0, o);
        g.sync();
        const double rctot = 1.0 / s.ctot, Tinv = s.recipT;
        for (int e = g.lane(); e < R; e += 32) {
            union { uint4 u; RxnRec r; } rr;
            rr.u = CTB_TAB(uint4, sm, v.o_rec)[e];
            const RxnRec& r = rr.r;
            const int i = perm[e];
            const double mult = CTB_TAB(double, sm, v.o_mult)[e];
            double sT, sPd = 0.0, sM = 0.0;
            if (r.cls == CLS_ARR_CONST || r.cls == CLS_ARR) {
                sT = (CTB_TAB(double, sm, v.o_EaR)[e] * s.recipT + CTB_TAB(double, sm, v.o_b)[e]) * s.recipT;
            } else {
                const double kf0 = mult != 0.0 ? rfn[e] / mult : evalSpecialRate(v, sm, r, e, s, concm[e]);
                sT = 0.0;
                if (kf0 != 0.0) {
                    sT = (evalSpecialRate(v, sm, r, e, s1, concm[e]) / kf0 - 1.0) / (s.T * delta);
                    if (r.cls >= CLS_PLOG) sPd = (evalSpecialRate(v, sm, r, e, sP, concm[e]) / kf0 - 1.0) / (s.P * delta);
                    if (r.cls <= CLS_TSANG && concm[e] > 0.0 && !a.skipThirdBodies && !a.skipFalloff) {
                        sM = (evalSpecialRate(v, sm, r, e, s, concm[e] * (1.0 + delta)) / kf0 - 1.0) / delta;
                    }
                }
            }
            // d ln(1/Kc) / dT (BulkKinetics.cpp:571-604)
            double fK = 0.0;
            double dg0 = 0.0 + sideSum(v, sm, r.np, r.p, mu0) - sideSum(v, sm, r.nr, r.r, mu0);
            if (r.flags & CTB_RXN_REVERSIBLE) {
                const double dg1 = 0.0 + sideSum(v, sm, r.np, r.p, mu1) - sideSum(v, sm, r.nr, r.r, mu1);
                fK = (dg1 - dg0) * (s.rrt * Tinv / delta) + CTB_TAB(double, sm, v.o_dn)[e] * Tinv - dg0 * (s.rrt * Tinv);
            }
            const double f = ropf[e], rv = ropr[e];
            // ddC factors (BulkKinetics.cpp:622-655)
            const bool tb = (r.flags & CTB_RXN_THIRD_BODY) && !a.skipThirdBodies;
            const double ma = (tb && (r.flags & CTB_RXN_MASS_ACTION)) ? 1.0 : 0.0;
            const double cF = (sideOrders(v, sm, r.nr, r.r, true) + ma + sM) * rctot;
            const double cR = (r.flags & CTB_RXN_REVERSIBLE) ? (sideOrders(v, sm, r.np, r.p, false) + ma + sM) * rctot : 0.0;
            dF[0][e] = f * sT; dF[1][e] = f * sPd; dF[2][e] = f * cF;
            dR[0][e] = rv * sT + rv * fK; dR[1][e] = rv * sPd; dR[2][e] = rv * cR;
            const long long o0 = n * a.ldR + i;
            if (a.kf[0]) a.kf[0][o0] = rfn[e] * sT;
            if (a.kf[1]) a.kf[1][o0] = rfn[e] * sPd;
            if (a.kf[2]) a.kf[2][o0] = rfn[e] * (sM * rctot);
            for (int d = 0; d < 3; d++) {
                if (a.ropf[d]) a.ropf[d][o0] = dF[d][e];
                if (a.ropr[d]) a.ropr[d][o0] = dR[d][e];
            }
            if (a.ropnet[0]) a.ropnet[0][o0] = (f - rv) * sT - rv * fK;
            if (a.ropnet[1]) a.ropnet[1][o0] = (f - rv) * sPd;
            if (a.ropnet[2]) a.ropnet[2][o0] = dF[2][e] - dR[2][e];
            // reaction deltas (getReactionDelta: every reaction)
            if (a.delta[3]) a.delta[3][o0] = dg0; // deltaSSGibbs
            if (a.delta[0] || a.delta[1] || a.delta[2] || a.delta[4] || a.delta[5]) {
                double dH = 0.0, dS0 = 0.0, dLnX = 0.0, dNu = 0.0;
                auto side = [&](int nn, const uint16_t* slot, double sign) {
                    if (nn == SIDE_GENERAL) {
                        const int32_t* ptr = CTB_TAB(int32_t, sm, v.o_genPtr);
                        const uint16_t* sp = CTB_TAB(uint16_t, sm, v.o_genSp);
                        const double* st = CTB_TAB(double, sm, v.o_genStoich);
                        for (int q = ptr[slot[0]]; q < ptr[slot[0] + 1]; q++) {
                            const int k = sp[q];
                            dH += sign * st[q] * hk[k]; dS0 += sign * st[q] * sk[k]; dNu += sign * st[q];
                            dLnX += sign * st[q] * log(fmax(c[k] * rctot, CTB_SMALL));
                        }
                    } else {
                        for (int q = 0; q < nn; q++) {
                            const int k = slot[q];
                            dH += sign * hk[k]; dS0 += sign * sk[k]; dNu += sign;
                            dLnX += sign * log(fmax(c[k] * rctot, CTB_SMALL));
                        }
                    }
                };
                side(r.np, r.p, 1.0);
                side(r.nr, r.r, -1.0);
                if (a.delta[0]) a.delta[0][o0] = dg0 + s.RT * dLnX;                                         // deltaGibbs
                if (a.delta[1]) a.delta[1][o0] = dH * s.RT;                                                 // deltaEnthalpy
                if (a.delta[2]) a.delta[2][o0] = CTB_GAS_CONSTANT * (dS0 - dNu * s.logPref - dLnX);         // deltaEntropy
                if (a.delta[4]) a.delta[4][o0] = dH * s.RT;                                                 // deltaSSEnthalpy
                if (a.delta[5]) a.delta[5][o0] = CTB_GAS_CONSTANT * (dS0 - dNu * s.logPref);                // deltaSSEntropy
            }
        }
        g.sync();
        // species rows: creation = nu'' ropf_dd + nu' ropr_dd, destruction = nu' ropf_dd + nu'' ropr_dd
        for (int d = 0; d < 3; d++) {
            if (!a.cdot[d] && !a.ddot[d] && !a.wdot[d]) continue;
            for (int k = g.lane(); k < K; k += 32) {
                const double cr = gatherSpecies<1>(v, sm, gmech, k, dF[d], dR[d]);
                const double ds = gatherSpecies<2>(v, sm, gmech, k, dF[d], dR[d]);
                if (a.cdot[d]) a.cdot[d][n * a.ldW + k] = cr;
                if (a.ddot[d]) a.ddot[d][n * a.ldW + k] = ds;
                if (a.wdot[d]) a.wdot[d][n * a.ldW + k] = cr - ds;
            }
        }
        g.sync();
    }
}

} // namespace

void launchDerivs(const DeviceMech& m, const DerivArgs& a, cudaStream_t stream)
{
    if (a.N <= 0) return;
    const MechView& v = m.view;
    const int workDoubles = (5 * v.K + 11 * v.R + 2 + 1) & ~1;
    const size_t fixed = v.hotBytes + 16;
    const size_t budget = (size_t) m.smemOptIn - 1024;
    if (fixed + (size_t) workDoubles * 8 > budget) throw std::runtime_error("mechanism too large for the derivatives kernel");
    const int warps = (int) std::min<size_t>(4, (budget - fixed) / ((size_t) workDoubles * 8));
    const size_t smemBytes = fixed + (size_t) warps * workDoubles * 8;
    cudaFuncSetAttribute(derivsKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smemBytes);
    const int grid = (int) std::max<long long>(1, std::min<long long>((a.N + warps - 1) / warps, 2LL * m.numSMs));
    derivsKernel<<<grid, warps * 32, smemBytes, stream>>>(v, m.dBlob, a, workDoubles);
}

} // namespace ctb
