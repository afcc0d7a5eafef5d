// jac_tile.cuh — the composition Jacobian d wdot_k / d c_j (BulkKinetics::netRatesOfProgress_ddCi
// contracted with the stoichiometric matrix: Kinetics.cpp:604-612, BulkKinetics.cpp:408-455,662-701,
// StoichManager.h:227-430, ThirdBodyCalc.h:97-100, MultiRate.h:135-165) for a TILE of 16 states.
//
// Two phases on top of the tile layout of rates_tile.cuh:
//   1. the reaction pass stores, per reaction and state, the forward and reverse COEFFICIENTS
//      kfM = kf [M]^(mass action), krM = kfM / Kc, the third-body weight we = d q / d [M], and
//      h2 = we default - h / c_tot with h = sum_j c_j dq/dc_j (q is homogeneous in c: h = orders ropf -
//      orders ropr + we [M]); the rows that do not fit in shared memory live in a small per-CTA
//      scratch in global memory that never leaves the L2;
//   2. the column pass: reaction lane l owns whole columns j (dealt by item count). A column is
//      zeroed in the lane's private shared-memory buffer [K][16 states], every (reaction, role) item
//      of species j adds its derivative times the reaction's stoichiometric column (no atomics: the
//      buffer is private, the half-warp's 16 states sit in 16 different banks), the explicit
//      third-body efficiencies follow, and the finished column leaves through a transposed,
//      coalesced write. The dense default-efficiency part (and, for the reactor, the volume-dilution
//      part) is one more species gather, D_k = sum_e nu_ke h2_e, added to every column.
// The group kernels of rates.cuh needed shared-memory atomicAdd(double) to balance this work, which
// sm_100a implements as a compare-and-swap loop (profiles/r02_sass_summary.txt).
#pragma once

#include "reactor_tile.cuh"

namespace ctb {

//! Global (L2-resident) scratch rows of one tile, already offset to the thread's state column.
struct JacRows {
    double* kfM; // [R][16]
    double* krM; // [R][16]
    double* h2;  // [R+1][16], row R = 0
    double* we;  // [nTb][16]
};
//! doubles of scratch per tile
__host__ __device__ inline size_t jacScratchDoubles(int R, int nTb)
{
    return ((size_t) 3 * R + 1 + (size_t) nTb) * CTB_TS;
}

//! C1/C2/C3::derivatives (StoichManager.h:227-236,300-312) and C_AnyN::derivatives (:406-430)
//! restricted to species j (byte offset joff of its row): d(r prod_side c^order) / d c_j.
__device__ __forceinline__ double sideDerivTile(const TileView& tv, const Tile& t, int n, const uint32_t* slot, double r,
                                                int j, uint32_t joff)
{
    double d = 0.0;
    if (n == 1) {
        d = r;
    } else if (n == 2) {
        const double a = ldOff(t.cS, slot[0]), b = ldOff(t.cS, slot[1]);
        if (slot[0] == joff && b > 0) d += r * b;
        if (slot[1] == joff && a > 0) d += r * a;
    } else if (n == 3) {
        const double a = ldOff(t.cS, slot[0]), b = ldOff(t.cS, slot[1]), c = ldOff(t.cS, slot[2]);
        if (slot[0] == joff && b > 0 && c > 0) d += r * b * c;
        if (slot[1] == joff && a > 0 && c > 0) d += r * a * c;
        if (slot[2] == joff && a > 0 && b > 0) d += r * a * b;
    } else if (n == SIDE_GENERAL) {
        const int32_t* ptr = CTB_TAB(int32_t, t.tp, tv.o_genPtr);
        const uint16_t* sp = CTB_TAB(uint16_t, t.tp, tv.o_genSp);
        const double* order = CTB_TAB(double, t.tp, tv.o_genOrder);
        const int b0 = ptr[slot[0]], e0 = ptr[slot[0] + 1];
        for (int i = b0; i < e0; i++) {
            if (sp[i] != j) continue;
            double prod = r;
            const double oi = order[i];
            const double ci = t.cS[sp[i] * CTB_LDC];
            if (ci > 0. && oi != 0.) {
                prod *= oi * pow(ci, oi - 1);
                for (int l = b0; l < e0; l++) {
                    if (l != i) {
                        const double cl = t.cS[sp[l] * CTB_LDC];
                        if (cl > 0) {
                            prod *= pow(cl, order[l]);
                        } else {
                            prod = 0.;
                            break;
                        }
                    }
                }
            } else {
                prod = 0.;
            }
            d += prod;
        }
    }
    return d;
}

//! col += val * (nu''_e - nu'_e): the reaction's stoichiometric column into a private column buffer.
__device__ __forceinline__ void scatterTile(const TileView& tv, const Tile& t, const Rx2Regs& r, double* colS, double val)
{
    unsigned char* cb = reinterpret_cast<unsigned char*>(colS);
    if (r.hints & RX2_HINT_DISTINCT) {
        // all slots name different rows: six independent read-modify-writes, loads first (the same
        // arithmetic as the chain below, without its serial shared-memory latency)
        double* q[6] = {reinterpret_cast<double*>(cb + r.ir[0]), reinterpret_cast<double*>(cb + r.ir[1]),
                        reinterpret_cast<double*>(cb + r.ir[2]), reinterpret_cast<double*>(cb + r.ip[0]),
                        reinterpret_cast<double*>(cb + r.ip[1]), reinterpret_cast<double*>(cb + r.ip[2])};
        const bool on[6] = {r.nr >= 1, r.nr >= 2, r.nr >= 3, r.np >= 1, r.np >= 2, r.np >= 3};
        double v[6];
#pragma unroll
        for (int i = 0; i < 6; i++) v[i] = on[i] ? *q[i] : 0.0;
#pragma unroll
        for (int i = 0; i < 3; i++) if (on[i]) *q[i] = v[i] - val;
#pragma unroll
        for (int i = 3; i < 6; i++) if (on[i]) *q[i] = v[i] + val;
        return;
    }
    if (r.nr == SIDE_GENERAL) {
        const int32_t* ptr = CTB_TAB(int32_t, t.tp, tv.o_genPtr);
        const uint16_t* sp = CTB_TAB(uint16_t, t.tp, tv.o_genSp);
        const double* st = CTB_TAB(double, t.tp, tv.o_genStoich);
        for (int i = ptr[r.ir[0]]; i < ptr[r.ir[0] + 1]; i++) colS[sp[i] * CTB_LDC] -= st[i] * val;
    } else {
        if (r.nr >= 1) *reinterpret_cast<double*>(cb + r.ir[0]) -= val;
        if (r.nr >= 2) *reinterpret_cast<double*>(cb + r.ir[1]) -= val;
        if (r.nr >= 3) *reinterpret_cast<double*>(cb + r.ir[2]) -= val;
    }
    if (r.np == SIDE_GENERAL) {
        const int32_t* ptr = CTB_TAB(int32_t, t.tp, tv.o_genPtr);
        const uint16_t* sp = CTB_TAB(uint16_t, t.tp, tv.o_genSp);
        const double* st = CTB_TAB(double, t.tp, tv.o_genStoich);
        for (int i = ptr[r.ip[0]]; i < ptr[r.ip[0] + 1]; i++) colS[sp[i] * CTB_LDC] += st[i] * val;
    } else {
        if (r.np >= 1) *reinterpret_cast<double*>(cb + r.ip[0]) += val;
        if (r.np >= 2) *reinterpret_cast<double*>(cb + r.ip[1]) += val;
        if (r.np >= 3) *reinterpret_cast<double*>(cb + r.ip[2]) += val;
    }
}

//! sum of the orders of one side (the degree of homogeneity of its rate of progress)
__device__ __forceinline__ double sideOrderSum(const TileView& tv, const Tile& t, int n, const uint32_t* slot, bool useOrder)
{
    if (n <= 3) return (double) n;
    const int32_t* ptr = CTB_TAB(int32_t, t.tp, tv.o_genPtr);
    const double* ex = CTB_TAB(double, t.tp, useOrder ? tv.o_genOrder : tv.o_genStoich);
    double acc = 0.0;
    for (int i = ptr[slot[0]]; i < ptr[slot[0] + 1]; i++) acc += ex[i];
    return acc;
}

//! One reaction of any kind for this thread's state, with everything the Jacobian needs (see the
//! header). The v rows hold w_k = exp(g'_k) (or g'_k in safe mode), not c_k exp(g'_k).
//! rctot: 1 / c_tot for the reactor's dilution term, 0 for the kinetics getters.
template <bool CHECK>
__device__ __forceinline__ void ropJacTile(const TileView& tv, const Tile& t, const TileScalars& sc, const int e,
                                           const bool safe, const bool skip3b, const bool skipFalloff, const double rctot,
                                           const JacRows& jr)
{
    const unsigned char* tp = t.tp;
    Rx2Regs r;
    loadRx<true, true>(tp + tv.o_rx, e, r);
    double concm = 0.0, tbDefault = 0.0;
    if (r.flags & CTB_RXN_THIRD_BODY) { // ThirdBodyCalc::update (ThirdBodyCalc.h:75-83)
        const int32_t* ptr = CTB_TAB(int32_t, tp, tv.o_tbPtr);
        const uint16_t* sp = CTB_TAB(uint16_t, tp, tv.o_tbSp);
        const double* eff = CTB_TAB(double, tp, tv.o_tbEff);
        double sum = 0.0;
        for (int i = ptr[r.tbIdx]; i < ptr[r.tbIdx + 1]; i++) sum += eff[i] * t.cS[sp[i] * CTB_LDC];
        tbDefault = CTB_TAB(double, tp, tv.o_tbDefault)[r.tbIdx];
        concm = tbDefault * sc.ctot + sum;
    }
    double kf, dk = 0.0;
    if (r.cls == CLS_ARR_CONST) {
        kf = r.A;
    } else if (r.cls == CLS_ARR) {
        kf = r.A * expSafe(r.b * sc.logT - r.EaR * sc.recipT);
    } else {
        GasScalars s;
        s.T = sc.T; s.logT = sc.logT; s.recipT = sc.recipT; s.P = sc.P; s.logP = sc.logP;
        if (r.cls <= CLS_TSANG) {
            const double* fc = CTB_TAB(double, tp, tv.o_fallCoef) + 5 * r.auxIdx;
            FalloffParts parts;
            kf = falloffRate(r.cls, r.flags & CTB_RXN_CHEM_ACTIVATED, CTB_TAB(double, tp, tv.o_fallLow) + 3 * r.auxIdx, fc,
                             r.A, r.b, r.EaR, s, concm, &parts);
            // exact d ln k / d [M] (the reference differences in [M], MultiRate.h:135-165)
            dk = (kf != 0.0 && concm > 0.0) ? falloffDlnkDlnM(r.cls, r.flags & CTB_RXN_CHEM_ACTIVATED, fc, s, parts) / concm : 0.0;
        } else {
            const MechView mv = tileAsMechView(tv);
            kf = r.cls == CLS_PLOG ? plogRate(mv, tp, r.auxIdx, s) : chebyshevRate(mv, tp, r.auxIdx, s);
        }
        kf *= CTB_TAB(double, tp, tv.o_auxMult)[r.auxMultIdx];
    }
    const double kfM = (r.flags & CTB_RXN_MASS_ACTION) ? kf * concm : kf;
    const double ropf = sideMulTile<CHECK>(tv, tp, r.nr, r.ir, t.cS, kfM, true);
    double rk = 0.0, ropr = 0.0;
    if (r.flags & CTB_RXN_REVERSIBLE) {
        if (!safe && r.nr >= 1 && r.nr <= 3 && r.np >= 1 && r.np <= 3) {
            double u[3], w[3];
            u[0] = ldOff(t.uS, r.ir[0]);
            u[1] = r.nr >= 2 ? ldOff(t.uS, r.ir[1]) : 1.0;
            u[2] = r.nr >= 3 ? ldOff(t.uS, r.ir[2]) : 1.0;
            w[0] = ldOff(t.vS, r.ip[0]);
            w[1] = r.np >= 2 ? ldOff(t.vS, r.ip[1]) : 1.0;
            w[2] = r.np >= 3 ? ldOff(t.vS, r.ip[2]) : 1.0;
            rk = fmin(pairedProduct<3, 3, false>(u, w), CTB_BIG);
        } else {
            double dg;
            if (safe) {
                dg = sideSumG(tv, t, r.np, r.ip, true) - sideSumG(tv, t, r.nr, r.ir, true);
            } else {
                // general sides with w = exp(g') in the v rows: g' = log w on the product side
                dg = -sideSumG(tv, t, r.nr, r.ir, false);
                if (r.np == SIDE_GENERAL) {
                    const int32_t* ptr = CTB_TAB(int32_t, tp, tv.o_genPtr);
                    const uint16_t* sp = CTB_TAB(uint16_t, tp, tv.o_genSp);
                    const double* st = CTB_TAB(double, tp, tv.o_genStoich);
                    for (int i = ptr[r.ip[0]]; i < ptr[r.ip[0] + 1]; i++) dg += st[i] * log(t.vS[sp[i] * CTB_LDC]);
                } else {
#pragma unroll
                    for (int i = 0; i < 3; i++) if (i < r.np) dg += log(ldOff(t.vS, r.ip[i])); // constant indices: registers
                }
            }
            rk = fmin(expSafe(dg), CTB_BIG);
        }
        ropr = sideMulTile<CHECK>(tv, tp, r.np, r.ip, t.cS, kfM * rk, false);
    }
    // third-body weight d q / d [M] (ThirdBodyCalc::derivatives weights, BulkKinetics.cpp:680-699)
    double we = 0.0;
    if ((r.flags & CTB_RXN_THIRD_BODY) && !skip3b) {
        const double wf = sideMulTile<CHECK>(tv, tp, r.nr, r.ir, t.cS, kf, true);
        const double wr = (r.flags & CTB_RXN_REVERSIBLE) ? sideMulTile<CHECK>(tv, tp, r.np, r.ip, t.cS, kf * rk, false) : 0.0;
        we = wf - wr;
        if (r.cls >= CLS_LINDEMANN && r.cls <= CLS_TSANG && !skipFalloff) we *= dk;
    }
    double h2 = we * tbDefault;
    if (rctot != 0.0) {
        const double h = sideOrderSum(tv, t, r.nr, r.ir, true) * ropf - sideOrderSum(tv, t, r.np, r.ip, false) * ropr + we * concm;
        h2 -= rctot * h;
    }
    t.qS[e * CTB_TS] = ropf - ropr;
    jr.kfM[e * CTB_TS] = kfM;
    jr.krM[e * CTB_TS] = kfM * rk;
    jr.h2[e * CTB_TS] = h2;
    if (r.flags & CTB_RXN_THIRD_BODY) jr.we[r.tbIdx * CTB_TS] = we;
}

//! The wdot gather of rates_tile.cuh with a second source: epi(k, sum nu q, sum nu h2).
template <int NL, class Epi>
__device__ __forceinline__ void gatherPhase2Tile(const TileView& tv, const Tile& t, const double* h2S, const int j, Epi epi)
{
    constexpr int DEAL = NL == 8 ? 0 : 1;
    const int32_t* lanePtr = CTB_TAB(int32_t, t.tp, tv.o_gLane[DEAL]);
    const uint4* gsp = CTB_TAB(uint4, t.tp, tv.o_gSp[DEAL]);
    const unsigned char* qb = reinterpret_cast<const unsigned char*>(t.qS);
    const unsigned char* hb = reinterpret_cast<const unsigned char*>(h2S);
    for (int i = lanePtr[j]; i < lanePtr[j + 1]; i++) {
        const uint4 g = gsp[i];
        const uint4* ent = reinterpret_cast<const uint4*>(t.tp + tv.o_gEnt + g.x);
        const int nPlus4 = g.y & 0xFFFFu, nMinus4 = g.y >> 16, k = g.z & 0xFFFFu, nGen = g.z >> 16;
        double a = 0.0, hsum = 0.0;
#pragma unroll 1
        for (int p = 0; p < nPlus4; p++) {
            const uint4 o = ent[p];
            a += *reinterpret_cast<const double*>(qb + o.x) + *reinterpret_cast<const double*>(qb + o.y);
            a += *reinterpret_cast<const double*>(qb + o.z) + *reinterpret_cast<const double*>(qb + o.w);
            hsum += *reinterpret_cast<const double*>(hb + o.x) + *reinterpret_cast<const double*>(hb + o.y);
            hsum += *reinterpret_cast<const double*>(hb + o.z) + *reinterpret_cast<const double*>(hb + o.w);
        }
        ent += nPlus4;
#pragma unroll 1
        for (int p = 0; p < nMinus4; p++) {
            const uint4 o = ent[p];
            a -= *reinterpret_cast<const double*>(qb + o.x) + *reinterpret_cast<const double*>(qb + o.y);
            a -= *reinterpret_cast<const double*>(qb + o.z) + *reinterpret_cast<const double*>(qb + o.w);
            hsum -= *reinterpret_cast<const double*>(hb + o.x) + *reinterpret_cast<const double*>(hb + o.y);
            hsum -= *reinterpret_cast<const double*>(hb + o.z) + *reinterpret_cast<const double*>(hb + o.w);
        }
        if (nGen) {
            const GGen2* gg = CTB_TAB(GGen2, t.tp, tv.o_gGen) + g.w;
            for (int p = 0; p < nGen; p++) {
                a += gg[p].nu * *reinterpret_cast<const double*>(qb + gg[p].off);
                hsum += gg[p].nu * *reinterpret_cast<const double*>(hb + gg[p].off);
            }
        }
        epi(k, a, hsum);
    }
}

//! The column pass: reaction lane j builds the columns the deal gave it in its private buffer colS
//! (= &col[0][s], rows of CTB_LDC) and hands each finished column (species jc) to `out`.
template <int NL, class Out>
__device__ __forceinline__ void jacColumnsTile(const TileView& tv, const Tile& t, const JacRows& jr, double* colS, const int j,
                                               const bool skip3b, Out out)
{
    constexpr int DEAL = NL == 4 ? 0 : 1;
    static_assert(NL == 4 || NL == 8, "no column deal for this number of reaction lanes");
    const int32_t* lanePtr = CTB_TAB(int32_t, t.tp, tv.o_colLane[DEAL]);
    const uint16_t* order = CTB_TAB(uint16_t, t.tp, tv.o_colOrder[DEAL]);
    const int32_t* cptr = CTB_TAB(int32_t, t.tp, tv.o_colPtr);
    const GatherEnt* cent = CTB_TAB(GatherEnt, t.tp, tv.o_colEnt);
    const int32_t* tptr = CTB_TAB(int32_t, t.tp, tv.o_tbTPtr);
    const uint16_t* trxn = CTB_TAB(uint16_t, t.tp, tv.o_tbTRxn);
    const double* teff = CTB_TAB(double, t.tp, tv.o_tbTEff);
    const uint16_t* tw = CTB_TAB(uint16_t, t.tp, tv.o_tbTW);
    const unsigned char* rx = t.tp + tv.o_rx;
    for (int ci = lanePtr[j]; ci < lanePtr[j + 1]; ci++) {
        const int jc = order[ci];
        const uint32_t joff = (uint32_t) jc * CTB_LDC * 8u;
        for (int k = 0; k < tv.K; k++) colS[k * CTB_LDC] = 0.0;
        // The coefficient rows live in the L2 scratch: an item's coefficient is requested two items ahead,
        // so that its latency is covered by the work on the items before it (the list holds no item of
        // an irreversible reaction's product side: mech_pack.cpp).
        const int iEnd = cptr[jc + 1];
        int i = cptr[jc];
        GatherEnt g1 = {0, 0}, g2 = {0, 0};
        double k1 = 0.0, k2 = 0.0;
        if (i < iEnd) {
            g1 = cent[i];
            k1 = (g1.nu > 0 ? jr.krM : jr.kfM)[g1.rxn * CTB_TS];
        }
        if (i + 1 < iEnd) {
            g2 = cent[i + 1];
            k2 = (g2.nu > 0 ? jr.krM : jr.kfM)[g2.rxn * CTB_TS];
        }
        for (; i < iEnd; i++) {
            const GatherEnt ge = g1;
            const double kM = k1;
            g1 = g2;
            k1 = k2;
            if (i + 2 < iEnd) {
                g2 = cent[i + 2];
                k2 = (g2.nu > 0 ? jr.krM : jr.kfM)[g2.rxn * CTB_TS];
            }
            Rx2Regs r;
            loadRx<true, true>(rx, ge.rxn, r);
            // product side: -d ropr / d c_j; reactant side: d ropf / d c_j
            const double d = ge.nu > 0 ? -sideDerivTile(tv, t, r.np, r.ip, kM, jc, joff)
                                       : sideDerivTile(tv, t, r.nr, r.ir, kM, jc, joff);
            scatterTile(tv, t, r, colS, d);
        }
        if (!skip3b) { // explicit efficiencies: ThirdBodyCalc::derivatives (ThirdBodyCalc.h:97-100)
            const int tEnd = tptr[jc + 1];
            int it = tptr[jc];
            double w1 = it < tEnd ? jr.we[tw[it] * CTB_TS] : 0.0; // one entry ahead, as above
            for (; it < tEnd; it++) {
                const double we = w1;
                if (it + 1 < tEnd) w1 = jr.we[tw[it + 1] * CTB_TS];
                Rx2Regs r;
                loadRx<true, true>(rx, trxn[it], r);
                scatterTile(tv, t, r, colS, we * teff[it]);
            }
        }
        out(jc);
    }
}

} // namespace ctb
