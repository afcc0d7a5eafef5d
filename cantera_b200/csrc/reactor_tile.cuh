// reactor_tile.cuh — the reactor right-hand side for a TILE of 16 reactors (rates_tile.cuh layout):
// IdealGasConstPressureMoleReactor::updateState + eval (IdealGasConstPressureMoleReactor.cpp:38-106,
// ReactorNet::eval ReactorNet.cpp:617-635) as reactorRhs of bdf.cuh, one half-warp = the 16 reactors
// of the tile on one species / reaction.
#pragma once

#include "rates_tile.cuh"
#include "reactor_types.h"

namespace ctb {

//! Shared-memory areas of one tile of the reactor kernels.
struct ReactorTileBufs {
    double* c;    // [K][17]  moles on entry, concentrations during the pass
    double* u;    // [K][17]
    double* v;    // [K][17]
    double* q;    // [R+1][16], row R = 0; its first 3 NL rows double as scratch before the rop phase
    double* red2; // [NL][16] scratch that is live while the q rows are read
    double* Trow; // [3][16] temperatures, volumes and masses of the tile's reactors (the last two by model)
    volatile int* safeFlag;
    int barId;
};

struct ReactorTileAux {
    double vol, totalCp, RT, ctot, dens, mmw, totalMass, f2;
    int type; // reactor model (reactor.h)
};

//! Mixture sums of the tile's reactors from the species rows of the solution vectors in the c rows (moles
//! for the mole reactors: Phase::setMolesNoTruncate, Phase.cpp:550-568; mass fractions as integrated for
//! the others: setMassFractions_NoNorm): fills aux and sc for temperature T. vol / mass: the V and m rows
//! of the models that have them. Two barriers.
template <int NL, int TT>
__device__ __forceinline__ void reactorStateTile(const TileView& tv, const Tile& t, const ReactorTileBufs& b, const int s,
                                                 const int j, const int type, const double T, const double P,
                                                 const double volRow, const double massRow, TileScalars& sc,
                                                 ReactorTileAux& aux)
{
    const double* molwt = CTB_TAB(double, t.tp, tv.o_molwt);
    const unsigned char* spTab = t.tp + tv.o_sp;
    double* red = b.q;
    const bool mole = reactorIsMole(type);
    double s0 = 0.0, s1 = 0.0; // mole: sum n, sum n W; mass fractions: sum Y / W
    for (int k = j; k < tv.K; k += NL) {
        const double xk = t.cS[k * CTB_LDC];
        if (mole) {
            s0 += xk;
            s1 += xk * molwt[k];
        } else {
            s0 += xk * reinterpret_cast<const double*>(spTab + (size_t) k * sizeof(Sp2))[15];
        }
    }
    red[j * CTB_TS + s] = s0;
    red[(NL + j) * CTB_TS + s] = s1;
    tileSync(b.barId, TT);
    s0 = 0.0; s1 = 0.0;
#pragma unroll
    for (int l = 0; l < NL; l++) { s0 += red[l * CTB_TS + s]; s1 += red[(NL + l) * CTB_TS + s]; }
    tileSync(b.barId, TT); // the scratch is re-used by the caller
    sc.T = T;
    aux.type = type;
    if (mole) {
        aux.totalMass = s1;
        aux.mmw = s1 / s0;
        aux.f2 = 1.0 / (aux.mmw * s0);
    } else {
        aux.totalMass = massRow;
        aux.mmw = 1.0 / s0;
        aux.f2 = 1.0; // unused: the species pass multiplies Y_k by 1 / W_k
    }
    if (reactorIsConstP(type)) {
        aux.dens = P * aux.mmw / (T * CTB_GAS_CONSTANT);
        aux.vol = aux.totalMass / aux.dens;
    } else {
        aux.vol = volRow;
        aux.dens = aux.totalMass / volRow; // setState_TD(T, m / V)
    }
    aux.ctot = aux.dens / aux.mmw;
    aux.RT = T * CTB_GAS_CONSTANT;
    sc.ctot = aux.ctot;
    sc.P = CTB_GAS_CONSTANT * sc.ctot * T;
    sc.logT = log(T);
    sc.recipT = 1.0 / T;
    sc.logP = tv.hasPdep ? log(sc.P) : 0.0;
}

//! One right-hand-side evaluation for the tile. On entry the c rows hold the moles n_k and sc / aux come
//! from reactorStateTile. On exit V wdot_k (the species rows of ydot) is at outS[k * CTB_LDC] and the
//! temperature row is returned (identical in every thread of a reactor). Ends with a barrier, after
//! which the c, u, v and q rows may be re-used.
template <int NL, int TT>
__device__ __forceinline__ double reactorRhsTile(const TileView& tv, const Tile& t, const ReactorTileBufs& b, const int tid,
                                                 const int s, const int j, const TileScalars& sc, ReactorTileAux& aux,
                                                 double* outS)
{
    double* red = b.q;
    double sumq = 0.0;
    const bool mole = reactorIsMole(aux.type), constP = reactorIsConstP(aux.type);
    if (mole ? speciesPhaseTile<NL, true>(tv, t, sc, aux.f2, aux.dens, j, sumq)
             : speciesPhaseTile<NL, true, false, true>(tv, t, sc, aux.f2, aux.dens, j, sumq)) *b.safeFlag = 1;
    red[(2 * NL + j) * CTB_TS + s] = sumq;
    tileSync(b.barId, TT);
    const bool safe = *b.safeFlag != 0;
    if (safe) speciesPhaseSafeTile<NL>(tv, t, sc, j);
    sumq = 0.0;
#pragma unroll
    for (int l = 0; l < NL; l++) sumq += red[(2 * NL + l) * CTB_TS + s];
    tileSync(b.barId, TT); // every lane has read the scratch (it aliases the q rows)
    if (3 * NL > tv.R && tid < CTB_TS) b.q[tv.R * CTB_TS + tid] = 0.0;
    ropPhaseTile<NL, true, 1>(tv, t, sc, j, safe);
    tileSync(b.barId, TT);
    if (tid == 0) *b.safeFlag = 0;
    // m cp (constant pressure) or m cv = m (cp - R) (constant volume: IdealGasMoleReactor.cpp:112, IdealGasReactor.cpp:113)
    const double cp_mole = CTB_GAS_CONSTANT * (aux.mmw * sumq);
    aux.totalCp = aux.totalMass * ((constP ? cp_mole : cp_mole - CTB_GAS_CONSTANT) / aux.mmw);
    double mcpdTdt = 0.0;
    const unsigned char* spTab = t.tp + tv.o_sp;
    const double* molwt = CTB_TAB(double, t.tp, tv.o_molwt);
    const double vol = aux.vol, RT = aux.RT, rmass = 1.0 / aux.totalMass;
    gatherPhaseTile<NL>(tv, t, j, [&](int k, double wdot) {
        double cp, h, sr, rmw;
        nasaTile(spTab, k, sc.T, sc.logT, sc.recipT, cp, h, sr, rmw);
        // partial molar enthalpy h_k, or internal energy u_k = h_k - RT (getPartialMolarIntEnergies_TV)
        mcpdTdt -= wdot * ((constP ? h : h - 1.0) * RT) * vol;
        // dn_k/dt = V wdot_k, or dY_k/dt = V wdot_k W_k / m
        outS[k * CTB_LDC] = mole ? wdot * vol : (wdot * vol * molwt[k]) * rmass;
    });
    b.red2[j * CTB_TS + s] = mcpdTdt;
    tileSync(b.barId, TT);
    mcpdTdt = 0.0;
#pragma unroll
    for (int l = 0; l < NL; l++) mcpdTdt += b.red2[l * CTB_TS + s];
    tileSync(b.barId, TT);
    return mcpdTdt / aux.totalCp;
}

} // namespace ctb
