// rates_thread.cuh — THREAD-PER-STATE evaluation of the ideal-gas kinetics hot path (sm_100a, FP64).
//
// One thread owns one thermochemical state. Its species arrays live in shared memory in a
// [species][thread] layout (consecutive threads -> consecutive banks: every gather is
// conflict-free), and ALL mechanism data (reaction records, Arrhenius parameters, third-body
// lists) is read at warp-uniform addresses (one broadcast wavefront, no divergence: every lane of
// a warp works on the same reaction at the same time). Production rates are accumulated by the
// owning thread into its own column w[k][tid] (the reference's scatter, StoichManager.h:732-744,
// without atomics). Reactions are walked as RUNS of equal "fast signature" (mech_pack.h): the
// common plain-Arrhenius reactions go through an unrolled branch-free body whose U independent
// exp chains give one warp per scheduler enough instruction-level parallelism to keep the FP64
// pipe busy; everything else (third bodies, falloff, Plog, Chebyshev, general sides) takes the
// general per-reaction path with warp-uniform branches.
//
// The expressions follow the reference's operation order exactly as rates.cuh does (same
// citations); only the order of the stoichiometric contraction differs.
#pragma once

#include "rates.cuh"

namespace ctb {

//! Per-thread view of the [species][thread] arrays: pointers are pre-offset by the thread index,
//! element k is at p[k * st].
struct TArr {
    double* __restrict__ c; // [K]   concentrations (holds Y / moles on entry)
    double* __restrict__ g; // [K]   standard chemical potentials
    double* __restrict__ w; // [K]   net production rates (accumulated)
    int st;
};

// The run bodies are separate (noinline) functions: one register allocation each instead of one
// giant kernel. Across that boundary shared-memory locations travel as 32-bit shared-window
// addresses (SPtr) and are turned back into pointers with cvta inside the callee, so that the
// compiler still emits LDS/STS with 32-bit address arithmetic instead of generic loads.
#ifdef __CUDA_ARCH__
typedef uint32_t SPtr;
#define CTB_SP(T, a) (reinterpret_cast<T*>(__cvta_shared_to_generic((size_t) (a))))
__device__ __forceinline__ SPtr toSPtr(const void* p) { return (SPtr) __cvta_generic_to_shared(p); }
#else
typedef const unsigned char* SPtr;
#define CTB_SP(T, a) (reinterpret_cast<T*>(const_cast<unsigned char*>(a)))
inline SPtr toSPtr(const void* p) { return static_cast<const unsigned char*>(p); }
#endif

struct TArrS {
    SPtr c, g, w;
    int st;
};
CTB_HD TArrS packT(const TArr& a)
{
    TArrS r;
    r.c = toSPtr(a.c); r.g = toSPtr(a.g); r.w = toSPtr(a.w); r.st = a.st;
    return r;
}
CTB_HD TArr unpackT(const TArrS& a)
{
    TArr r;
    r.c = CTB_SP(double, a.c); r.g = CTB_SP(double, a.g); r.w = CTB_SP(double, a.w); r.st = a.st;
    return r;
}

//! Everything a fast run needs, by value.
struct FastCtx {
    SPtr tA, tb, tE, tmult, tdn, trec;
    double logT, recipT, rrt, lsc;
    TArrS a;
};
//! Pointer form, rebuilt inside the callee.
struct FastPtrs {
    const double* tA; const double* tb; const double* tE; const double* tmult; const double* tdn;
    const uint4* trec;
    double logT, recipT, rrt, lsc;
    TArr a;
};
CTB_HD FastPtrs unpackCtx(const FastCtx& x)
{
    FastPtrs f;
    f.tA = CTB_SP(const double, x.tA); f.tb = CTB_SP(const double, x.tb); f.tE = CTB_SP(const double, x.tE);
    f.tmult = CTB_SP(const double, x.tmult); f.tdn = CTB_SP(const double, x.tdn);
    f.trec = CTB_SP(const uint4, x.trec);
    f.logT = x.logT; f.recipT = x.recipT; f.rrt = x.rrt; f.lsc = x.lsc;
    f.a = unpackT(x.a);
    return f;
}

CTB_HD void fillScalarsT(const MechView& v, GasScalars& s)
{
    s.recipT = 1.0 / s.T;
    s.RT = s.T * CTB_GAS_CONSTANT;
    s.rrt = 1.0 / s.RT;
    s.logT = log(s.T);
    s.logPref = log(s.P / v.p_ref);
    s.logStandConc = log(s.P / s.RT);
    s.logP = (v.nPlogRx > 0) ? log(s.P) : 0.0;
}

//! stateFromTPY of rates.cuh for one thread: c holds the raw mass fractions on entry and
//! max(Y,0)/sum/W (the mole numbers per unit mass) on exit; dens is returned so that the caller
//! can fuse the final scaling c_k = dens * ym_k into its next sweep.
CTB_HD bool stateFromTPY_T(const MechView& v, const unsigned char* sm, double T, double P,
                                               const TArr& a, GasScalars& s, double& dens)
{
    const int K = v.K, st = a.st;
    const double* rmolwt = CTB_TAB(double, sm, v.o_rmolwt);
    double sum = 0.0;
    for (int k = 0; k < K; k++) {
        double y = a.c[k * st];
        y = y > 0.0 ? y : 0.0;
        a.c[k * st] = y;
        sum += y;
    }
    const double invSum = 1.0 / sum;
    const bool ok = isfinite(invSum);
    double sym = 0.0;
    for (int k = 0; k < K; k++) {
        const double ym = (a.c[k * st] * invSum) * rmolwt[k];
        a.c[k * st] = ym;
        sym += ym;
    }
    const double mmw = 1.0 / sym;
    dens = P * mmw / (T * CTB_GAS_CONSTANT);
    s.T = T;
    s.ctot = dens / mmw;
    s.P = CTB_GAS_CONSTANT * s.ctot * T;
    fillScalarsT(v, s);
    return ok;
}

//! thermoPhase of rates.cuh for one thread, fused with c_k <- dens * c_k and w_k <- 0.
//! Returns sum_k c_k(before scaling) * cp_k/R (the reactor's mixture heat capacity sum).
CTB_HD double thermoT(const MechView& v, const unsigned char* sm, const GasScalars& s,
                                          const TArr& a, double dens)
{
    const int K = v.K, st = a.st;
    const double tmp = s.logPref * s.RT;
    double sumcp = 0.0;
#pragma unroll 2
    for (int k = 0; k < K; k++) {
        double cp, h, sr;
        nasaEval(v, sm, k, s.T, s.logT, s.recipT, cp, h, sr);
        const double gg = (h - sr) * s.RT;
        a.g[k * st] = gg + tmp;
        const double ck = a.c[k * st];
        sumcp += ck * cp;
        a.c[k * st] = dens * ck;
        a.w[k * st] = 0.0;
    }
    return sumcp;
}

// ------------------------------------------------------------------------------------------------
// Fast runs
// ------------------------------------------------------------------------------------------------

template <int PAT> struct PatInfo;
template <> struct PatInfo<PAT_A> { static constexpr int N = 1; static constexpr int NS = 1; };
template <> struct PatInfo<PAT_AB> { static constexpr int N = 2; static constexpr int NS = 2; };
template <> struct PatInfo<PAT_2A> { static constexpr int N = 2; static constexpr int NS = 1; };
template <> struct PatInfo<PAT_ABC> { static constexpr int N = 3; static constexpr int NS = 3; };

//! C1/C2/C3::multiply (StoichManager.h:147-149,205-211,276-283) on already gathered values.
template <int N, bool NEG>
CTB_HD double mulGathered(const double (&x)[3], double r)
{
    if (N == 1) {
        return r * x[0];
    } else if (N == 2) {
        if (NEG && x[0] < 0 && x[1] < 0) return 0.0;
        return r * (x[0] * x[1]);
    } else {
        if (NEG && ((x[0] < 0 && (x[1] < 0 || x[2] < 0)) || (x[1] < 0 && x[2] < 0))) return 0.0;
        return r * (x[0] * x[1] * x[2]);
    }
}

//! One group of U reactions [e, e+U) of signature (REV, PR, PP): q[u] = ropf - ropr and the
//! species slots to scatter into. The scatter of the PREVIOUS group (qp, ip) is issued between the
//! gathers and the exp chains of this one, so that its load-add-store latency hides under them.
template <int PR, int PP, bool REV, bool CONSTK, bool NEG, int U>
CTB_HD void fastGroup(const FastPtrs& x, const int e, double (&qp)[U], int (&ip)[U][6],
                                          const bool havePrev)
{
    constexpr int NR = PatInfo<PR>::N, NP = PatInfo<PP>::N;
    constexpr int SR = PatInfo<PR>::NS, SP = PatInfo<PP>::NS;
    const double* tA = x.tA;
    const double* tb = x.tb;
    const double* tE = x.tE;
    const double* tmult = x.tmult;
    const double* tdn = x.tdn;
    const uint4* trec = x.trec;
    const TArr& a = x.a;
    const int st = a.st;

    int idx[U][6];
    double cr[U][3], cp[U][3], dg[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
        const uint4 rc = trec[e + u];
        idx[u][0] = (int) (rc.x & 0xFFFFu) * st;
        idx[u][1] = (int) (rc.x >> 16) * st;
        idx[u][2] = (int) (rc.y & 0xFFFFu) * st;
        idx[u][3] = (int) (rc.y >> 16) * st;
        idx[u][4] = (int) (rc.z & 0xFFFFu) * st;
        idx[u][5] = (int) (rc.z >> 16) * st;
#pragma unroll
        for (int i = 0; i < NR; i++) cr[u][i] = a.c[idx[u][i]];
        if (REV) {
#pragma unroll
            for (int i = 0; i < NP; i++) cp[u][i] = a.c[idx[u][3 + i]];
            // getRevReactionDelta (Kinetics.cpp:408-418)
            double gp = a.g[idx[u][3]];
#pragma unroll
            for (int i = 1; i < NP; i++) gp += a.g[idx[u][3 + i]];
            double gr = a.g[idx[u][0]];
#pragma unroll
            for (int i = 1; i < NR; i++) gr += a.g[idx[u][i]];
            dg[u] = (0.0 + gp) - gr;
        }
    }
    // ---- scatter of the previous group: distinct species within a reaction (pack-time
    // guarantee), reactions in order
    if (havePrev) {
#pragma unroll
        for (int u = 0; u < U; u++) {
            double wr[3], wp[3];
#pragma unroll
            for (int i = 0; i < SR; i++) wr[i] = a.w[ip[u][i]];
#pragma unroll
            for (int i = 0; i < SP; i++) wp[i] = a.w[ip[u][3 + i]];
            const double q = qp[u];
            const double qr = (PR == PAT_2A) ? 2.0 * q : q;
            const double qq = (PP == PAT_2A) ? 2.0 * q : q;
#pragma unroll
            for (int i = 0; i < SR; i++) a.w[ip[u][i]] = wr[i] - qr;
#pragma unroll
            for (int i = 0; i < SP; i++) a.w[ip[u][3 + i]] = wp[i] + qq;
        }
    }
    // ---- rate constants and rates of progress (BulkKinetics.cpp:466-534)
#pragma unroll
    for (int u = 0; u < U; u++) {
        const double kf = CONSTK ? tA[e + u] : arrhenius(tA[e + u], tb[e + u], tE[e + u], x.logT, x.recipT);
        const double rfn = kf * tmult[e + u];
        double q = mulGathered<NR, NEG>(cr[u], rfn);
        if (REV) {
            const double rkc = fmin(exp(dg[u] * x.rrt - tdn[e + u] * x.lsc), CTB_BIG);
            q -= mulGathered<NP, NEG>(cp[u], rfn * rkc);
        }
        qp[u] = q;
#pragma unroll
        for (int i = 0; i < 6; i++) ip[u][i] = idx[u][i];
    }
}

template <int PR, int PP, bool NEG, int U>
CTB_HD void scatterTail(const TArr& a, const double (&qp)[U], const int (&ip)[U][6], int cnt)
{
    constexpr int SR = PatInfo<PR>::NS, SP = PatInfo<PP>::NS;
#pragma unroll
    for (int u = 0; u < U; u++) {
        if (u < cnt) {
            const double q = qp[u];
            const double qr = (PR == PAT_2A) ? 2.0 * q : q;
            const double qq = (PP == PAT_2A) ? 2.0 * q : q;
            double wr[3], wp[3];
#pragma unroll
            for (int i = 0; i < SR; i++) wr[i] = a.w[ip[u][i]];
#pragma unroll
            for (int i = 0; i < SP; i++) wp[i] = a.w[ip[u][3 + i]];
#pragma unroll
            for (int i = 0; i < SR; i++) a.w[ip[u][i]] = wr[i] - qr;
#pragma unroll
            for (int i = 0; i < SP; i++) a.w[ip[u][3 + i]] = wp[i] + qq;
        }
    }
}

template <int PR, int PP, bool REV, bool CONSTK, bool NEG, int U>
__host__ __device__ __noinline__ void fastRun(const FastCtx xs, int e0, int e1)
{
    const FastPtrs x = unpackCtx(xs);
    const TArr& a = x.a;
    double qp[U];
    int ip[U][6];
    int e = e0;
    bool have = false;
    for (; e + U <= e1; e += U) {
        fastGroup<PR, PP, REV, CONSTK, NEG, U>(x, e, qp, ip, have);
        have = true;
    }
    if (have) scatterTail<PR, PP, NEG, U>(a, qp, ip, U);
    // remainder, one at a time
    for (; e < e1; e++) {
        double q1[1];
        int i1[1][6];
        fastGroup<PR, PP, REV, CONSTK, NEG, 1>(x, e, q1, i1, false);
        scatterTail<PR, PP, NEG, 1>(a, q1, i1, 1);
    }
}

// ------------------------------------------------------------------------------------------------
// General path: any reaction, warp-uniform branches
// ------------------------------------------------------------------------------------------------

//! sideMultiply of rates.cuh on a strided array
template <bool NEG>
CTB_HD double sideMultiplyS(const MechView& v, const unsigned char* sm, int n,
                                                const uint16_t* slot, const double* S, int st, double r)
{
    if (n == 1) {
        return r * S[slot[0] * st];
    } else if (n == 2) {
        const double a = S[slot[0] * st], b = S[slot[1] * st];
        if (NEG && a < 0 && b < 0) return 0.0;
        return r * (a * b);
    } else if (n == 3) {
        const double a = S[slot[0] * st], b = S[slot[1] * st], c = S[slot[2] * st];
        if (NEG && ((a < 0 && (b < 0 || c < 0)) || (b < 0 && c < 0))) return 0.0;
        return r * (a * b * c);
    } else if (n == SIDE_GENERAL) {
        const int32_t* ptr = CTB_TAB(int32_t, sm, v.o_genPtr);
        const uint16_t* sp = CTB_TAB(uint16_t, sm, v.o_genSp);
        const double* order = CTB_TAB(double, sm, v.o_genOrder);
        for (int j = ptr[slot[0]]; j < ptr[slot[0] + 1]; j++) {
            const double o = order[j];
            if (o != 0.0) {
                const double c = S[sp[j] * st];
                if (c > 0.0) {
                    r *= pow(c, o);
                } else {
                    r = 0.0;
                }
            }
        }
        return r;
    }
    return r;
}

CTB_HD double sideSumS(const MechView& v, const unsigned char* sm, int n,
                                           const uint16_t* slot, const double* S, int st)
{
    if (n == 1) {
        return S[slot[0] * st];
    } else if (n == 2) {
        return S[slot[0] * st] + S[slot[1] * st];
    } else if (n == 3) {
        return S[slot[0] * st] + S[slot[1] * st] + S[slot[2] * st];
    } else if (n == SIDE_GENERAL) {
        const int32_t* ptr = CTB_TAB(int32_t, sm, v.o_genPtr);
        const uint16_t* sp = CTB_TAB(uint16_t, sm, v.o_genSp);
        const double* stc = CTB_TAB(double, sm, v.o_genStoich);
        double acc = 0.0;
        for (int j = ptr[slot[0]]; j < ptr[slot[0] + 1]; j++) acc += stc[j] * S[sp[j] * st];
        return acc;
    }
    return 0.0;
}

//! w[species of the side] += sign * q, one molecule slot after the other (duplicates allowed)
CTB_HD void sideScatterS(const MechView& v, const unsigned char* sm, int n,
                                             const uint16_t* slot, double* W, int st, double q)
{
    if (n == SIDE_GENERAL) {
        const int32_t* ptr = CTB_TAB(int32_t, sm, v.o_genPtr);
        const uint16_t* sp = CTB_TAB(uint16_t, sm, v.o_genSp);
        const double* stc = CTB_TAB(double, sm, v.o_genStoich);
        for (int j = ptr[slot[0]]; j < ptr[slot[0] + 1]; j++) W[sp[j] * st] += stc[j] * q;
    } else {
        if (n >= 1) W[slot[0] * st] += q;
        if (n >= 2) W[slot[1] * st] += q;
        if (n >= 3) W[slot[2] * st] += q;
    }
}

template <bool NEG>
__host__ __device__ __noinline__ void generalRun(const SPtr smS, const uint32_t oView, int e0, int e1,
                                                 const GasScalars s, const TArrS as)
{
    const unsigned char* sm = CTB_SP(const unsigned char, smS);
    const MechView& v = *reinterpret_cast<const MechView*>(sm + oView); // the blob carries its own view
    const TArr a = unpackT(as);
    const double* tA = CTB_TAB(double, sm, v.o_A);
    const double* tb = CTB_TAB(double, sm, v.o_b);
    const double* tE = CTB_TAB(double, sm, v.o_EaR);
    const double* tmult = CTB_TAB(double, sm, v.o_mult);
    const double* tdn = CTB_TAB(double, sm, v.o_dn);
    const uint4* trec = CTB_TAB(uint4, sm, v.o_rec);
    const int16_t* tbIdx = CTB_TAB(int16_t, sm, v.o_tbIdx);
    const int16_t* auxIdx = CTB_TAB(int16_t, sm, v.o_auxIdx);
    const int st = a.st;
    for (int e = e0; e < e1; e++) {
        union { uint4 u; RxnRec r; } rr;
        rr.u = trec[e];
        const RxnRec& r = rr.r;
        const int flags = r.flags;
        double concm = 0.0;
        if (flags & CTB_RXN_THIRD_BODY) {
            const int t = tbIdx[e];
            const int32_t* ptr = CTB_TAB(int32_t, sm, v.o_tbPtr);
            const uint16_t* sp = CTB_TAB(uint16_t, sm, v.o_tbSp);
            const double* eff = CTB_TAB(double, sm, v.o_tbEff);
            double sum = 0.0;
            const int j1 = ptr[t + 1];
#pragma unroll 4
            for (int j = ptr[t]; j < j1; j++) sum += eff[j] * a.c[sp[j] * st];
            concm = CTB_TAB(double, sm, v.o_tbDefault)[t] * s.ctot + sum;
        }
        double kf0;
        const int cls = r.cls;
        if (cls == CLS_ARR_CONST) {
            kf0 = tA[e];
        } else if (cls == CLS_ARR) {
            kf0 = arrhenius(tA[e], tb[e], tE[e], s.logT, s.recipT);
        } else if (cls <= CLS_TSANG) {
            const int ax = auxIdx[e];
            kf0 = falloffRate(cls, flags & CTB_RXN_CHEM_ACTIVATED, CTB_TAB(double, sm, v.o_fallLow) + 3 * ax,
                              CTB_TAB(double, sm, v.o_fallCoef) + 5 * ax, tA[e], tb[e], tE[e], s, concm, nullptr);
        } else if (cls == CLS_PLOG) {
            kf0 = plogRate(v, sm, auxIdx[e], s);
        } else {
            kf0 = chebyshevRate(v, sm, auxIdx[e], s);
        }
        double rkc = 0.0;
        if (flags & CTB_RXN_REVERSIBLE) {
            double dg = 0.0 + sideSumS(v, sm, r.np, r.p, a.g, st);
            dg -= sideSumS(v, sm, r.nr, r.r, a.g, st);
            rkc = fmin(exp(dg * s.rrt - tdn[e] * s.logStandConc), CTB_BIG);
        }
        const double rfn = kf0 * tmult[e];
        double ropf = rfn;
        if (flags & CTB_RXN_MASS_ACTION) ropf *= concm;
        double ropr = ropf * rkc;
        ropf = sideMultiplyS<NEG>(v, sm, r.nr, r.r, a.c, st, ropf);
        if (flags & CTB_RXN_REVERSIBLE) ropr = sideMultiplyS<NEG>(v, sm, r.np, r.p, a.c, st, ropr);
        const double q = ropf - ropr;
        sideScatterS(v, sm, r.nr, r.r, a.w, st, -q);
        sideScatterS(v, sm, r.np, r.p, a.w, st, q);
    }
}

// ------------------------------------------------------------------------------------------------
// All reactions of one state
// ------------------------------------------------------------------------------------------------

#define CTB_FAST_CASE(ck, rev, pr, pp) \
    case ((((ck) * 2 + (rev)) * PAT_COUNT + (pr)) * PAT_COUNT + (pp)): \
        fastRun<pr, pp, (rev) != 0, (ck) != 0, NEG, U>(x, b, e); break;
#define CTB_FAST_CASES_P(ck, rev, pr) \
    CTB_FAST_CASE(ck, rev, pr, PAT_A) CTB_FAST_CASE(ck, rev, pr, PAT_AB) CTB_FAST_CASE(ck, rev, pr, PAT_2A) \
    CTB_FAST_CASE(ck, rev, pr, PAT_ABC)
#define CTB_FAST_CASES(ck, rev) \
    CTB_FAST_CASES_P(ck, rev, PAT_A) CTB_FAST_CASES_P(ck, rev, PAT_AB) CTB_FAST_CASES_P(ck, rev, PAT_2A)

//! BulkKinetics::updateROP + Kinetics::getNetProductionRates for the state of this thread:
//! a.c, a.g filled, a.w zeroed on entry; a.w = wdot on exit. NEG: keep the negative-concentration
//! guards of C2/C3::multiply (needed only when concentrations can be negative, i.e. for reactor
//! states set without truncation).
template <bool NEG, int U>
CTB_HD void reactionsT(const MechView& v, const unsigned char* sm, const GasScalars& s,
                                           const TArr& a)
{
    FastCtx x;
    x.tA = toSPtr(sm + v.o_A);
    x.tb = toSPtr(sm + v.o_b);
    x.tE = toSPtr(sm + v.o_EaR);
    x.tmult = toSPtr(sm + v.o_mult);
    x.tdn = toSPtr(sm + v.o_dn);
    x.trec = toSPtr(sm + v.o_rec);
    x.logT = s.logT; x.recipT = s.recipT; x.rrt = s.rrt; x.lsc = s.logStandConc;
    x.a = packT(a);
    const SPtr smS = toSPtr(sm);
    const RunRec* runs = CTB_TAB(RunRec, sm, v.o_runs);
    for (int i = 0; i < v.nRuns; i++) {
        const int4 rr = *reinterpret_cast<const int4*>(runs + i);
        const int b = rr.x, e = rr.y;
        switch (rr.z) {
            CTB_FAST_CASES(0, 0)
            CTB_FAST_CASES(0, 1)
            CTB_FAST_CASES(1, 0)
            CTB_FAST_CASES(1, 1)
        default:
            generalRun<NEG>(smS, v.o_view, b, e, s, x.a);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Transposed staging between row-per-state global arrays and the [k][thread] shared layout
// ------------------------------------------------------------------------------------------------

//! dst[k * st + row] = src[row * ld + k] for row < rows, k < width (coalesced global reads; the
//! odd stride st makes the shared-memory writes conflict-free).
__device__ __forceinline__ void loadTransposed(double* dst, int st, const double* __restrict__ src, long long ld,
                                               int rows, int width)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int row = warp; row < rows; row += nw) {
        const double* p = src + (long long) row * ld;
        for (int k = lane; k < width; k += 32) dst[k * st + row] = p[k];
    }
}

__device__ __forceinline__ void storeTransposed(double* __restrict__ dst, long long ld, const double* src, int st,
                                                int rows, int width)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
    for (int row = warp; row < rows; row += nw) {
        double* p = dst + (long long) row * ld;
        for (int k = lane; k < width; k += 32) p[k] = src[k * st + row];
    }
}

} // namespace ctb
