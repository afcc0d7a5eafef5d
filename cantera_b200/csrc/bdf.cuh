// bdf.cuh — per-reactor device code of BatchReactorNet:
//   * the IdealGasConstPressureMoleReactor right-hand side and its Newton matrix,
//   * a dense LU with partial pivoting in shared memory,
//   * a variable-order (1-5), variable-step BDF integrator in Nordsieck form with a fixed leading
//     coefficient — the published VODE/CVODE algorithm (Brown, Byrne, Hindmarsh 1989) with the
//     controller constants the reference's CVODES uses (ext/sundials/src/cvodes/cvodes_impl.h,
//     SURVEY.md Appendix C), so that step counts are comparable with ReactorNet::advance.
// One cooperative GROUP of lanes (a warp, or NT threads on a named barrier) owns one reactor; all
// scalar controller state is replicated in the registers of every lane of the group (reductions
// return bit-identical values to every lane), so control flow is uniform inside a group and free
// between groups.
#pragma once

#include "rates.cuh"
#include "reactor.h"

namespace ctb {

// ------------------------------------------------------------------------------------------------
// A group of NT threads (multiple of 32) inside a larger CTA, synchronised by a named barrier.
// ------------------------------------------------------------------------------------------------
template <int NT_>
struct NamedGroup {
    static constexpr int NT = NT_;
    double* scratch; // >= NT/32 + 1 doubles of shared memory owned by this group
    int id;          // barrier id 1..15
    __device__ __forceinline__ NamedGroup(double* s, int barrierId) : scratch(s), id(barrierId) {}
    __device__ __forceinline__ int lane() const { return threadIdx.x % NT; }
    __device__ __forceinline__ void sync() const {
        if (NT == 32) {
            __syncwarp();
        } else {
            asm volatile("bar.sync %0, %1;" ::"r"(id), "n"(NT) : "memory");
        }
    }
    __device__ __forceinline__ double sum(double x) const {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        if (NT == 32) return x;
        sync();
        if ((threadIdx.x & 31) == 0) scratch[lane() >> 5] = x;
        sync();
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < NT / 32; w++) s += scratch[w];
        return s;
    }
    __device__ __forceinline__ double max(double x) const {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x = fmax(x, __shfl_xor_sync(0xffffffffu, x, o));
        if (NT == 32) return x;
        sync();
        if ((threadIdx.x & 31) == 0) scratch[lane() >> 5] = x;
        sync();
        double s = scratch[0];
#pragma unroll
        for (int w = 1; w < NT / 32; w++) s = fmax(s, scratch[w]);
        return s;
    }
};

// ------------------------------------------------------------------------------------------------
// Shared-memory work area of one reactor
// ------------------------------------------------------------------------------------------------
struct ReactorWork {
    int n;        // K + 1 equations
    int ldm;      // odd leading dimension of M (conflict-free row AND column sweeps)
    double* M;    // [n * ldm] Newton matrix I - gamma*J, then its LU factors (column-major)
    double* zn;   // [6 * n] Nordsieck array
    double* acor; // [n] accumulated corrector
    double* ewt;  // [n] error weights
    double* y;    // [n] current iterate
    double* f;    // [n] right-hand side
    double* tmp;  // [n]
    double* b;    // [n] linear-system right-hand side / solution
    double* c;    // [K] concentrations
    double* grt;  // [K] standard chemical potentials
    double* hrt;  // [K]
    double* cpr;  // [K]
    double* q;    // [R] net rates of progress
    double* rfn;  // [R]   } only filled when the Jacobian is being built
    double* rkc;  // [R]   }
    double* concm;// [R]   }
    double* dk;   // [R]   }
    double* w;    // [R]   }
    double* D;    // [K + R] }
    int* perm;    // [n] row permutation of the LU
};

__host__ __device__ inline size_t reactorWorkDoubles(int K, int R)
{
    const int n = K + 1, ldm = n | 1;
    size_t d = (size_t) n * ldm + 6 * n + 6 * n + 4 * K + 6 * R + (K + R);
    d += (n + 1) / 2; // perm (int32)
    return (d + 1) & ~(size_t) 1;
}

__device__ inline ReactorWork carveWork(double* base, int K, int R)
{
    ReactorWork w;
    w.n = K + 1;
    w.ldm = w.n | 1;
    double* p = base;
    w.M = p; p += (size_t) w.n * w.ldm;
    w.zn = p; p += 6 * w.n;
    w.acor = p; p += w.n;
    w.ewt = p; p += w.n;
    w.y = p; p += w.n;
    w.f = p; p += w.n;
    w.tmp = p; p += w.n;
    w.b = p; p += w.n;
    w.c = p; p += K;
    w.grt = p; p += K;
    w.hrt = p; p += K;
    w.cpr = p; p += K;
    w.q = p; p += R;
    w.rfn = p; p += R;
    w.rkc = p; p += R;
    w.concm = p; p += R;
    w.dk = p; p += R;
    w.w = p; p += R;
    w.D = p; p += K + R;
    w.perm = reinterpret_cast<int*>(p);
    return w;
}

// ------------------------------------------------------------------------------------------------
// Reactor right-hand side
// ------------------------------------------------------------------------------------------------

struct RhsAux {
    double ctot, vol, totalCp, RT, Tdot;
};

//! IdealGasConstPressureMoleReactor::updateState + eval, ReactorNet::eval
//! (IdealGasConstPressureMoleReactor.cpp:38-106, ReactorNet.cpp:617-635) for y = [T, n_1..n_K]:
//! ydot[0] = -V sum_k wdot_k h_k / (m cp), ydot[1+k] = V wdot_k. Phase state from moles without
//! clipping (Phase::setMolesNoTruncate, Phase.cpp:550-568). WITH_JAC also leaves the per-reaction
//! intermediates the Jacobian needs in the work area.
template <class G, bool WITH_JAC>
__device__ __forceinline__ void reactorRhs(const G& g, const MechView& v, const unsigned char* sm,
                                           const unsigned char* gm, const ReactorWork& wk, double P,
                                           const double* y, double* ydot, RhsAux& aux)
{
    const int K = v.K;
    const double* molwt = CTB_TAB(double, sm, v.o_molwt);
    const double T = y[0];
    double totalMoles = 0.0, totalMass = 0.0;
    for (int k = g.lane(); k < K; k += G::NT) {
        const double nk = y[1 + k];
        totalMoles += nk;
        totalMass += nk * molwt[k];
    }
    totalMoles = g.sum(totalMoles);
    totalMass = g.sum(totalMass);
    const double mmw = totalMass / totalMoles;
    const double f2 = 1.0 / (mmw * totalMoles);
    const double dens = P * mmw / (T * CTB_GAS_CONSTANT);
    GasScalars s;
    s.T = T;
    s.ctot = dens / mmw;
    s.P = CTB_GAS_CONSTANT * s.ctot * T;
    fillScalars(g, v, s);
    thermoPhase<G, true>(g, v, sm, s, wk.grt, wk.hrt, wk.cpr);
    double sumq = 0.0;
    for (int k = g.lane(); k < K; k += G::NT) {
        const double ym = y[1 + k] * f2;
        sumq += ym * wk.cpr[k]; // same lane wrote cpr[k]
        wk.c[k] = dens * ym;
    }
    sumq = g.sum(sumq);
    g.sync();
    RopOut o = {};
    o.q = wk.q;
    if (WITH_JAC) {
        o.rfn = wk.rfn; o.rkc = wk.rkc; o.concm = wk.concm; o.dlnkdM = wk.dk;
    }
    ropPhase(g, v, sm, s, wk.c, wk.grt, o);
    g.sync();
    const double vol = totalMass / dens;
    const double cp_mole = CTB_GAS_CONSTANT * (mmw * sumq);
    const double totalCp = totalMass * (cp_mole / mmw);
    double mcpdTdt = 0.0;
    if (G::NT == 8) {
        // balanced deal of the species to the 8 lanes (mech_pack.h o_gl8*)
        const int32_t* lp = CTB_TAB(int32_t, sm, v.o_gl8Ptr);
        const uint16_t* ls = CTB_TAB(uint16_t, sm, v.o_gl8Sp);
        for (int i = lp[g.lane()]; i < lp[g.lane() + 1]; i++) {
            const int k = ls[i];
            const double wdot = gatherSpecies<0>(v, sm, gm, k, wk.q, nullptr);
            const double hk = wk.hrt[k] * s.RT;
            mcpdTdt -= wdot * hk * vol;
            ydot[1 + k] = wdot * vol;
        }
    } else {
        for (int k = g.lane(); k < K; k += G::NT) {
            const double wdot = gatherSpecies<0>(v, sm, gm, k, wk.q, nullptr);
            const double hk = wk.hrt[k] * s.RT;
            mcpdTdt -= wdot * hk * vol;
            ydot[1 + k] = wdot * vol;
        }
    }
    mcpdTdt = g.sum(mcpdTdt);
    const double Tdot = mcpdTdt / totalCp;
    if (g.lane() == 0) ydot[0] = Tdot;
    aux.ctot = s.ctot; aux.vol = vol; aux.totalCp = totalCp; aux.RT = s.RT; aux.Tdot = Tdot;
    g.sync();
}

// ------------------------------------------------------------------------------------------------
// Newton matrix  M = I - gamma * d(ydot)/dy
// ------------------------------------------------------------------------------------------------

//! Builds J = d(ydot)/dy of the const-pressure mole reactor at wk.y into wk.M, and f(y) into wk.f.
//!   species block: d(V wdot_k)/d n_j = ddCi_kj - sum_l ddCi_kl X_l + wdot_k/c_tot   (exact: the
//!       reference's preconditioner template IdealGasConstPressureMoleReactor.cpp:108-123 keeps only
//!       ddCi; the two other terms are the volume-dilution part V = n_tot R T / P)
//!   T row: -(sum_k h_k J_kj + Tdot cp_j) / (m cp)                                   (:153-173)
//!   T column: one-sided finite difference with dT = T sqrt(eps)                    (:126-152)
//! ddCi is the analytic composition Jacobian of rates.cuh (falloff [M]-derivative exact).
//! Then M <- I - gamma*J. Costs two RHS evaluations.
template <class G>
__device__ __forceinline__ void buildNewtonMatrix(const G& g, const MechView& v, const unsigned char* sm,
                                                  const unsigned char* gm, const ReactorWork& wk, double P,
                                                  double gamma, long long* prof = nullptr,
                                                  const bool formNewton = true)
{
    const int K = v.K, n = wk.n, ldm = wk.ldm;
    RhsAux aux;
    long long tk = prof ? clock64() : 0;
    auto tick = [&](int slot) {
        if (prof) {
            const long long t = clock64();
            if (g.lane() == 0) atomicAdd(reinterpret_cast<unsigned long long*>(prof + slot), (unsigned long long) (t - tk));
            tk = t;
        }
    };
    reactorRhs<G, true>(g, v, sm, gm, wk, P, wk.y, wk.f, aux);
    tick(0);
    double* A = wk.M + ldm + 1; // species block, A[j*ldm + k]
    jacobianCi<G, true>(g, v, sm, gm, wk.c, wk.rfn, wk.rkc, wk.concm, wk.dk, wk.w, wk.D, A, ldm, 1.0, false, false);
    tick(1);
    // add_k = wdot_k / ctot - sum_l A_kl X_l
    const double rctot = 1.0 / aux.ctot;
    const double rN = 1.0 / (aux.vol * aux.ctot);
    for (int k = g.lane(); k < K; k += G::NT) {
        double acc = 0.0;
        for (int l = 0; l < K; l++) acc += A[(size_t) l * ldm + k] * wk.c[l];
        wk.tmp[1 + k] = wk.f[1 + k] * rN - acc * rctot;
    }
    g.sync();
    const double rC = 1.0 / aux.totalCp;
    for (int j = g.lane(); j < K; j += G::NT) {
        double* col = A + (size_t) j * ldm;
        double acc = 0.0;
        for (int k = 0; k < K; k++) {
            const double val = col[k] + wk.tmp[1 + k];
            col[k] = val;
            acc += wk.hrt[k] * val;
        }
        // h_k = hrt*RT, cp_j = R*cpr_j
        wk.M[(size_t) (1 + j) * ldm] = -(acc * aux.RT + aux.Tdot * (CTB_GAS_CONSTANT * wk.cpr[j])) * rC;
    }
    g.sync();
    tick(2);
    // temperature column
    const double T = wk.y[0];
    const double dT = T * 1.4901161193847656e-08; // sqrt(DBL_EPSILON)
    for (int i = g.lane(); i < n; i += G::NT) wk.b[i] = wk.y[i];
    g.sync();
    if (g.lane() == 0) wk.b[0] = T + dT;
    g.sync();
    RhsAux aux2;
    reactorRhs<G, false>(g, v, sm, gm, wk, P, wk.b, wk.tmp, aux2);
    tick(3);
    const double rdT = 1.0 / dT;
    for (int i = g.lane(); i < n; i += G::NT) wk.M[i] = (wk.tmp[i] - wk.f[i]) * rdT;
    g.sync();
    if (formNewton) {
        for (int j = 0; j < n; j++) {
            double* col = wk.M + (size_t) j * ldm;
            for (int i = g.lane(); i < n; i += G::NT) col[i] = (i == j ? 1.0 : 0.0) - gamma * col[i];
        }
        g.sync();
    }
    tick(4);
}

// ------------------------------------------------------------------------------------------------
// Dense LU with partial pivoting, in place in shared memory (the role of SUNLinSol_Dense,
// CVodesIntegrator.cpp:382-411). Lanes own rows in the column sweeps (consecutive addresses) and
// columns in the row swap (stride ldm, odd => conflict-free).
// ------------------------------------------------------------------------------------------------

//! Returns 0, or k+1 if a zero pivot was met in column k.
template <class G>
__device__ __forceinline__ int luFactor(const G& g, const ReactorWork& wk)
{
    const int n = wk.n, ldm = wk.ldm;
    double* M = wk.M;
    const int rowLane = g.lane() & 31;
    const int colGrp = g.lane() >> 5;
    constexpr int NCG = G::NT / 32;
    int fail = 0;
    for (int k = 0; k < n; k++) {
        double* ck = M + (size_t) k * ldm;
        // pivot search
        double best = -1.0;
        int bi = n;
        for (int i = k + g.lane(); i < n; i += G::NT) {
            const double a = fabs(ck[i]);
            if (a > best) { best = a; bi = i; }
        }
        const double mx = g.max(best);
        const int p = -(int) g.max(best == mx ? (double) (-bi) : (double) (-n));
        if (g.lane() == 0) wk.perm[k] = p;
        if (!(mx > 0.0)) {
            if (!fail) fail = k + 1;
            continue;
        }
        if (p != k) {
            for (int j = g.lane(); j < n; j += G::NT) {
                double* cj = M + (size_t) j * ldm;
                const double t = cj[k];
                cj[k] = cj[p];
                cj[p] = t;
            }
        }
        g.sync();
        const double rp = 1.0 / ck[k];
        for (int i0 = k + 1; i0 < n; i0 += 32) {
            const int i = i0 + rowLane;
            if (i < n) {
                const double li = ck[i] * rp;
                if (NCG == 1) ck[i] = li;
                for (int j = k + 1 + colGrp; j < n; j += NCG) {
                    double* cj = M + (size_t) j * ldm;
                    cj[i] -= li * cj[k];
                }
            }
        }
        if (NCG > 1) {
            // several warps read column k above: store the multipliers only once all have
            g.sync();
            for (int i = k + 1 + g.lane(); i < n; i += G::NT) ck[i] *= rp;
        }
        g.sync();
    }
    return fail;
}

//! Solves (LU) x = b in place in wk.b.
template <class G>
__device__ __forceinline__ void luSolve(const G& g, const ReactorWork& wk)
{
    const int n = wk.n, ldm = wk.ldm;
    const double* M = wk.M;
    double* b = wk.b;
    // permutation (sequential swaps as recorded) then unit-lower forward substitution
    if (g.lane() == 0) {
        for (int k = 0; k < n; k++) {
            const int p = wk.perm[k];
            if (p != k) { const double t = b[k]; b[k] = b[p]; b[p] = t; }
        }
    }
    g.sync();
    for (int k = 0; k < n - 1; k++) {
        const double bk = b[k];
        const double* ck = M + (size_t) k * ldm;
        for (int i = k + 1 + g.lane(); i < n; i += G::NT) b[i] -= ck[i] * bk;
        g.sync();
    }
    for (int k = n - 1; k >= 0; k--) {
        const double* ck = M + (size_t) k * ldm;
        const double xk = b[k] / ck[k];
        g.sync();
        if (g.lane() == 0) b[k] = xk;
        for (int i = g.lane(); i < k; i += G::NT) b[i] -= ck[i] * xk;
        g.sync();
    }
}

// ------------------------------------------------------------------------------------------------
// Variable-order variable-step BDF (fixed-leading-coefficient Nordsieck form)
// ------------------------------------------------------------------------------------------------

// Controller constants: ext/sundials/src/cvodes/cvodes_impl.h:53-126, cvodes.c:305-310
// (SURVEY.md Appendix C).
#define BDF_QMAX 5
#define BDF_MSBP 20        // max steps between Newton-matrix setups
#define BDF_MSBJ 50        // max steps between Jacobian evaluations (cvodes_ls_impl.h CVLS_MSBJ)
#define BDF_DGMAX_JBAD 0.2 // gamma change below which a failure is blamed on the Jacobian (CVLS_DGMAX)
#define BDF_DGMAX 0.3      // |gamma/gammap - 1| that forces a setup
#define BDF_MAXCOR 3       // Newton iterations per attempt
#define BDF_CRDOWN 0.3
#define BDF_RDIV 2.0
#define BDF_CORTES 0.1     // Newton convergence coefficient
#define BDF_MXNCF 10
#define BDF_MXNEF 7
#define BDF_MXNEF1 3
#define BDF_SMALL_NEF 2
#define BDF_LONG_WAIT 10
#define BDF_ETAMX1 10000.0 // first step
#define BDF_ETAMX2 10.0    // later steps
#define BDF_ETAMXF 0.2
#define BDF_ETAMIN 0.1
#define BDF_ETACF 0.25
#define BDF_THRESH 1.5
#define BDF_ADDON 1e-6
#define BDF_BIAS1 6.0
#define BDF_BIAS2 6.0
#define BDF_BIAS3 10.0
#define BDF_HIST_SCALARS 32

enum { NFLAG_FIRST = 0, NFLAG_PREV_CONV_FAIL = 1, NFLAG_PREV_ERR_FAIL = 2 };

template <class G>
struct Bdf {
    const G& g;
    const MechView& v;
    const unsigned char* sm;
    const unsigned char* gm;
    ReactorWork wk;
    double P, rtol, atolT, atolS, crateFloor;
    // integrator memory (identical in every lane of the group)
    int q, L, qwait, qprime, nst, nstlp;
    bool forceSetup; // the Newton matrix is not kept between advance() calls
    double tn, h, hscale, hprime, eta, etamax, gamma, gammap, gamrat, crate, acnrm, savedTq5, rl1;
    double tau[BDF_QMAX + 1], tq[BDF_QMAX + 1], l[BDF_QMAX + 1];
    // counters (ReactorNet::solverStats)
    long long nfe, nsetups, nni, ncfn, netf;

    __device__ __forceinline__ Bdf(const G& g_, const MechView& v_, const unsigned char* sm_,
                                   const unsigned char* gm_, const ReactorWork& wk_)
        : g(g_), v(v_), sm(sm_), gm(gm_), wk(wk_) {}

    __device__ __forceinline__ double* zn(int j) const { return wk.zn + (size_t) j * wk.n; }

    __device__ __forceinline__ double wrms(const double* x) const {
        double s = 0.0;
        for (int i = g.lane(); i < wk.n; i += G::NT) {
            const double t = x[i] * wk.ewt[i];
            s += t * t;
        }
        return sqrt(g.sum(s) / wk.n);
    }

    //! ewt_i = 1 / (rtol |y_i| + atol_i)  (cvEwtSetSV)
    __device__ __forceinline__ void setEwt(const double* y) const {
        for (int i = g.lane(); i < wk.n; i += G::NT) {
            wk.ewt[i] = 1.0 / (rtol * fabs(y[i]) + (i == 0 ? atolT : atolS));
        }
        g.sync();
    }

    __device__ __forceinline__ void rhs(const double* y, double* ydot) {
        RhsAux aux;
        reactorRhs<G, false>(g, v, sm, gm, wk, P, y, ydot, aux);
        nfe++;
    }

    // ---- Nordsieck array manipulation (each lane owns its elements in every vector: no syncs)
    __device__ __forceinline__ void predict() {
        tn += h;
        for (int i = g.lane(); i < wk.n; i += G::NT) {
            for (int k = 1; k <= q; k++)
                for (int j = q; j >= k; j--) zn(j - 1)[i] += zn(j)[i];
        }
    }
    __device__ __forceinline__ void restore(double savedT) {
        tn = savedT;
        for (int i = g.lane(); i < wk.n; i += G::NT) {
            for (int k = 1; k <= q; k++)
                for (int j = q; j >= k; j--) zn(j - 1)[i] -= zn(j)[i];
        }
    }
    __device__ __forceinline__ void rescale() {
        double factor = eta;
        for (int j = 1; j <= q; j++) {
            for (int i = g.lane(); i < wk.n; i += G::NT) zn(j)[i] *= factor;
            factor *= eta;
        }
        h = hscale * eta;
        hscale = h;
    }

    //! Coefficients l[] of the corrector polynomial and the test quantities tq[] for the current
    //! order and step history (VODE eqs. for the fixed-leading-coefficient BDF).
    __device__ __forceinline__ void setCoefficients() {
        for (int i = 0; i <= BDF_QMAX; i++) l[i] = 0.0;
        l[0] = l[1] = 1.0;
        double xi_inv = 1.0, xistar_inv = 1.0, alpha0 = -1.0, alpha0_hat = -1.0, hsum = h;
        if (q > 1) {
            for (int j = 2; j < q; j++) {
                hsum += tau[j - 1];
                xi_inv = h / hsum;
                alpha0 -= 1.0 / j;
                for (int i = j; i >= 1; i--) l[i] += l[i - 1] * xi_inv;
            }
            alpha0 -= 1.0 / q;
            xistar_inv = -l[1] - alpha0;
            hsum += tau[q - 1];
            xi_inv = h / hsum;
            alpha0_hat = -l[1] - xi_inv;
            for (int i = q; i >= 1; i--) l[i] += l[i - 1] * xistar_inv;
        }
        const double A1 = 1.0 - alpha0_hat + alpha0;
        const double A2 = 1.0 + q * A1;
        tq[2] = fabs(A1 / (alpha0 * A2));
        tq[5] = fabs(A2 * xistar_inv / (l[q] * xi_inv));
        if (qwait == 1) {
            if (q > 1) {
                const double C = xistar_inv / l[q];
                const double A3 = alpha0 + 1.0 / q;
                const double A4 = alpha0_hat + xi_inv;
                const double Cpinv = (1.0 - A4 + A3) / A3;
                tq[1] = fabs(C * Cpinv);
            } else {
                tq[1] = 1.0;
            }
            hsum += tau[q];
            xi_inv = h / hsum;
            const double A5 = alpha0 - 1.0 / (q + 1);
            const double A6 = alpha0_hat - xi_inv;
            const double Cppinv = (1.0 - A6 + A5) / A2;
            tq[3] = fabs(Cppinv / (xi_inv * (q + 2) * A5));
        }
        tq[4] = BDF_CORTES / tq[2];
        rl1 = 1.0 / l[1];
        gamma = h * rl1;
        if (nst == 0) gammap = gamma;
        gamrat = (nst > 0) ? gamma / gammap : 1.0;
    }

    //! Order change by +-1: adjust the history so that it represents the new-order polynomial.
    __device__ __forceinline__ void adjustOrder(int dq) {
        if (q == 2 && dq != 1) return;
        for (int i = 0; i <= BDF_QMAX; i++) l[i] = 0.0;
        if (dq == 1) {
            l[2] = 1.0;
            double alpha1 = 1.0, prod = 1.0, xiold = 1.0, alpha0 = -1.0, hsum = hscale;
            if (q > 1) {
                for (int j = 1; j < q; j++) {
                    hsum += tau[j + 1];
                    const double xi = hsum / hscale;
                    prod *= xi;
                    alpha0 -= 1.0 / (j + 1);
                    alpha1 += 1.0 / xi;
                    for (int i = j + 2; i >= 2; i--) l[i] = l[i] * xiold + l[i - 1];
                    xiold = xi;
                }
            }
            const double A1 = (-alpha0 - alpha1) / prod;
            for (int i = g.lane(); i < wk.n; i += G::NT) {
                const double zL = A1 * zn(BDF_QMAX)[i];
                zn(L)[i] = zL;
                for (int j = 2; j <= q; j++) zn(j)[i] += l[j] * zL;
            }
        } else {
            l[2] = 1.0;
            double hsum = 0.0;
            for (int j = 1; j <= q - 2; j++) {
                hsum += tau[j];
                const double xi = hsum / hscale;
                for (int i = j + 2; i >= 2; i--) l[i] = l[i] * xi + l[i - 1];
            }
            for (int i = g.lane(); i < wk.n; i += G::NT) {
                const double zq = zn(q)[i];
                for (int j = 2; j < q; j++) zn(j)[i] -= l[j] * zq;
            }
        }
    }

    //! Modified Newton iteration on G(y) = y - zn[0] - gamma f(y) + rl1 zn[1]... in the accumulated
    //! correction acor = y - zn[0]. Returns 0 when converged, 1 on a (recoverable) failure.
    __device__ __forceinline__ int nonlinearSolve(int nflag) {
        const int n = wk.n;
        bool callSetup = (nflag != NFLAG_FIRST) || nst == 0 || nst >= nstlp + BDF_MSBP
                         || fabs(gamrat - 1.0) > BDF_DGMAX || forceSetup;
        bool jcur = false;
        for (int pass = 0; pass < 2; pass++) {
            for (int i = g.lane(); i < n; i += G::NT) {
                wk.y[i] = zn(0)[i];
                wk.acor[i] = 0.0;
            }
            g.sync();
            if (callSetup) {
                buildNewtonMatrix(g, v, sm, gm, wk, P, gamma); // also f(y) -> wk.f
                nfe += 2;
                nsetups++;
                const int sing = luFactor(g, wk);
                callSetup = false;
                forceSetup = false;
                gamrat = 1.0;
                crate = 1.0;
                gammap = gamma;
                nstlp = nst;
                jcur = true;
                if (sing) return 1;
            } else {
                rhs(wk.y, wk.f);
            }
            double delp = 0.0;
            int m = 0;
            for (;;) {
                for (int i = g.lane(); i < n; i += G::NT) {
                    wk.b[i] = gamma * wk.f[i] - (rl1 * zn(1)[i] + wk.acor[i]);
                }
                g.sync();
                luSolve(g, wk);
                nni++;
                const double sc = (gamrat != 1.0) ? 2.0 / (1.0 + gamrat) : 1.0;
                double s = 0.0;
                for (int i = g.lane(); i < n; i += G::NT) {
                    const double d = wk.b[i] * sc;
                    const double t = d * wk.ewt[i];
                    s += t * t;
                    const double a = wk.acor[i] + d;
                    wk.acor[i] = a;
                    wk.y[i] = zn(0)[i] + a;
                }
                const double del = sqrt(g.sum(s) / n);
                g.sync();
                if (m > 0) crate = fmax(BDF_CRDOWN * crate, del / delp);
                const double dcon = del * fmin(1.0, fmax(crate, crateFloor)) / tq[4];
                if (dcon <= 1.0) {
                    acnrm = (m == 0) ? del : wrms(wk.acor);
                    return 0;
                }
                m++;
                if (m == BDF_MAXCOR || (m >= 2 && del > BDF_RDIV * delp) || !(del == del)) break;
                delp = del;
                rhs(wk.y, wk.f);
            }
            if (jcur) return 1;
            callSetup = true; // the matrix was stale: rebuild it and try once more
        }
        return 1;
    }

    __device__ __forceinline__ void completeStep() {
        nst++;
        for (int i = q; i >= 2; i--) tau[i] = tau[i - 1];
        if (q == 1 && nst > 1) tau[2] = tau[1];
        tau[1] = h;
        const bool saveAcor = (qwait - 1 == 1) && (q != BDF_QMAX);
        for (int i = g.lane(); i < wk.n; i += G::NT) {
            const double a = wk.acor[i];
            for (int j = 0; j <= q; j++) zn(j)[i] += l[j] * a;
            if (saveAcor) zn(BDF_QMAX)[i] = a;
        }
        qwait--;
        if (saveAcor) savedTq5 = tq[5];
        g.sync();
    }

    __device__ __forceinline__ void setEta() {
        if (eta < BDF_THRESH) {
            eta = 1.0;
            hprime = h;
        } else {
            eta = fmin(eta, etamax);
            hprime = h * eta;
        }
    }

    //! Step-size ratio and order for the next step from the local error estimates at orders
    //! q-1, q, q+1.
    __device__ __forceinline__ void prepareNextStep(double dsm) {
        if (etamax == 1.0) {
            qwait = qwait > 2 ? qwait : 2;
            qprime = q;
            hprime = h;
            eta = 1.0;
            return;
        }
        const double etaq = 1.0 / (pow(BDF_BIAS2 * dsm, 1.0 / L) + BDF_ADDON);
        if (qwait != 0) {
            eta = etaq;
            qprime = q;
            setEta();
            return;
        }
        qwait = 2;
        double etaqm1 = 0.0, etaqp1 = 0.0;
        if (q > 1) {
            const double ddn = wrms(zn(q)) * tq[1];
            etaqm1 = 1.0 / (pow(BDF_BIAS1 * ddn, 1.0 / q) + BDF_ADDON);
        }
        if (q != BDF_QMAX && savedTq5 != 0.0) {
            const double cquot = (tq[5] / savedTq5) * pow(h / tau[2], (double) L);
            double s = 0.0;
            for (int i = g.lane(); i < wk.n; i += G::NT) {
                const double t = (wk.acor[i] - cquot * zn(BDF_QMAX)[i]) * wk.ewt[i];
                s += t * t;
            }
            const double dup = sqrt(g.sum(s) / wk.n) * tq[3];
            etaqp1 = 1.0 / (pow(BDF_BIAS3 * dup, 1.0 / (L + 1)) + BDF_ADDON);
        }
        const double etam = fmax(etaqm1, fmax(etaq, etaqp1));
        if (etam < BDF_THRESH) {
            eta = 1.0;
            qprime = q;
            hprime = h;
            return;
        }
        if (etam == etaq) {
            eta = etaq;
            qprime = q;
        } else if (etam == etaqm1) {
            eta = etaqm1;
            qprime = q - 1;
        } else {
            eta = etaqp1;
            qprime = q + 1;
            for (int i = g.lane(); i < wk.n; i += G::NT) zn(BDF_QMAX)[i] = wk.acor[i];
        }
        setEta();
    }

    //! One internal step (CVODE's cvStep). Returns CTB_STATUS_OK or a failure code.
    __device__ __forceinline__ int step() {
        const double savedT = tn;
        int ncf = 0, nef = 0, nflag = NFLAG_FIRST;
        if (nst > 0 && hprime != h) {
            if (qprime != q) {
                adjustOrder(qprime - q);
                q = qprime;
                L = q + 1;
                qwait = L;
            }
            rescale();
        }
        double dsm;
        for (;;) {
            if (!(fabs(h) > 1e-290) || !(h == h)) return CTB_STATUS_STEP_TOO_SMALL;
            predict();
            setCoefficients();
            if (nonlinearSolve(nflag) != 0) {
                ncfn++;
                restore(savedT);
                ncf++;
                etamax = 1.0;
                if (ncf == BDF_MXNCF) return CTB_STATUS_CONV_FAIL;
                eta = BDF_ETACF;
                nflag = NFLAG_PREV_CONV_FAIL;
                rescale();
                continue;
            }
            dsm = acnrm * tq[2];
            if (dsm <= 1.0) break;
            nef++;
            netf++;
            nflag = NFLAG_PREV_ERR_FAIL;
            restore(savedT);
            if (nef == BDF_MXNEF) return CTB_STATUS_ERR_TEST_FAIL;
            etamax = 1.0;
            if (nef <= BDF_MXNEF1) {
                eta = 1.0 / (pow(BDF_BIAS2 * dsm, 1.0 / L) + BDF_ADDON);
                eta = fmax(BDF_ETAMIN, eta);
                if (nef >= BDF_SMALL_NEF) eta = fmin(eta, BDF_ETAMXF);
                if (!(eta == eta)) eta = BDF_ETAMIN;
                rescale();
                continue;
            }
            if (q > 1) {
                eta = BDF_ETAMIN;
                adjustOrder(-1);
                L = q;
                q--;
                qwait = L;
                rescale();
                continue;
            }
            // order 1 and still failing: restart the history from f(tn, zn[0])
            eta = BDF_ETAMIN;
            h *= eta;
            hscale = h;
            qwait = BDF_LONG_WAIT;
            g.sync();
            rhs(zn(0), wk.tmp);
            for (int i = g.lane(); i < wk.n; i += G::NT) zn(1)[i] = h * wk.tmp[i];
            g.sync();
        }
        completeStep();
        prepareNextStep(dsm);
        etamax = BDF_ETAMX2;
        const double t2 = tq[2];
        for (int i = g.lane(); i < wk.n; i += G::NT) wk.acor[i] *= t2;
        g.sync();
        return CTB_STATUS_OK;
    }

    //! Initial step size (CVODE's cvHin): h ~ sqrt(2 / ||y''||) iterated a few times, bounded
    //! below by 100 ulp(t) and above by 0.1 |tout - t0| and 0.1 |y|/|y'|. Requires zn[0] = y0,
    //! zn[1] = f(t0, y0) and ewt. Returns 0 if tout is too close to t0.
    __device__ __forceinline__ double initialStep(double t0, double tout) {
        const int n = wk.n;
        const double tdist = fabs(tout - t0);
        const double tround = 2.220446049250313e-16 * fmax(fabs(t0), fabs(tout));
        if (tdist < 2.0 * tround) return 0.0;
        const double hlb = 100.0 * tround;
        double hubInv = 0.0;
        for (int i = g.lane(); i < n; i += G::NT) {
            const double den = 0.1 * fabs(zn(0)[i]) + 1.0 / wk.ewt[i];
            hubInv = fmax(hubInv, fabs(zn(1)[i]) / den);
        }
        hubInv = g.max(hubInv);
        double hub = 0.1 * tdist;
        if (hub * hubInv > 1.0) hub = 1.0 / hubInv;
        if (hub < hlb) return sqrt(hlb * hub);
        double hg = sqrt(hlb * hub);
        double hnew = hg;
        for (int count = 1; count <= 4; count++) {
            for (int i = g.lane(); i < n; i += G::NT) wk.y[i] = hg * zn(1)[i] + zn(0)[i];
            g.sync();
            rhs(wk.y, wk.tmp);
            double s = 0.0;
            for (int i = g.lane(); i < n; i += G::NT) {
                const double t = (wk.tmp[i] - zn(1)[i]) / hg * wk.ewt[i];
                s += t * t;
            }
            const double yddnrm = sqrt(g.sum(s) / n);
            if (!(yddnrm == yddnrm)) { // RHS not finite at the trial point: shrink and retry
                hg *= 0.2;
                hnew = hg;
                continue;
            }
            hnew = (yddnrm * hub * hub > 2.0) ? sqrt(2.0 / yddnrm) : sqrt(hg * hub);
            if (count == 4) break;
            const double hrat = hnew / hg;
            if (hrat > 0.5 && hrat < 2.0) break;
            if (count > 1 && hrat > 2.0) {
                hnew = hg;
                break;
            }
            hg = hnew;
        }
        double h0 = 0.5 * hnew;
        if (h0 < hlb) h0 = hlb;
        if (h0 > hub) h0 = hub;
        return h0;
    }

    //! (Re)initialise at (t0, zn[0]) — CVodeInit/CVodeReInit plus the first-call part of CVode.
    __device__ __forceinline__ int start(double t0, double tout) {
        q = 1; L = 2; qwait = L; qprime = 1;
        nst = 0; nstlp = 0; forceSetup = false;
        etamax = BDF_ETAMX1;
        eta = 1.0; crate = 1.0; acnrm = 0.0; savedTq5 = 0.0; gammap = 0.0; gamrat = 1.0; gamma = 0.0; rl1 = 1.0;
        for (int i = 0; i <= BDF_QMAX; i++) { tau[i] = 0.0; tq[i] = 0.0; l[i] = 0.0; }
        tn = t0;
        setEwt(zn(0));
        rhs(zn(0), zn(1));
        const double h0 = initialStep(t0, tout);
        if (!(h0 > 0.0)) return CTB_STATUS_STEP_TOO_SMALL;
        h = h0;
        hscale = h;
        hprime = h;
        for (int i = g.lane(); i < wk.n; i += G::NT) zn(1)[i] *= h;
        g.sync();
        return CTB_STATUS_OK;
    }

    //! y(t) from the Nordsieck polynomial (CVodeGetDky, k = 0) into out[].
    __device__ __forceinline__ void interpolate(double t, double* out) const {
        const double s = (t - tn) / h;
        for (int i = g.lane(); i < wk.n; i += G::NT) {
            double acc = zn(q)[i];
            for (int j = q - 1; j >= 0; j--) acc = acc * s + zn(j)[i];
            out[i] = acc;
        }
    }

    // ---- persistence between advance() calls: [6n Nordsieck][BDF_HIST_SCALARS scalars]
    __device__ __forceinline__ void save(double* hist) const {
        const int n = wk.n;
        for (int i = g.lane(); i < 6 * n; i += G::NT) hist[i] = wk.zn[i];
        if (g.lane() == 0) {
            double* s = hist + 6 * n;
            s[0] = q; s[1] = qwait; s[2] = qprime; s[3] = nst; s[4] = nstlp;
            s[5] = tn; s[6] = h; s[7] = hscale; s[8] = hprime; s[9] = eta; s[10] = etamax;
            s[11] = gammap; s[12] = savedTq5; s[13] = crate; s[14] = atolS;
            for (int i = 0; i <= BDF_QMAX; i++) s[16 + i] = tau[i];
            for (int i = 0; i <= BDF_QMAX; i++) s[22 + i] = tq[i];
        }
    }
    __device__ __forceinline__ void load(const double* hist) {
        const int n = wk.n;
        for (int i = g.lane(); i < 6 * n; i += G::NT) wk.zn[i] = hist[i];
        const double* s = hist + 6 * n;
        q = (int) s[0]; qwait = (int) s[1]; qprime = (int) s[2]; nst = (int) s[3]; nstlp = (int) s[4];
        L = q + 1;
        tn = s[5]; h = s[6]; hscale = s[7]; hprime = s[8]; eta = s[9]; etamax = s[10];
        gammap = s[11]; savedTq5 = s[12]; crate = s[13]; atolS = s[14];
        for (int i = 0; i <= BDF_QMAX; i++) tau[i] = s[16 + i];
        for (int i = 0; i <= BDF_QMAX; i++) tq[i] = s[22 + i];
        for (int i = 0; i <= BDF_QMAX; i++) l[i] = 0.0;
        acnrm = 0.0; gamma = gammap; gamrat = 1.0; rl1 = 1.0;
        forceSetup = true;
        g.sync();
    }
};

} // namespace ctb
