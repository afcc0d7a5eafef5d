// bdf_batch.cuh — the BDF/Newton controller of BatchReactorNet as a RESUMABLE per-reactor state
// machine, for the batch-synchronous integrator of reactor_batch.cu.
//
// The integrator is cut at the points where it needs the chemistry — a right-hand side, or a new
// Newton matrix — and everything between two such points is ONE invocation of BatchBdf::run() by
// ONE WARP:
//
//     control (warp/slot)  ->  rhs (8 lanes/slot)  ->  jacobian (group/flagged slot)  ->  invert
//            ^----------------------------------------------------------------------------|
//
// so that the expensive part (rates) runs as a throughput kernel over all slots at once, and the
// per-slot data (Nordsieck array, Newton matrix: ~28 KB for gri30) lives in HBM, one contiguous
// block per slot. The Newton systems are solved with the EXPLICIT INVERSE of I - gamma J (computed
// once per setup by a register-resident Gauss-Jordan kernel): a solve is then a matrix-vector
// product whose matrix is streamed from HBM/L2 with coalesced loads and used exactly once, with no
// dependent chain — the triangular solves of the first version (two shuffles per pivot, 108 pivots)
// were two thirds of a controller visit.
// A slot that finishes its reactor pulls the next one from a global queue (reactors differ ~10x in
// step count), so all slots stay busy until the queue is empty.
//
// PROVENANCE. The step-size / order controller (setCoefficients, adjustOrder, prepareNextStep, the
// initial-step iteration, the error-test and convergence-failure handling) is a port of the scalar
// controller arithmetic of SUNDIALS CVODES, ext/sundials/src/cvodes/cvodes.c (cvSetBDF, cvSetTqBDF,
// cvIncreaseBDF, cvDecreaseBDF, cvHin, cvUpperBoundH0, cvPrepareNextStep, cvComputeEtaqm1/p1,
// cvChooseEta, cvNlsNewton's convergence test), re-expressed as a resumable state machine so that
// step counts are comparable with ReactorNet::advance. SUNDIALS is
//   Copyright (c) 2002-2025, Lawrence Livermore National Security and Southern Methodist University.
//   All rights reserved. SPDX-License-Identifier: BSD-3-Clause (see ext/sundials/LICENSE).
// What is original here: the state-machine inversion, the slot layout, the lane-cyclic vector
// arithmetic, the explicit-inverse Newton solve and the analytic Newton matrix.
// The arithmetic is __host__ __device__: tests/emul runs it on the CPU.
#pragma once

#include "bdf.cuh"

namespace ctb {

enum BatchPc : int {
    PC_IDLE = 0,      // no reactor left for this slot
    PC_NEW = 1,       // take the next reactor from the queue
    PC_START_F0 = 2,  // f(y0) requested (CVodeInit)
    PC_HIN = 3,       // f(y0 + hg f0) requested (cvHin)
    PC_NEWTON = 4,    // f(y) requested inside the Newton iteration (and LU factors if flagged)
    PC_RESTART1 = 5,  // f(zn0) requested by the order-1 restart after repeated error-test failures
    // transient states (never stored)
    X_ADVANCE = 10, X_ATTEMPT = 11, X_NLS_PASS = 12, X_NLS_OK = 13, X_NLS_FAIL = 14, X_STEP_OK = 15,
    X_START_DONE = 16, X_FINISH = 17
};

//! Per-slot controller memory.
struct BatchCtl {
    int pc, status, q, L, qwait, qprime, nst, nstlp, nst0, ncf, nef, nflag, m, pass, jcur, callSetup,
        needSing, forceSetup, hinCount, ignPending;
    int nfe, nsetups, nni, ncfn, netf;
    int njac; // Jacobian evaluations (one per setup)
    int sing; // written by the invert kernel: non-zero = singular Newton matrix
    long long r;
    double P, atolS, tn, h, hscale, hprime, eta, etamax, gamma, gammap, gamrat, crate, acnrm, savedTq5, rl1,
        delp, savedT, dsm, hg, hnew, hub, hlb, h0, tPrev, TPrev, Tign;
    double tau[BDF_QMAX + 1], tq[BDF_QMAX + 1], l[BDF_QMAX + 1];
};

//! Slot-major integrator memory. Every slot owns ONE contiguous, 128-byte aligned block
//!     [ BatchCtl | vec: zn[0..5], acor, ewt, y, f (10 n doubles) | M (n*n doubles) ]
//! so that the control kernel can bring its working set ([0, oM): controller + vectors) into shared
//! memory with a single TMA bulk copy and write it back with another. y = state the next right-hand
//! side is evaluated at, f = right-hand side at y. M, column-major (col * n + row): the jacobian
//! kernel leaves J = d(ydot)/dy there, the invert kernel replaces it in place by (I - gamma J)^-1.
#define SLOT_VECS 10
struct SlotMem {
    unsigned char* slots;
    unsigned long long slotBytes; // multiple of 128
    unsigned int oVec, oM;  // byte offsets inside a slot block (multiples of 16)
    int* pc;        // [S] copy of ctl.pc (lets kernels skip idle slots without touching the block)
    int* setupCount; // [2]: the control kernel of round r appends to [r & 1] and clears [(r + 1) & 1]
    int* setupList; // [S] slots that need a new Newton matrix in this round
    int* nActive;   // slots that are not idle
    long long* prof; // optional cycle counters of the setup kernel phases (CTB_PROFILE), or null
    int S, n;
    int rtype;      // reactor model (reactor.h)
    int parity;     // round & 1
    CTB_HD unsigned char* block(int s) const { return slots + (size_t) s * slotBytes; }
    CTB_HD BatchCtl* ctl(int s) const { return reinterpret_cast<BatchCtl*>(block(s)); }
    CTB_HD double* vec(int s) const { return reinterpret_cast<double*>(block(s) + oVec); }
    CTB_HD double* M(int s) const { return reinterpret_cast<double*>(block(s) + oM); }
};

//! Fills the layout fields of a SlotMem for n equations; returns the bytes per slot.
inline unsigned long long slotLayout(int n, SlotMem& m)
{
    auto up = [](unsigned long long x, unsigned long long a) { return (x + a - 1) / a * a; };
    m.n = n;
    m.oVec = (unsigned int) up(sizeof(BatchCtl), 16);
    m.oM = (unsigned int) up(m.oVec + sizeof(double) * SLOT_VECS * n, 128);
    m.slotBytes = up(m.oM + sizeof(double) * (unsigned long long) n * n, 128);
    return m.slotBytes;
}

CTB_HD long long fetchAddLL(unsigned long long* p)
{
#ifdef __CUDA_ARCH__
    return (long long) atomicAdd(p, 1ULL);
#else
    return (long long) ((*p)++);
#endif
}
CTB_HD int fetchAddInt(int* p, int v)
{
#ifdef __CUDA_ARCH__
    return atomicAdd(p, v);
#else
    const int o = *p;
    *p += v;
    return o;
#endif
}

// The elements a lane owns (i = lane, lane + NT, ...). (A fully unrolled, predicated form measured
// slower in the control kernel than this plain loop.)
#define CTB_FOR_OWN(i) for (int i = g.lane(); i < n; i += G::NT)

//! The host stand-in for a cooperative group: one "lane" (tests/emul).
struct Solo {
    static constexpr int NT = 1;
    CTB_HD int lane() const { return 0; }
    CTB_HD void sync() const {}
    CTB_HD double sum(double x) const { return x; }
    CTB_HD double max(double x) const { return x; }
};

#ifdef __CUDACC__
// ---- warp-level TMA helpers (1-D bulk copies completing on an mbarrier)
__device__ __forceinline__ void bulkLoad(void* dstSmem, const void* srcGlobal, uint32_t bytes, unsigned long long* bar)
{
    const uint32_t barAddr = (uint32_t) __cvta_generic_to_shared(bar);
    const uint32_t dstAddr = (uint32_t) __cvta_generic_to_shared(dstSmem);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(barAddr), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dstAddr), "l"(srcGlobal), "r"(bytes), "r"(barAddr)
                 : "memory");
}
__device__ __forceinline__ void barWait(unsigned long long* bar, uint32_t parity)
{
    const uint32_t barAddr = (uint32_t) __cvta_generic_to_shared(bar);
    uint32_t done = 0;
    while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(barAddr), "r"(parity)
            : "memory");
    }
}
#endif

//! Per-warp shared-memory ring through which the control kernel streams the slot's Newton-matrix inverse
//! for the matrix-vector product of the Newton solve: NBUF chunks of chunkCols columns in flight as TMA
//! bulk copies (no registers held across the DRAM latency), the first NBUF requested together with the
//! slot's prefix at the start of the visit.
struct MatRing {
    static constexpr int NBUF = 2;
    double* buf;              // NBUF x chunkCols x n doubles of shared memory
    unsigned long long* bar;  // NBUF mbarriers
    const unsigned char* src; // the slot's matrix in HBM
    int chunkCols, nChunks;
    uint32_t chunkBytes, lastBytes; // multiples of 16
    uint32_t parity;          // bit i: phase parity of barrier i (persists across the warp's visits)
#ifdef __CUDACC__
    __device__ __forceinline__ void request(int ch) const {
        const int bi = ch % NBUF;
        bulkLoad(reinterpret_cast<unsigned char*>(buf) + (size_t) bi * chunkBytes, src + (size_t) ch * chunkBytes,
                 ch == nChunks - 1 ? lastBytes : chunkBytes, bar + bi);
    }
    //! lane 0, at the start of a visit that resumes in the Newton solve
    __device__ __forceinline__ void prime(const unsigned char* matrix) {
        src = matrix;
        for (int ch = 0; ch < NBUF && ch < nChunks; ch++) request(ch);
    }
    __device__ __forceinline__ void wait(int bi) {
        barWait(bar + bi, (parity >> bi) & 1u);
        parity ^= 1u << bi;
    }
    //! consumes the primed chunks without using them (the visit leaves the Newton solve early)
    __device__ __forceinline__ void drain() {
        for (int ch = 0; ch < NBUF && ch < nChunks; ch++) wait(ch);
    }
#endif
};

//! One cooperative group G (a warp on the device) advances slot s until it needs a right-hand
//! side (returns with ctl.pc set) or runs out of reactors. Controller scalars are replicated in
//! every lane (reductions return identical values), vectors are distributed lane-cyclically:
//! NPL = ceil(n / G::NT) elements per lane.
template <class G, int NPL>
struct BatchBdf {
    const G& g;
    const ReactorBatch& p;
    const SlotMem& mem;
    const int n, s;
    BatchCtl* const ctlPtr;  // this slot's controller memory: in HBM, or its shared-memory copy
    double* const vec;       // the slot's vectors, likewise
    const double* const MPtr; // (I - gamma J)^-1 of the slot, column-major, in HBM
    double* const xs;  // n doubles of scratch (only used by the one-lane solve)
    MatRing* const ring; // shared-memory ring of the matrix-vector product, or null: loads through registers
#ifdef CTB_CTL_IN_SMEM
    BatchCtl& c;       // controller memory is used where it lies (shared memory): fewer registers
#else
    BatchCtl c;        // controller memory in registers for the duration of run()
#endif

    //! blk: the slot's [ctl | vec] prefix wherever it currently lives; Minv: the slot's matrix
    CTB_HD BatchBdf(const G& g_, const ReactorBatch& p_, const SlotMem& m_, int s_, unsigned char* blk,
                    const double* Minv, double* xs_, MatRing* ring_ = nullptr)
        : g(g_), p(p_), mem(m_), n(p_.neq), s(s_), ctlPtr(reinterpret_cast<BatchCtl*>(blk)),
          vec(reinterpret_cast<double*>(blk + m_.oVec)), MPtr(Minv), xs(xs_), ring(ring_)
#ifdef CTB_CTL_IN_SMEM
          , c(*reinterpret_cast<BatchCtl*>(blk))
#endif
    {}

    CTB_HD double& ZN(int j, int i) const { return vec[j * n + i]; }
    CTB_HD double& AC(int i) const { return vec[6 * n + i]; }
    CTB_HD double& EW(int i) const { return vec[7 * n + i]; }
    CTB_HD double& Y(int i) const { return vec[8 * n + i]; }
    CTB_HD double& F(int i) const { return vec[9 * n + i]; }

    CTB_HD void setEwt() const {
        // species rows: c.atolS (MoleReactor::updateDefaultTolerances, MoleReactor.cpp:64-76); T, V, m rows: atol
        CTB_FOR_OWN(i) EW(i) = 1.0 / (p.rtol * fabs(ZN(0, i)) + (i < p.sIdx ? p.atol : c.atolS));
    }
    //! zn <- Pascal-triangle product (SUB = false: the prediction) or its inverse (SUB = true), for order Q. The
    //! order is a compile-time constant here (run() dispatches on c.q): Q + 1 loads, Q (Q + 1) / 2 additions in the
    //! order of cvPredict / cvRestore, Q stores, no predicates.
    template <int Q, bool SUB>
    CTB_HD void pascalQ() {
        CTB_FOR_OWN(i) {
            double z[Q + 1];
#pragma unroll
            for (int j = 0; j <= Q; j++) z[j] = ZN(j, i);
#pragma unroll
            for (int k = 1; k <= Q; k++) {
#pragma unroll
                for (int j = Q; j >= k; j--) z[j - 1] = SUB ? z[j - 1] - z[j] : z[j - 1] + z[j];
            }
#pragma unroll
            for (int j = 0; j < Q; j++) ZN(j, i) = z[j];
        }
    }
    template <bool SUB>
    CTB_HD void pascal() {
        switch (c.q) {
        case 1: pascalQ<1, SUB>(); break;
        case 2: pascalQ<2, SUB>(); break;
        case 3: pascalQ<3, SUB>(); break;
        case 4: pascalQ<4, SUB>(); break;
        default: pascalQ<5, SUB>(); break;
        }
    }
    CTB_HD void predict() {
        c.tn += c.h;
        pascal<false>();
    }
    CTB_HD void restore() {
        c.tn = c.savedT;
        pascal<true>();
    }
    CTB_HD void rescale() {
        double factor = c.eta;
        for (int j = 1; j <= c.q; j++) {
            CTB_FOR_OWN(i) ZN(j, i) *= factor;
            factor *= c.eta;
        }
        c.h = c.hscale * c.eta;
        c.hscale = c.h;
    }
    CTB_HD double wrmsAcor() const {
        double sum = 0.0;
        CTB_FOR_OWN(i) {
            const double t = AC(i) * EW(i);
            sum += t * t;
        }
        return sqrt(g.sum(sum) / n);
    }

    // The history arrays tau, tq, l are only ever indexed with compile-time constants below (fully
    // unrolled, predicated loops; get6 for the few run-time indices), so that the whole controller
    // memory lives in registers: a per-lane copy in local memory would not fit the L1 that the
    // slot buffers leave.
    CTB_HD static double get6(const double (&a)[BDF_QMAX + 1], int i) {
        double r = a[0];
#pragma unroll
        for (int k = 1; k <= BDF_QMAX; k++) if (i == k) r = a[k];
        return r;
    }

    // Bdf::setCoefficients of bdf.cuh
    CTB_HD void setCoefficients() {
        double (&l)[BDF_QMAX + 1] = c.l;
        double (&tq)[BDF_QMAX + 1] = c.tq;
        const double (&tau)[BDF_QMAX + 1] = c.tau;
        const int q = c.q;
        const double h = c.h;
#pragma unroll
        for (int i = 0; i <= BDF_QMAX; i++) l[i] = 0.0;
        l[0] = l[1] = 1.0;
        double xi_inv = 1.0, xistar_inv = 1.0, alpha0 = -1.0, alpha0_hat = -1.0, hsum = h;
        if (q > 1) {
#pragma unroll
            for (int j = 2; j < BDF_QMAX; j++) {
                if (j < q) {
                    hsum += tau[j - 1];
                    xi_inv = h / hsum;
                    alpha0 -= 1.0 / j;
#pragma unroll
                    for (int i = BDF_QMAX; i >= 1; i--) if (i <= j) l[i] += l[i - 1] * xi_inv;
                }
            }
            alpha0 -= 1.0 / q;
            xistar_inv = -l[1] - alpha0;
            hsum += get6(tau, q - 1);
            xi_inv = h / hsum;
            alpha0_hat = -l[1] - xi_inv;
#pragma unroll
            for (int i = BDF_QMAX; i >= 1; i--) if (i <= q) l[i] += l[i - 1] * xistar_inv;
        }
        const double lq = get6(l, q);
        const double A1 = 1.0 - alpha0_hat + alpha0;
        const double A2 = 1.0 + q * A1;
        tq[2] = fabs(A1 / (alpha0 * A2));
        tq[5] = fabs(A2 * xistar_inv / (lq * xi_inv));
        if (c.qwait == 1) {
            if (q > 1) {
                const double C = xistar_inv / lq;
                const double A3 = alpha0 + 1.0 / q;
                const double A4 = alpha0_hat + xi_inv;
                const double Cpinv = (1.0 - A4 + A3) / A3;
                tq[1] = fabs(C * Cpinv);
            } else {
                tq[1] = 1.0;
            }
            hsum += get6(tau, q);
            xi_inv = h / hsum;
            const double A5 = alpha0 - 1.0 / (q + 1);
            const double A6 = alpha0_hat - xi_inv;
            const double Cppinv = (1.0 - A6 + A5) / A2;
            tq[3] = fabs(Cppinv / (xi_inv * (q + 2) * A5));
        }
        tq[4] = BDF_CORTES / tq[2];
        c.rl1 = 1.0 / l[1];
        c.gamma = h * c.rl1;
        if (c.nst == 0) c.gammap = c.gamma;
        c.gamrat = (c.nst > 0) ? c.gamma / c.gammap : 1.0;
    }

    // Bdf::adjustOrder of bdf.cuh
    CTB_HD void adjustOrder(int dq) {
        double (&l)[BDF_QMAX + 1] = c.l;
        const double (&tau)[BDF_QMAX + 1] = c.tau;
        const int q = c.q;
        if (q == 2 && dq != 1) return;
#pragma unroll
        for (int i = 0; i <= BDF_QMAX; i++) l[i] = 0.0;
        if (dq == 1) {
            l[2] = 1.0;
            double alpha1 = 1.0, prod = 1.0, xiold = 1.0, alpha0 = -1.0, hsum = c.hscale;
            if (q > 1) {
#pragma unroll
                for (int j = 1; j < BDF_QMAX - 1; j++) { // q + 1 <= BDF_QMAX here, so j <= q - 1 <= 3
                    if (j < q) {
                        hsum += tau[j + 1];
                        const double xi = hsum / c.hscale;
                        prod *= xi;
                        alpha0 -= 1.0 / (j + 1);
                        alpha1 += 1.0 / xi;
#pragma unroll
                        for (int i = BDF_QMAX; i >= 2; i--) if (i <= j + 2) l[i] = l[i] * xiold + l[i - 1];
                        xiold = xi;
                    }
                }
            }
            const double A1 = (-alpha0 - alpha1) / prod;
            CTB_FOR_OWN(i) {
                const double zL = A1 * ZN(BDF_QMAX, i);
                ZN(c.L, i) = zL;
#pragma unroll
                for (int j = 2; j <= BDF_QMAX; j++) if (j <= q) ZN(j, i) += l[j] * zL;
            }
        } else {
            l[2] = 1.0;
            double hsum = 0.0;
#pragma unroll
            for (int j = 1; j <= BDF_QMAX - 2; j++) {
                if (j <= q - 2) {
                    hsum += tau[j];
                    const double xi = hsum / c.hscale;
#pragma unroll
                    for (int i = BDF_QMAX; i >= 2; i--) if (i <= j + 2) l[i] = l[i] * xi + l[i - 1];
                }
            }
            CTB_FOR_OWN(i) {
                const double zq = ZN(q, i);
#pragma unroll
                for (int j = 2; j < BDF_QMAX; j++) if (j < q) ZN(j, i) -= l[j] * zq;
            }
        }
    }

    CTB_HD void completeStep() {
        const int q = c.q;
        c.nst++;
#pragma unroll
        for (int i = BDF_QMAX; i >= 2; i--) if (i <= q) c.tau[i] = c.tau[i - 1];
        if (q == 1 && c.nst > 1) c.tau[2] = c.tau[1];
        c.tau[1] = c.h;
        const bool saveAcor = (c.qwait - 1 == 1) && (q != BDF_QMAX);
        CTB_FOR_OWN(i) {
            const double a = AC(i);
#pragma unroll
            for (int j = 0; j <= BDF_QMAX; j++) if (j <= q) ZN(j, i) += c.l[j] * a;
            if (saveAcor) ZN(BDF_QMAX, i) = a;
        }
        c.qwait--;
        if (saveAcor) c.savedTq5 = c.tq[5];
    }
    //! x^(1/k) for the step-size selection (k = 1 .. 7, x >= 0): square roots where k is a power of two, else
    //! exp(log(x) / k). The library pow() costs about twice as many instructions and the result only steers
    //! the step size (cvodes.c cvComputeEtaq*: SUNRpowerR).
    CTB_HD static double rootK(double x, int k) {
        if (k == 1) return x;
        if (k == 2) return sqrt(x);
        if (k == 4) return sqrt(sqrt(x));
        return x > 0.0 ? exp(log(x) / k) : 0.0;
    }
    CTB_HD void setEta() {
        if (c.eta < BDF_THRESH) {
            c.eta = 1.0;
            c.hprime = c.h;
        } else {
            c.eta = fmin(c.eta, c.etamax);
            c.hprime = c.h * c.eta;
        }
    }
    CTB_HD void prepareNextStep(double dsm) {
        const int q = c.q, L = c.L;
        if (c.etamax == 1.0) {
            c.qwait = c.qwait > 2 ? c.qwait : 2;
            c.qprime = q;
            c.hprime = c.h;
            c.eta = 1.0;
            return;
        }
        const double etaq = 1.0 / (rootK(BDF_BIAS2 * dsm, L) + BDF_ADDON);
        if (c.qwait != 0) {
            c.eta = etaq;
            c.qprime = q;
            setEta();
            return;
        }
        c.qwait = 2;
        double etaqm1 = 0.0, etaqp1 = 0.0;
        if (q > 1) {
            double sum = 0.0;
            CTB_FOR_OWN(i) {
                const double t = ZN(q, i) * EW(i);
                sum += t * t;
            }
            const double ddn = sqrt(g.sum(sum) / n) * c.tq[1];
            etaqm1 = 1.0 / (rootK(BDF_BIAS1 * ddn, q) + BDF_ADDON);
        }
        if (q != BDF_QMAX && c.savedTq5 != 0.0) {
            const double hr = c.h / c.tau[2];
            double hrL = hr; // (h / tau_2)^L, L = 2 .. 6
            for (int k = 1; k < L; k++) hrL *= hr;
            const double cquot = (c.tq[5] / c.savedTq5) * hrL;
            double sum = 0.0;
            CTB_FOR_OWN(i) {
                const double t = (AC(i) - cquot * ZN(BDF_QMAX, i)) * EW(i);
                sum += t * t;
            }
            const double dup = sqrt(g.sum(sum) / n) * c.tq[3];
            etaqp1 = 1.0 / (rootK(BDF_BIAS3 * dup, L + 1) + BDF_ADDON);
        }
        const double etam = fmax(etaqm1, fmax(etaq, etaqp1));
        if (etam < BDF_THRESH) {
            c.eta = 1.0;
            c.qprime = q;
            c.hprime = c.h;
            return;
        }
        if (etam == etaq) {
            c.eta = etaq;
            c.qprime = q;
        } else if (etam == etaqm1) {
            c.eta = etaqm1;
            c.qprime = q - 1;
        } else {
            c.eta = etaqp1;
            c.qprime = q + 1;
            CTB_FOR_OWN(i) ZN(BDF_QMAX, i) = AC(i);
        }
        setEta();
    }

    CTB_HD long long bcastLL(long long v) const {
#ifdef __CUDA_ARCH__
        if (G::NT > 1) return __shfl_sync(0xffffffffu, v, 0);
#endif
        return v;
    }

    //! One lane (host emulation): xs[0..n) <- Minv * F.
    CTB_HD void solveSerial() const {
        const double* M = MPtr;
        for (int i = 0; i < n; i++) xs[i] = 0.0;
        for (int j = 0; j < n; j++) {
            const double bj = F(j);
            const double* cj = M + (size_t) j * n;
            for (int i = 0; i < n; i++) xs[i] += cj[i] * bj;
        }
    }

#ifdef __CUDA_ARCH__
    //! A warp: x = Minv * F with row i = lane + 32 j of x in xr[j]. The right-hand side F is read
    //! from shared memory (broadcast), the matrix is streamed from global memory column by column
    //! (consecutive lanes, consecutive addresses; every element is used exactly once, so the loads
    //! bypass L1 allocation), eight columns in flight per lane and two accumulators per row.
    __device__ __forceinline__ void gemvWarp(double (&xr)[NPL]) const {
        const int lane = g.lane();
        const double* __restrict__ Mg = MPtr + lane;
        const double* __restrict__ b = vec + 9 * n;
        const bool lastOk = lane + 32 * (NPL - 1) < n;
        double a0[NPL], a1[NPL];
#pragma unroll
        for (int j = 0; j < NPL; j++) { a0[j] = 0.0; a1[j] = 0.0; }
        int col = 0;
#pragma unroll 1
        for (; col + 8 <= n; col += 8) {
            double m[8][NPL];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const double* ck = Mg + (size_t) (col + u) * n;
#pragma unroll
                for (int j = 0; j < NPL; j++) m[u][j] = (j < NPL - 1 || lastOk) ? __ldcs(ck + 32 * j) : 0.0;
            }
#pragma unroll
            for (int u = 0; u < 8; u += 2) {
                const double b0 = b[col + u], b1 = b[col + u + 1];
#pragma unroll
                for (int j = 0; j < NPL; j++) {
                    a0[j] = fma(m[u][j], b0, a0[j]);
                    a1[j] = fma(m[u + 1][j], b1, a1[j]);
                }
            }
        }
        for (; col < n; col++) {
            const double* ck = Mg + (size_t) col * n;
            const double bc = b[col];
#pragma unroll
            for (int j = 0; j < NPL; j++) {
                const double mv = (j < NPL - 1 || lastOk) ? __ldcs(ck + 32 * j) : 0.0;
                a0[j] = fma(mv, bc, a0[j]);
            }
        }
#pragma unroll
        for (int j = 0; j < NPL; j++) xr[j] = a0[j] + a1[j];
    }

    //! The same product with the matrix arriving through the shared-memory ring (MatRing): chunk ch + NBUF is
    //! requested as soon as chunk ch has been consumed. chunkCols is even: even columns feed a0, odd ones a1.
    //! Pointers advance by additions only (n is a run-time value), and rows past n read a clamped,
    //! in-range element whose result is never used, so the inner loop carries no predicates.
    __device__ __forceinline__ void gemvRing(double (&xr)[NPL]) const {
        const int lane = g.lane();
        MatRing& rg = *ring;
        int off[NPL];
#pragma unroll
        for (int j = 0; j < NPL; j++) off[j] = min(lane + 32 * j, n - 1);
        double a0[NPL], a1[NPL];
#pragma unroll
        for (int j = 0; j < NPL; j++) { a0[j] = 0.0; a1[j] = 0.0; }
        const double* __restrict__ bp = vec + 9 * n;
        const int cc = rg.chunkCols;
        const uint32_t chunkDoubles = rg.chunkBytes >> 3;
        const double* bufp = rg.buf;
        int bi = 0, left = n;
#pragma unroll 1
        for (int ch = 0; ch < rg.nChunks; ch++) {
            rg.wait(bi);
            const double* __restrict__ col = bufp;
            const int nc = min(cc, left);
            int u = 0;
#pragma unroll 2
            for (; u + 2 <= nc; u += 2) {
                const double b0 = bp[0], b1 = bp[1];
                const double* __restrict__ col1 = col + n;
#pragma unroll
                for (int j = 0; j < NPL; j++) {
                    a0[j] = fma(col[off[j]], b0, a0[j]);
                    a1[j] = fma(col1[off[j]], b1, a1[j]);
                }
                col = col1 + n;
                bp += 2;
            }
            if (u < nc) {
                const double b0 = bp[0];
#pragma unroll
                for (int j = 0; j < NPL; j++) a0[j] = fma(col[off[j]], b0, a0[j]);
                bp += 1;
            }
            left -= nc;
            __syncwarp(); // every lane has read the buffer before it is refilled
            if (lane == 0 && ch + MatRing::NBUF < rg.nChunks) rg.request(ch + MatRing::NBUF);
            bufp += chunkDoubles;
            if (++bi == MatRing::NBUF) { bi = 0; bufp = rg.buf; }
        }
#pragma unroll
        for (int j = 0; j < NPL; j++) xr[j] = a0[j] + a1[j];
    }
#endif

    CTB_HD void finish(int status) {
        const long long r = c.r;
        double* yg = p.y + r * n;
        if (status == CTB_STATUS_OK) {
            // CVodeGetDky, k = 0
            const double sc = (p.tout - c.tn) / c.h;
            CTB_FOR_OWN(i) {
                double acc = ZN(c.q, i);
                for (int j = c.q - 1; j >= 0; j--) acc = acc * sc + ZN(j, i);
                yg[i] = acc;
            }
        } else {
            CTB_FOR_OWN(i) yg[i] = ZN(0, i);
        }
        double* hist = p.hist + r * p.histStride;
        for (int j = 0; j < 6; j++)
            CTB_FOR_OWN(i) hist[j * n + i] = ZN(j, i);
        if (g.lane() != 0) return;
        double* hs = hist + 6 * n;
        hs[0] = c.q; hs[1] = c.qwait; hs[2] = c.qprime; hs[3] = c.nst; hs[4] = c.nstlp;
        hs[5] = c.tn; hs[6] = c.h; hs[7] = c.hscale; hs[8] = c.hprime; hs[9] = c.eta; hs[10] = c.etamax;
        hs[11] = c.gammap; hs[12] = c.savedTq5; hs[13] = c.crate; hs[14] = c.atolS;
#pragma unroll
        for (int i = 0; i <= BDF_QMAX; i++) { hs[16 + i] = c.tau[i]; hs[22 + i] = c.tq[i]; }
        p.status[r] = status;
        long long* st = p.stats + r * CTB_NSTATS;
        st[STAT_STEPS] += c.nst - c.nst0;
        st[STAT_RHS] += c.nfe;
        st[STAT_JAC] += c.njac;
        st[STAT_SETUPS] += c.nsetups;
        st[STAT_NLITERS] += c.nni;
        st[STAT_NCFAILS] += c.ncfn;
        st[STAT_ETFAILS] += c.netf;
        st[STAT_ORDER] = c.q;
    }

    //! Returns when the slot waits for a right-hand side (c.pc >= PC_START_F0) or is idle.
    CTB_HD void run() {
#ifndef CTB_CTL_IN_SMEM
        c = *ctlPtr;
#endif
        if (c.pc == PC_IDLE) return;
        int st = c.pc;
        int finishStatus = CTB_STATUS_OK;
        for (;;) {
            if (st == PC_NEW) {
                long long r = 0;
                for (;;) {
                    if (g.lane() == 0) r = fetchAddLL(p.counter);
                    r = bcastLL(r);
                    if (r >= p.N) break;
                    if (p.status[r] == CTB_STATUS_OK || p.restart) break; // failed earlier: stays failed
                }
                if (r >= p.N) {
                    c.pc = PC_IDLE;
                    if (g.lane() == 0) fetchAddInt(mem.nActive, -1);
                    break;
                }
                c.r = r;
                c.P = p.P[r];
                c.nfe = c.nsetups = c.nni = c.ncfn = c.netf = c.njac = 0;
                c.Tign = p.ignDT > 0.0 ? p.T0[r] + p.ignDT : 0.0;
                c.ignPending = p.ignDT > 0.0 && !(p.tIgn[r] == p.tIgn[r]);
                c.needSing = 0;
                if (p.restart) {
                    const double* yg = p.y + r * n;
                    double moles = 0.0;
                    CTB_FOR_OWN(i) {
                        const double yi = yg[i];
                        ZN(0, i) = yi;
                        Y(i) = yi;
                        if (i >= p.sIdx) moles += yi;
                    }
                    // MoleReactor::updateDefaultTolerances (MoleReactor.cpp:64-76): mole reactors only
                    moles = g.sum(moles);
                    c.atolS = (p.atolUser || !reactorIsMole(p.type) || !(moles > 0.0)) ? p.atol : p.atol * moles;
                    c.q = 1; c.L = 2; c.qwait = 2; c.qprime = 1;
                    c.nst = 0; c.nstlp = 0; c.nst0 = 0; c.forceSetup = 0;
                    c.etamax = BDF_ETAMX1;
                    c.eta = 1.0; c.crate = 1.0; c.acnrm = 0.0; c.savedTq5 = 0.0; c.gammap = 0.0; c.gamrat = 1.0;
                    c.gamma = 0.0; c.rl1 = 1.0;
#pragma unroll
                    for (int i = 0; i <= BDF_QMAX; i++) { c.tau[i] = 0.0; c.tq[i] = 0.0; c.l[i] = 0.0; }
                    c.tn = p.t0;
                    c.h = c.hscale = c.hprime = 0.0;
                    setEwt();
                    c.tPrev = c.tn;
                    g.sync();
                    c.TPrev = ZN(0, p.tIdx);
                    c.nfe++;
                    c.pc = PC_START_F0;
                    break;
                }
                const double* hist = p.hist + r * p.histStride;
                for (int j = 0; j < 6; j++)
                    CTB_FOR_OWN(i) ZN(j, i) = hist[j * n + i];
                const double* hs = hist + 6 * n;
                c.q = (int) hs[0]; c.qwait = (int) hs[1]; c.qprime = (int) hs[2]; c.nst = (int) hs[3];
                c.nstlp = (int) hs[4];
                c.L = c.q + 1;
                c.tn = hs[5]; c.h = hs[6]; c.hscale = hs[7]; c.hprime = hs[8]; c.eta = hs[9]; c.etamax = hs[10];
                c.gammap = hs[11]; c.savedTq5 = hs[12]; c.crate = hs[13]; c.atolS = hs[14];
#pragma unroll
                for (int i = 0; i <= BDF_QMAX; i++) { c.tau[i] = hs[16 + i]; c.tq[i] = hs[22 + i]; c.l[i] = 0.0; }
                c.acnrm = 0.0; c.gamma = c.gammap; c.gamrat = 1.0; c.rl1 = 1.0;
                c.forceSetup = 1;
                c.nst0 = c.nst;
                c.tPrev = c.tn;
                g.sync();
                c.TPrev = ZN(0, p.tIdx);
                st = X_ADVANCE;
            } else if (st == PC_START_F0) {
                // cvHin (Bdf::initialStep of bdf.cuh), first part
                CTB_FOR_OWN(i) ZN(1, i) = F(i);
                const double tdist = fabs(p.tout - p.t0);
                const double tround = 2.220446049250313e-16 * fmax(fabs(p.t0), fabs(p.tout));
                if (tdist < 2.0 * tround) {
                    finishStatus = CTB_STATUS_STEP_TOO_SMALL;
                    st = X_FINISH;
                    continue;
                }
                c.hlb = 100.0 * tround;
                double hubInv = 0.0;
                CTB_FOR_OWN(i) {
                    const double den = 0.1 * fabs(ZN(0, i)) + 1.0 / EW(i);
                    hubInv = fmax(hubInv, fabs(ZN(1, i)) / den);
                }
                hubInv = g.max(hubInv);
                double hub = 0.1 * tdist;
                if (hub * hubInv > 1.0) hub = 1.0 / hubInv;
                c.hub = hub;
                if (hub < c.hlb) {
                    c.h0 = sqrt(c.hlb * hub);
                    st = X_START_DONE;
                    continue;
                }
                c.hg = sqrt(c.hlb * hub);
                c.hnew = c.hg;
                c.hinCount = 1;
                CTB_FOR_OWN(i) Y(i) = c.hg * ZN(1, i) + ZN(0, i);
                c.nfe++;
                c.pc = PC_HIN;
                break;
            } else if (st == PC_HIN) {
                double sum = 0.0;
                CTB_FOR_OWN(i) {
                    const double t = (F(i) - ZN(1, i)) / c.hg * EW(i);
                    sum += t * t;
                }
                const double yddnrm = sqrt(g.sum(sum) / n);
                bool next;
                if (!(yddnrm == yddnrm)) { // RHS not finite at the trial point: shrink and retry
                    c.hg *= 0.2;
                    c.hnew = c.hg;
                    next = c.hinCount < 4;
                } else {
                    c.hnew = (yddnrm * c.hub * c.hub > 2.0) ? sqrt(2.0 / yddnrm) : sqrt(c.hg * c.hub);
                    if (c.hinCount == 4) {
                        next = false;
                    } else {
                        const double hrat = c.hnew / c.hg;
                        if (hrat > 0.5 && hrat < 2.0) {
                            next = false;
                        } else if (c.hinCount > 1 && hrat > 2.0) {
                            c.hnew = c.hg;
                            next = false;
                        } else {
                            c.hg = c.hnew;
                            next = true;
                        }
                    }
                }
                if (next) {
                    c.hinCount++;
                    CTB_FOR_OWN(i) Y(i) = c.hg * ZN(1, i) + ZN(0, i);
                    c.nfe++;
                    c.pc = PC_HIN;
                    break;
                }
                double h0 = 0.5 * c.hnew;
                if (h0 < c.hlb) h0 = c.hlb;
                if (h0 > c.hub) h0 = c.hub;
                c.h0 = h0;
                st = X_START_DONE;
            } else if (st == X_START_DONE) {
                if (!(c.h0 > 0.0)) {
                    finishStatus = CTB_STATUS_STEP_TOO_SMALL;
                    st = X_FINISH;
                    continue;
                }
                c.h = c.hscale = c.hprime = c.h0;
                CTB_FOR_OWN(i) ZN(1, i) *= c.h;
                st = X_ADVANCE;
            } else if (st == X_ADVANCE) {
                if (!(c.tn < p.tout)) {
                    finishStatus = CTB_STATUS_OK;
                    st = X_FINISH;
                    continue;
                }
                if (c.nst - c.nst0 >= p.maxSteps) {
                    finishStatus = CTB_STATUS_TOO_MANY_STEPS;
                    st = X_FINISH;
                    continue;
                }
                if (c.nst > 0) setEwt();
                // Bdf::step
                c.savedT = c.tn;
                c.ncf = 0; c.nef = 0; c.nflag = NFLAG_FIRST;
                if (c.nst > 0 && c.hprime != c.h) {
                    if (c.qprime != c.q) {
                        adjustOrder(c.qprime - c.q);
                        c.q = c.qprime;
                        c.L = c.q + 1;
                        c.qwait = c.L;
                    }
                    rescale();
                }
                st = X_ATTEMPT;
            } else if (st == X_ATTEMPT) {
                if (!(fabs(c.h) > 1e-290) || !(c.h == c.h)) {
                    finishStatus = CTB_STATUS_STEP_TOO_SMALL;
                    st = X_FINISH;
                    continue;
                }
                predict();
                setCoefficients();
                // Bdf::nonlinearSolve
                c.callSetup = (c.nflag != NFLAG_FIRST) || c.nst == 0 || c.nst >= c.nstlp + BDF_MSBP
                              || fabs(c.gamrat - 1.0) > BDF_DGMAX || c.forceSetup;
                c.jcur = 0;
                c.pass = 0;
                st = X_NLS_PASS;
            } else if (st == X_NLS_PASS) {
                CTB_FOR_OWN(i) {
                    Y(i) = ZN(0, i);
                    AC(i) = 0.0;
                }
                c.m = 0;
                c.delp = 0.0;
                if (c.callSetup) {
                    // cvLsSetup (ext/sundials/src/cvodes/cvodes_ls.c:1614-1700) with the Jacobian
                    // re-evaluated at every setup: re-using a saved Jacobian like CVODES was measured
                    // (round 1: ~28 % more steps on the golden gri30 cases, 8.4 k instead of 10.4 k
                    // reactors/s) and dropped. The jacobian kernel evaluates f(y) and f(y + dT e_T).
                    if (g.lane() == 0) mem.setupList[fetchAddInt(mem.setupCount + mem.parity, 1)] = s;
                    c.nfe += 2;
                    c.njac++;
                    c.nsetups++;
                    c.callSetup = 0;
                    c.forceSetup = 0;
                    c.gamrat = 1.0;
                    c.crate = 1.0;
                    c.gammap = c.gamma;
                    c.nstlp = c.nst;
                    c.jcur = 1;
                    c.needSing = 1;
                } else {
                    c.nfe++;
                    c.needSing = 0;
                }
                c.pc = PC_NEWTON;
                break;
            } else if (st == PC_NEWTON) {
                if (c.needSing) {
                    c.needSing = 0;
                    if (c.sing) {
#ifdef __CUDA_ARCH__
                        if (ring) ring->drain();
#endif
                        st = X_NLS_FAIL;
                        continue;
                    }
                }
                c.nni++;
                const double sc = (c.gamrat != 1.0) ? 2.0 / (1.0 + c.gamrat) : 1.0;
                double sum = 0.0;
#ifdef __CUDA_ARCH__
                if (G::NT > 1) {
                    // residual b = gamma f - (rl1 zn1 + acor) overwrites f (its only use), then
                    // x = (I - gamma J)^-1 b as a matrix-vector product
                    CTB_FOR_OWN(i) F(i) = c.gamma * F(i) - (c.rl1 * ZN(1, i) + AC(i));
                    g.sync();
                    double xr[NPL];
                    if (ring) gemvRing(xr); else gemvWarp(xr);
#pragma unroll
                    for (int j = 0; j < NPL; j++) {
                        const int i = g.lane() + G::NT * j;
                        if (i < n) {
                            const double d = xr[j] * sc;
                            const double t = d * EW(i);
                            sum += t * t;
                            const double a = AC(i) + d;
                            AC(i) = a;
                            Y(i) = ZN(0, i) + a;
                        }
                    }
                } else
#endif
                {
                    for (int i = 0; i < n; i++) F(i) = c.gamma * F(i) - (c.rl1 * ZN(1, i) + AC(i));
                    solveSerial();
                    for (int i = 0; i < n; i++) {
                        const double d = xs[i] * sc;
                        const double t = d * EW(i);
                        sum += t * t;
                        const double a = AC(i) + d;
                        AC(i) = a;
                        Y(i) = ZN(0, i) + a;
                    }
                }
                const double del = sqrt(g.sum(sum) / n);
                if (c.m > 0) c.crate = fmax(BDF_CRDOWN * c.crate, del / c.delp);
                const double dcon = del * fmin(1.0, fmax(c.crate, p.crateFloor)) / c.tq[4];
                if (dcon <= 1.0) {
                    c.acnrm = (c.m == 0) ? del : wrmsAcor();
                    st = X_NLS_OK;
                    continue;
                }
                c.m++;
                if (c.m == BDF_MAXCOR || (c.m >= 2 && del > BDF_RDIV * c.delp) || !(del == del)) {
                    if (c.jcur || c.pass == 1) {
                        st = X_NLS_FAIL;
                    } else {
                        c.callSetup = 1; // the matrix was stale: rebuild it and try once more
                        c.pass = 1;
                        st = X_NLS_PASS;
                    }
                    continue;
                }
                c.delp = del;
                c.nfe++;
                c.pc = PC_NEWTON;
                break;
            } else if (st == X_NLS_FAIL) {
                c.ncfn++;
                restore();
                c.ncf++;
                c.etamax = 1.0;
                if (c.ncf == BDF_MXNCF) {
                    finishStatus = CTB_STATUS_CONV_FAIL;
                    st = X_FINISH;
                    continue;
                }
                c.eta = BDF_ETACF;
                c.nflag = NFLAG_PREV_CONV_FAIL;
                rescale();
                st = X_ATTEMPT;
            } else if (st == X_NLS_OK) {
                c.dsm = c.acnrm * c.tq[2];
                if (c.dsm <= 1.0) {
                    st = X_STEP_OK;
                    continue;
                }
                c.nef++;
                c.netf++;
                c.nflag = NFLAG_PREV_ERR_FAIL;
                restore();
                if (c.nef == BDF_MXNEF) {
                    finishStatus = CTB_STATUS_ERR_TEST_FAIL;
                    st = X_FINISH;
                    continue;
                }
                c.etamax = 1.0;
                if (c.nef <= BDF_MXNEF1) {
                    c.eta = 1.0 / (rootK(BDF_BIAS2 * c.dsm, c.L) + BDF_ADDON);
                    c.eta = fmax(BDF_ETAMIN, c.eta);
                    if (c.nef >= BDF_SMALL_NEF) c.eta = fmin(c.eta, BDF_ETAMXF);
                    if (!(c.eta == c.eta)) c.eta = BDF_ETAMIN;
                    rescale();
                    st = X_ATTEMPT;
                    continue;
                }
                if (c.q > 1) {
                    c.eta = BDF_ETAMIN;
                    adjustOrder(-1);
                    c.L = c.q;
                    c.q--;
                    c.qwait = c.L;
                    rescale();
                    st = X_ATTEMPT;
                    continue;
                }
                // order 1 and still failing: restart the history from f(tn, zn[0])
                c.eta = BDF_ETAMIN;
                c.h *= c.eta;
                c.hscale = c.h;
                c.qwait = BDF_LONG_WAIT;
                CTB_FOR_OWN(i) Y(i) = ZN(0, i);
                c.nfe++;
                c.pc = PC_RESTART1;
                break;
            } else if (st == PC_RESTART1) {
                CTB_FOR_OWN(i) ZN(1, i) = c.h * F(i);
                st = X_ATTEMPT;
            } else if (st == X_STEP_OK) {
                completeStep();
                prepareNextStep(c.dsm);
                c.etamax = BDF_ETAMX2;
                const double t2 = c.tq[2];
                CTB_FOR_OWN(i) AC(i) *= t2;
                g.sync();
                const double Tnow = ZN(0, p.tIdx);
                if (!(Tnow == Tnow)) {
                    finishStatus = CTB_STATUS_NOT_FINITE;
                    st = X_FINISH;
                    continue;
                }
                if (c.ignPending && Tnow >= c.Tign) {
                    // first crossing of T0 + dT, linear in the step (openmp_ignition.cpp:70-73)
                    if (g.lane() == 0) p.tIgn[c.r] = c.tPrev + (c.tn - c.tPrev) * (c.Tign - c.TPrev) / (Tnow - c.TPrev);
                    c.ignPending = 0;
                }
                c.tPrev = c.tn;
                c.TPrev = Tnow;
                st = X_ADVANCE;
            } else { // X_FINISH
                finish(finishStatus);
                st = PC_NEW;
            }
        }
#ifdef CTB_CTL_IN_SMEM
        if (g.lane() == 0) mem.pc[s] = c.pc;
#else
        if (g.lane() == 0) {
            *ctlPtr = c;
            mem.pc[s] = c.pc;
        }
#endif
    }
};

} // namespace ctb
