// TEST INFRASTRUCTURE — runs the resumable BDF controller of cantera_b200/csrc/bdf_batch.cuh and
// the thread-per-reactor right-hand side of reactor_thread.cuh ON THE HOST, in the same
// control -> rhs -> setup rounds as the GPU integrator (reactor_batch.cu), with a few slots so that
// the queue refill is exercised. The Newton matrix is built here by finite differences of the host
// RHS (the GPU builds it analytically), which changes the Newton convergence rate but not the
// converged steps.
//   emul_batch <mech.ctbm> <in.bin> <out.bin> <slots>
// in.bin:  int64 N, int64 K, double tEnd, double dTign, T[N], P[N], Y[N*K]
// out.bin: int64 N, int64 K, y[N*(K+1)], tIgn[N], status[N] (as double), steps[N] (as double)
#include "../../cantera_b200/csrc/bdf_batch.cuh"
#include "../../cantera_b200/csrc/reactor_thread.cuh"
#include "../../include/cantera_b200/MechTables.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace ctb;

static void hostRhs(const PackedMech& pm, double P, const double* y, double* ydot)
{
    const int K = pm.v.K;
    std::vector<double> c(y + 1, y + 1 + K), g(K), w(K);
    TArr t;
    t.c = c.data(); t.g = g.data(); t.w = w.data(); t.st = 1;
    ydot[0] = reactorRhsT<4>(pm.v, pm.blob.data(), y[0], P, t);
    for (int k = 0; k < K; k++) ydot[1 + k] = w[k];
}

int main(int argc, char** argv)
{
    if (argc < 5) return 2;
    cantera_b200::MechTables mt = cantera_b200::MechTables::load(argv[1]);
    ctb_mech_tables view = mt.view();
    PackedMech pm = packMechanism(view);
    FILE* f = std::fopen(argv[2], "rb");
    if (!f) return 1;
    long long hdr[2];
    double th[2];
    if (std::fread(hdr, 8, 2, f) != 2 || std::fread(th, 8, 2, f) != 2) return 1;
    const long long N = hdr[0];
    const int K = (int) hdr[1], n = K + 1;
    if (K != pm.v.K) return 3;
    std::vector<double> T(N), P(N), Y(N * K);
    if (std::fread(T.data(), 8, N, f) != (size_t) N || std::fread(P.data(), 8, N, f) != (size_t) N
        || std::fread(Y.data(), 8, N * K, f) != (size_t) (N * K)) return 1;
    std::fclose(f);
    const int S = std::atoi(argv[4]);

    // reactor initial states (reactorInitKernel of reactor_kernels.cu)
    std::vector<double> y(N * n), T0(T), tIgn(N, NAN);
    for (long long r = 0; r < N; r++) {
        double sum = 0.0, sym = 0.0;
        for (int k = 0; k < K; k++) sum += std::fmax(Y[r * K + k], 0.0);
        for (int k = 0; k < K; k++) sym += (std::fmax(Y[r * K + k], 0.0) / sum) * (1.0 / mt.mol_weights[k]);
        const double mmw = 1.0 / sym;
        const double dens = P[r] * mmw / (T[r] * CTB_GAS_CONSTANT);
        y[r * n] = T[r];
        for (int k = 0; k < K; k++) y[r * n + 1 + k] = dens * (1.0 / mt.mol_weights[k]) * (std::fmax(Y[r * K + k], 0.0) / sum);
    }
    const long long histStride = 6LL * n + BDF_HIST_SCALARS;
    std::vector<double> hist(N * histStride);
    std::vector<int32_t> status(N, 0);
    std::vector<long long> stats(N * CTB_NSTATS, 0);
    unsigned long long counter = 0;
    ReactorBatch b = {};
    b.type = 0; b.tIdx = 0; b.sIdx = 1;
    b.N = N; b.neq = n; b.y = y.data(); b.P = P.data(); b.T0 = T0.data(); b.hist = hist.data();
    b.histStride = histStride; b.tIgn = tIgn.data(); b.status = status.data(); b.stats = stats.data();
    b.counter = &counter; b.t0 = 0.0; b.tout = th[0]; b.rtol = 1e-9; b.atol = 1e-15 * 0.1 /* reactor.cu kDefaultAtolSafety */; b.atolUser = 0;
    b.maxSteps = 20000; b.restart = 1; b.ignDT = th[1]; b.crateFloor = 0.0;

    SlotMem mem = {};
    const unsigned long long slotBytes = slotLayout(n, mem);
    std::vector<double> slotStore((size_t) slotBytes * S / 8 + 32, 0.0);
    std::vector<int> setupList(S), pcArr(S, PC_NEW);
    int setupCount[2] = {0, 0}, nActive = S;
    mem.slots = reinterpret_cast<unsigned char*>(slotStore.data());
    mem.pc = pcArr.data();
    mem.setupCount = setupCount; mem.setupList = setupList.data(); mem.nActive = &nActive; mem.S = S;
    mem.parity = 0;
    for (int s = 0; s < S; s++) mem.ctl(s)->pc = PC_NEW;

    const Solo solo;
    std::vector<double> xs(n), yy(n), f0(n), f1(n), J((size_t) n * n);
    long long rounds = 0;
    for (;; rounds++) {
        setupCount[0] = 0;
        for (int s = 0; s < S; s++) {
            if (pcArr[s] == PC_IDLE) continue;
            BatchBdf<Solo, 1> bdf(solo, b, mem, s, mem.block(s), mem.M(s), xs.data());
            bdf.run();
        }
        if (nActive == 0) break;
        for (int s = 0; s < S; s++) {
            if (pcArr[s] == PC_IDLE) continue;
            double* sv = mem.vec(s);
            hostRhs(pm, mem.ctl(s)->P, sv + 8 * n, sv + 9 * n);
        }
        // flagged slots: Jacobian by finite differences here (analytic on the GPU), then
        // (I - gamma J)^-1 by Gauss-Jordan elimination with partial pivoting (batchInvertKernel)
        for (int e = 0; e < setupCount[0]; e++) {
            const int s = setupList[e];
            const double Ps = mem.ctl(s)->P;
            for (int i = 0; i < n; i++) yy[i] = mem.vec(s)[8 * n + i];
            hostRhs(pm, Ps, yy.data(), f0.data());
            double ntot = 0.0;
            for (int i = 1; i < n; i++) ntot += yy[i];
            const double gamma = mem.ctl(s)->gamma;
            for (int j = 0; j < n; j++) {
                const double save = yy[j];
                const double d = (j == 0) ? save * 1.4901161193847656e-08
                                          : std::fmax(std::fabs(save) * 1e-7, 1e-11 * ntot);
                yy[j] = save + d;
                hostRhs(pm, Ps, yy.data(), f1.data());
                yy[j] = save;
                for (int i = 0; i < n; i++) J[(size_t) j * n + i] = (i == j ? 1.0 : 0.0) - gamma * (f1[i] - f0[i]) / d;
            }
            // in-place Gauss-Jordan without row exchanges: step k takes the largest |a_ik| among the
            // rows not used yet; row r_k then holds row k of the inverse, column r_j at position j
            int fail = 0;
            std::vector<int> piv(n), stepOf(n, -1);
            std::vector<char> used(n, 0);
            for (int k = 0; k < n && !fail; k++) {
                int pv = -1;
                double best = 0.0;
                for (int i = 0; i < n; i++) {
                    const double a = std::fabs(J[(size_t) k * n + i]);
                    if (!used[i] && a > best) { best = a; pv = i; }
                }
                if (pv < 0 || !std::isfinite(best)) { fail = 1; break; }
                used[pv] = 1; piv[k] = pv; stepOf[pv] = k;
                const double pinv = 1.0 / J[(size_t) k * n + pv];
                for (int j = 0; j < n; j++) J[(size_t) j * n + pv] = (j == k) ? pinv : J[(size_t) j * n + pv] * pinv;
                for (int i = 0; i < n; i++) {
                    if (i == pv) continue;
                    const double fct = J[(size_t) k * n + i];
                    for (int j = 0; j < n; j++) {
                        const double r = J[(size_t) j * n + pv];
                        J[(size_t) j * n + i] = (j == k) ? -fct * r : J[(size_t) j * n + i] - fct * r;
                    }
                }
            }
            mem.ctl(s)->sing = fail;
            if (!fail) {
                double* Minv = mem.M(s);
                for (int i = 0; i < n; i++)
                    for (int j = 0; j < n; j++) Minv[(size_t) piv[j] * n + stepOf[i]] = J[(size_t) j * n + i];
            }
        }
    }
    std::fprintf(stderr, "rounds %lld\n", rounds);
    f = std::fopen(argv[3], "wb");
    if (!f) return 1;
    std::fwrite(hdr, 8, 2, f);
    std::fwrite(y.data(), 8, N * n, f);
    std::fwrite(tIgn.data(), 8, N, f);
    std::vector<double> tmp(N);
    for (long long r = 0; r < N; r++) tmp[r] = status[r];
    std::fwrite(tmp.data(), 8, N, f);
    for (long long r = 0; r < N; r++) tmp[r] = (double) stats[r * CTB_NSTATS + STAT_STEPS];
    std::fwrite(tmp.data(), 8, N, f);
    std::fclose(f);
    return 0;
}
