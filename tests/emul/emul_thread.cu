// TEST INFRASTRUCTURE — runs the thread-per-state device code of cantera_b200/csrc/rates_thread.cuh
// ON THE HOST (the per-state arithmetic is __host__ __device__), so that the packer's run table and
// the fast/general reaction paths can be checked against the oracle without a GPU.
//   emul_thread <mech.ctbm> <nStates> <seed> <out.bin> [U]
// out.bin: int64 n, int64 K, then T[n], P[n], Y[n*K], wdot[n*K] (float64).
#include "../../cantera_b200/csrc/rates_thread.cuh"
#include "../../include/cantera_b200/MechTables.h"

#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

using namespace ctb;

template <int U>
static void evalState(const PackedMech& pm, double T, double P, const double* Y, double* wdot, bool neg)
{
    const MechView& v = pm.v;
    const unsigned char* sm = pm.blob.data();
    const int K = v.K;
    std::vector<double> c(Y, Y + K), g(K), w(K);
    TArr t;
    t.c = c.data(); t.g = g.data(); t.w = w.data(); t.st = 1;
    GasScalars s;
    double dens;
    stateFromTPY_T(v, sm, T, P, t, s, dens);
    thermoT(v, sm, s, t, dens);
    if (neg) reactionsT<true, U>(v, sm, s, t); else reactionsT<false, U>(v, sm, s, t);
    for (int k = 0; k < K; k++) wdot[k] = w[k];
}

int main(int argc, char** argv)
{
    if (argc < 5) { std::fprintf(stderr, "usage: emul_thread mech.ctbm n seed out.bin [U]\n"); return 2; }
    cantera_b200::MechTables mt = cantera_b200::MechTables::load(argv[1]);
    const long long n = std::atoll(argv[2]);
    const unsigned long long seed = std::strtoull(argv[3], nullptr, 10);
    const int U = argc > 5 ? std::atoi(argv[5]) : 4;
    ctb_mech_tables view = mt.view();
    PackedMech pm = packMechanism(view);
    const int K = pm.v.K;
    int fast = 0;
    const RunRec* runs = reinterpret_cast<const RunRec*>(pm.blob.data() + pm.v.o_runs);
    for (int i = 0; i < pm.v.nRuns; i++) if (runs[i].code >= 0) fast += runs[i].end - runs[i].begin;
    std::fprintf(stderr, "K=%d R=%d runs=%d fast=%d hotBytes=%u\n", K, pm.v.R, pm.v.nRuns, fast, pm.v.hotBytes);
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<double> u01(0.0, 1.0);
    std::vector<double> T(n), P(n), Y(n * K), W(n * K);
    for (long long i = 0; i < n; i++) {
        T[i] = 300.0 + 2700.0 * u01(rng);
        P[i] = 101325.0 * (0.01 + 50.0 * u01(rng));
        double sum = 0.0;
        for (int k = 0; k < K; k++) { Y[i * K + k] = u01(rng) < 0.15 ? 0.0 : u01(rng); sum += Y[i * K + k]; }
        for (int k = 0; k < K; k++) Y[i * K + k] /= sum;
        if (U == 2) evalState<2>(pm, T[i], P[i], &Y[i * K], &W[i * K], i & 1);
        else evalState<4>(pm, T[i], P[i], &Y[i * K], &W[i * K], i & 1);
    }
    FILE* f = std::fopen(argv[4], "wb");
    if (!f) return 1;
    long long hdr[2] = {n, K};
    std::fwrite(hdr, 8, 2, f);
    std::fwrite(T.data(), 8, n, f); std::fwrite(P.data(), 8, n, f);
    std::fwrite(Y.data(), 8, n * K, f); std::fwrite(W.data(), 8, n * K, f);
    std::fclose(f);
    return 0;
}
