// TEST / BASELINE INFRASTRUCTURE — times the UNMODIFIED reference (Cantera 4.0.0a2 compiled by
// build_ref.sh into oracle/_ref/libcantera_ref.a) on the hot path, one Solution (+ reactor +
// ReactorNet) per worker PROCESS, exactly like samples/cxx/openmp_ignition/openmp_ignition.cpp:26-76
// does per thread. Used by bench.py's cpu_baseline and --impl reference legs only.
//
//   ref_bench wdot   <yaml> <phase> <N> <seed> <procs> <datadir>
//   ref_bench ddx    <yaml> <phase> <N> <seed> <procs> <datadir>
//   ref_bench ignite <yaml> <phase> <N> <seed> <procs> <datadir> <fuel> <nO2> <O2> <N2> <Tlo> <Thi> <Plo> <Phi> <tEnd>
//
// Prints one JSON line: {"mode":..., "n":N, "procs":P, "seconds":wall, "per_second":N/wall, "checksum":...}
// wall = max(end) - min(start) over the workers' loops (CLOCK_MONOTONIC; mechanism load excluded).
#include "cantera/base/Solution.h"
#include "cantera/base/global.h"
#include "cantera/kinetics/Kinetics.h"
#include "cantera/thermo/ThermoPhase.h"
#include "cantera/zeroD/IdealGasConstPressureMoleReactor.h"
#include "cantera/zeroD/ReactorNet.h"
#include "cantera/numerics/eigen_sparse.h"

#include <sys/mman.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <cmath>
#include <cstdio>
#include <iostream>
#include <random>

using namespace Cantera;

static double now()
{
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

struct Shared {
    std::atomic<int> ready;
    std::atomic<int> go;
    double t0[256], t1[256], sum[256];
    int fail[256];
};

int main(int argc, char** argv)
{
    if (argc < 8) {
        std::cerr << "usage: ref_bench wdot|ddx|ignite yaml phase N seed procs datadir [...]\n";
        return 2;
    }
    std::string mode = argv[1], yaml = argv[2], phase = argv[3];
    if (phase == "-") phase = "";
    size_t N = std::stoul(argv[4]);
    uint64_t seed = std::stoull(argv[5]);
    int procs = std::max(1, std::min(256, std::atoi(argv[6])));
    std::string datadir = argv[7];
    Shared* sh = (Shared*) mmap(nullptr, sizeof(Shared), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    new (sh) Shared();
    sh->ready = 0; sh->go = 0;
    for (int p = 0; p < procs; p++) {
        pid_t pid = fork();
        if (pid != 0) continue;
        // ---- worker p: contiguous slice [b, e)
        int rc = 0;
        try {
            addDataDirectory(datadir);
            auto sol = newSolution(yaml, phase, "none");
            auto gas = sol->thermo();
            auto kin = sol->kinetics();
            size_t K = gas->nSpecies();
            size_t b = N * p / procs, e = N * (p + 1) / procs, n = e - b;
            std::mt19937_64 rng(seed + 7919 * p);
            std::uniform_real_distribution<double> U(0.0, 1.0);
            std::vector<double> T(n), P(n), Y(n * K, 0.0), out(K);
            double acc = 0.0;
            if (mode == "wdot" || mode == "ddx") {
                for (size_t i = 0; i < n; i++) {
                    T[i] = 800.0 + 1700.0 * U(rng);
                    P[i] = OneAtm * (1.0 + 49.0 * U(rng));
                    double s = 0;
                    for (size_t k = 0; k < K; k++) { Y[i * K + k] = U(rng); s += Y[i * K + k]; }
                    for (size_t k = 0; k < K; k++) Y[i * K + k] /= s;
                }
                sh->ready++;
                while (!sh->go.load()) usleep(100);
                sh->t0[p] = now();
                for (size_t i = 0; i < n; i++) {
                    gas->setState_TPY(T[i], P[i], span<const double>(&Y[i * K], K));
                    if (mode == "wdot") {
                        kin->getNetProductionRates(out);
                        acc += out[K / 2];
                    } else {
                        auto J = kin->netProductionRates_ddX();
                        acc += J.coeff(K / 2, K / 2);
                    }
                }
                sh->t1[p] = now();
            } else if (mode == "ignite") {
                if (argc < 17) throw std::runtime_error("ignite needs fuel nO2 O2 N2 Tlo Thi Plo Phi tEnd");
                std::string fuel = argv[8], o2 = argv[10], n2 = argv[11];
                double nO2 = std::stod(argv[9]), Tlo = std::stod(argv[12]), Thi = std::stod(argv[13]);
                double Plo = std::stod(argv[14]), Phi = std::stod(argv[15]), tEnd = std::stod(argv[16]);
                auto reactor = std::make_shared<IdealGasConstPressureMoleReactor>(sol, false);
                ReactorNet net(reactor);
                size_t kf = gas->speciesIndex(fuel), ko = gas->speciesIndex(o2), kn = gas->speciesIndex(n2);
                std::vector<double> X(K, 0.0);
                for (size_t i = 0; i < n; i++) {
                    T[i] = Tlo + (Thi - Tlo) * U(rng);
                    P[i] = OneAtm * (Plo + (Phi - Plo) * U(rng));
                    double phi = 0.5 + 1.5 * U(rng);
                    std::fill(X.begin(), X.end(), 0.0);
                    X[kf] = phi; X[ko] = nO2; X[kn] = 3.76 * nO2;
                    gas->setState_TPX(T[i], P[i], X);
                    gas->getMassFractions(span<double>(&Y[i * K], K));
                }
                sh->ready++;
                while (!sh->go.load()) usleep(100);
                sh->t0[p] = now();
                for (size_t i = 0; i < n; i++) {
                    gas->setState_TPY(T[i], P[i], span<const double>(&Y[i * K], K));
                    net.setInitialTime(0.0);
                    net.advance(tEnd);
                    acc += gas->temperature();
                }
                sh->t1[p] = now();
            } else {
                throw std::runtime_error("unknown mode " + mode);
            }
            sh->sum[p] = acc;
        } catch (std::exception& ex) {
            std::cerr << "worker " << p << ": " << ex.what() << "\n";
            sh->ready++;
            rc = 1;
        }
        sh->fail[p] = rc;
        _exit(rc);
    }
    while (sh->ready.load() < procs) usleep(1000);
    sh->go = 1;
    int bad = 0;
    for (int p = 0; p < procs; p++) {
        int st = 0;
        wait(&st);
        if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) bad++;
    }
    if (bad) {
        std::printf("{\"mode\": \"%s\", \"error\": \"%d worker(s) failed\"}\n", mode.c_str(), bad);
        return 1;
    }
    double t0 = 1e300, t1 = 0, sum = 0;
    for (int p = 0; p < procs; p++) {
        t0 = std::min(t0, sh->t0[p]);
        t1 = std::max(t1, sh->t1[p]);
        sum += sh->sum[p];
    }
    std::printf("{\"mode\": \"%s\", \"n\": %zu, \"procs\": %d, \"seconds\": %.6f, \"per_second\": %.6g, \"checksum\": %.17g}\n",
                mode.c_str(), N, procs, t1 - t0, N / (t1 - t0), sum);
    return 0;
}
