// TEST / BASELINE INFRASTRUCTURE — drives the UNMODIFIED reference (Cantera 4.0.0a2 compiled by
// build_ref.sh into oracle/_ref/libcantera_ref.a) on the hot path, one Solution (+ reactor +
// ReactorNet) per worker PROCESS, exactly like samples/cxx/openmp_ignition/openmp_ignition.cpp:26-76
// does per thread. Used by bench.py's cpu_baseline / --impl reference / same-run parity legs and by
// the parity tests only.
//
//   ref_bench wdot   <yaml> <phase> <N> <seed> <procs> <datadir> [options]
//   ref_bench ddx    <yaml> <phase> <N> <seed> <procs> <datadir> [options]
//   ref_bench ignite <yaml> <phase> <N> <seed> <procs> <datadir> <fuel> <nO2> <O2> <N2> <Tlo> <Thi> <Plo> <Phi> <tEnd> [options]
//
// options:
//   --states FILE   read T[N], P[N], Y[N*K] from a CTBM0001 array file instead of generating them
//                   (N is taken from the file; the composition arguments of `ignite` are ignored)
//   --dump FILE     write inputs and reference outputs as a CTBM0001 array file:
//                   wdot: wdot[N*K]; ddx: ddX[N*K*K] dense column-major; ignite: Tend, Yend, tign, stats
//   --tign DT       ignite + dump: first crossing of T0 + DT by the openmp_ignition.cpp:70-73 step loop
//                   (a separate pass before the timed advance)
//   --reactor TYPE  IdealGasConstPressureMoleReactor (default) | IdealGasMoleReactor | IdealGasReactor |
//                   IdealGasConstPressureReactor
//   --rtol R --atol A   ReactorNet::setTolerances
//
// Prints one JSON line: {"mode":..., "n":N, "procs":P, "seconds":wall, "per_second":N/wall, "checksum":...}
// wall = max(end) - min(start) over the workers' timed loops (CLOCK_MONOTONIC; mechanism load,
// state generation and the --tign pass excluded).
#include "ctb_arrays.h"

#include "cantera/base/Solution.h"
#include "cantera/base/global.h"
#include "cantera/kinetics/Kinetics.h"
#include "cantera/thermo/ThermoPhase.h"
#include "cantera/zeroD/IdealGasConstPressureMoleReactor.h"
#include "cantera/zeroD/IdealGasConstPressureReactor.h"
#include "cantera/zeroD/IdealGasMoleReactor.h"
#include "cantera/zeroD/IdealGasReactor.h"
#include "cantera/zeroD/ReactorNet.h"
#include "cantera/numerics/eigen_sparse.h"

#include <sys/mman.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
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

template <class T>
static T* sharedArray(size_t n)
{
    if (n == 0) n = 1;
    void* p = mmap(nullptr, n * sizeof(T), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (p == MAP_FAILED) throw std::runtime_error("mmap failed");
    return static_cast<T*>(p);
}

static shared_ptr<ReactorBase> makeReactor(const std::string& type, shared_ptr<Solution> sol)
{
    if (type == "IdealGasConstPressureMoleReactor") return std::make_shared<IdealGasConstPressureMoleReactor>(sol, false);
    if (type == "IdealGasMoleReactor") return std::make_shared<IdealGasMoleReactor>(sol, false);
    if (type == "IdealGasReactor") return std::make_shared<IdealGasReactor>(sol, false);
    if (type == "IdealGasConstPressureReactor") return std::make_shared<IdealGasConstPressureReactor>(sol, false);
    throw std::runtime_error("unknown reactor type " + type);
}

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
    // positional ignite arguments, then options
    std::vector<std::string> pos;
    std::string statesFile, dumpFile, reactorType = "IdealGasConstPressureMoleReactor";
    double dTign = 0.0, rtol = -1.0, atol = -1.0;
    for (int i = 8; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> std::string {
            if (i + 1 >= argc) throw std::runtime_error("missing value after " + a);
            return argv[++i];
        };
        if (a == "--states") statesFile = next();
        else if (a == "--dump") dumpFile = next();
        else if (a == "--tign") dTign = std::stod(next());
        else if (a == "--reactor") reactorType = next();
        else if (a == "--rtol") rtol = std::stod(next());
        else if (a == "--atol") atol = std::stod(next());
        else pos.push_back(a);
    }
    size_t K = 0, R = 0;
    try {
        addDataDirectory(datadir);
        auto probe = newSolution(yaml, phase, "none");
        K = probe->thermo()->nSpecies();
        R = probe->kinetics()->nReactions();
    } catch (std::exception& ex) {
        std::cerr << ex.what() << "\n";
        return 1;
    }
    ArrayReader in;
    if (!statesFile.empty()) {
        in.load(statesFile);
        N = in.d("T").size();
        if (in.d("P").size() != N || in.d("Y").size() != N * K) {
            std::cerr << "ref_bench: --states arrays do not match the mechanism\n";
            return 2;
        }
    }
    if (mode == "ignite" && statesFile.empty() && pos.size() < 9) {
        std::cerr << "ignite needs fuel nO2 O2 N2 Tlo Thi Plo Phi tEnd\n";
        return 2;
    }
    if (mode == "ignite" && pos.size() < 9) pos.resize(9, "0");
    const bool dump = !dumpFile.empty();
    Shared* sh = sharedArray<Shared>(1);
    new (sh) Shared();
    sh->ready = 0; sh->go = 0;
    double* T = sharedArray<double>(N);
    double* P = sharedArray<double>(N);
    double* Y = sharedArray<double>(N * K);
    if (!statesFile.empty()) {
        std::memcpy(T, in.d("T").data(), N * 8);
        std::memcpy(P, in.d("P").data(), N * 8);
        std::memcpy(Y, in.d("Y").data(), N * K * 8);
    }
    const size_t outPer = !dump ? 0 : mode == "wdot" ? K : mode == "ddx" ? K * K : K;
    double* out = sharedArray<double>(N * outPer);
    double* Tend = sharedArray<double>(dump ? N : 0);
    double* tign = sharedArray<double>(dump ? N : 0);
    double* stats = sharedArray<double>(dump ? N * 8 : 0);
    const double tEnd = mode == "ignite" ? std::stod(pos[8]) : 0.0;

    for (int p = 0; p < procs; p++) {
        pid_t pid = fork();
        if (pid != 0) continue;
        // ---- worker p: contiguous slice [b, e)
        int rc = 0;
        try {
            auto sol = newSolution(yaml, phase, "none");
            auto gas = sol->thermo();
            auto kin = sol->kinetics();
            const size_t b = N * p / procs, e = N * (p + 1) / procs;
            std::mt19937_64 rng(seed + 7919 * p);
            std::uniform_real_distribution<double> U(0.0, 1.0);
            std::vector<double> w(K);
            double acc = 0.0;
            if (mode == "wdot" || mode == "ddx") {
                if (statesFile.empty()) {
                    for (size_t i = b; i < e; i++) {
                        T[i] = 800.0 + 1700.0 * U(rng);
                        P[i] = OneAtm * (1.0 + 49.0 * U(rng));
                        double s = 0;
                        for (size_t k = 0; k < K; k++) { Y[i * K + k] = U(rng); s += Y[i * K + k]; }
                        for (size_t k = 0; k < K; k++) Y[i * K + k] /= s;
                    }
                }
                sh->ready++;
                while (!sh->go.load()) usleep(100);
                sh->t0[p] = now();
                for (size_t i = b; i < e; i++) {
                    gas->setState_TPY(T[i], P[i], span<const double>(&Y[i * K], K));
                    if (mode == "wdot") {
                        kin->getNetProductionRates(w);
                        acc += w[K / 2];
                        if (dump) std::memcpy(out + i * K, w.data(), K * 8);
                    } else {
                        auto J = kin->netProductionRates_ddX();
                        acc += J.coeff(K / 2, K / 2);
                        if (dump) {
                            double* o = out + i * K * K;
                            std::fill(o, o + K * K, 0.0);
                            for (int c = 0; c < J.outerSize(); c++) {
                                for (Eigen::SparseMatrix<double>::InnerIterator it(J, c); it; ++it) {
                                    o[(size_t) it.col() * K + it.row()] = it.value();
                                }
                            }
                        }
                    }
                }
                sh->t1[p] = now();
            } else if (mode == "ignite") {
                auto reactor = makeReactor(reactorType, sol);
                ReactorNet net(reactor);
                if (rtol > 0 || atol > 0) net.setTolerances(rtol, atol);
                if (statesFile.empty()) {
                    std::string fuel = pos[0], o2 = pos[2], n2 = pos[3];
                    double nO2 = std::stod(pos[1]), Tlo = std::stod(pos[4]), Thi = std::stod(pos[5]);
                    double Plo = std::stod(pos[6]), Phi = std::stod(pos[7]);
                    size_t kf = gas->speciesIndex(fuel), ko = gas->speciesIndex(o2), kn = gas->speciesIndex(n2);
                    std::vector<double> X(K, 0.0);
                    for (size_t i = b; i < e; i++) {
                        T[i] = Tlo + (Thi - Tlo) * U(rng);
                        P[i] = OneAtm * (Plo + (Phi - Plo) * U(rng));
                        double phi = 0.5 + 1.5 * U(rng);
                        std::fill(X.begin(), X.end(), 0.0);
                        X[kf] = phi; X[ko] = nO2; X[kn] = 3.76 * nO2;
                        gas->setState_TPX(T[i], P[i], X);
                        gas->getMassFractions(span<double>(&Y[i * K], K));
                    }
                }
                if (dump && dTign > 0) {
                    // untimed pass: ignition delay by the step loop of openmp_ignition.cpp:70-73, crossing
                    // time by linear interpolation between internal steps
                    for (size_t i = b; i < e; i++) {
                        gas->setState_TPY(T[i], P[i], span<const double>(&Y[i * K], K));
                        net.setInitialTime(0.0);
                        tign[i] = std::nan("");
                        double tPrev = 0.0, TPrev = T[i];
                        while (net.time() < tEnd && std::isnan(tign[i])) {
                            double t = net.step();
                            double Tn = reactor->temperature();
                            if (Tn >= T[i] + dTign) tign[i] = tPrev + (t - tPrev) * (T[i] + dTign - TPrev) / (Tn - TPrev);
                            tPrev = t; TPrev = Tn;
                        }
                    }
                } else if (dump) {
                    for (size_t i = b; i < e; i++) tign[i] = std::nan("");
                }
                sh->ready++;
                while (!sh->go.load()) usleep(100);
                sh->t0[p] = now();
                for (size_t i = b; i < e; i++) {
                    gas->setState_TPY(T[i], P[i], span<const double>(&Y[i * K], K));
                    net.setInitialTime(0.0);
                    net.advance(tEnd);
                    acc += gas->temperature();
                    if (dump) {
                        Tend[i] = gas->temperature();
                        gas->getMassFractions(span<double>(out + i * K, K));
                        AnyMap st = net.solverStats();
                        const char* keys[8] = {"steps", "rhs_evals", "jac_evals", "lin_solve_setups",
                                               "nonlinear_iters", "nonlinear_conv_fails", "err_test_fails", "last_order"};
                        for (int j = 0; j < 8; j++) stats[i * 8 + j] = st.hasKey(keys[j]) ? (double) st[keys[j]].asInt() : -1.0;
                    }
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
    if (dump) {
        ArrayFile f;
        f.add("dims", std::vector<int32_t>{(int32_t) N, (int32_t) K, (int32_t) R});
        f.add("T", std::vector<double>(T, T + N));
        f.add("P", std::vector<double>(P, P + N));
        f.add("Y", std::vector<double>(Y, Y + N * K));
        if (mode == "wdot") f.add("wdot", std::vector<double>(out, out + N * K));
        if (mode == "ddx") f.add("ddX", std::vector<double>(out, out + N * K * K));
        if (mode == "ignite") {
            f.add("tEnd", std::vector<double>{tEnd});
            f.add("dTign", std::vector<double>{dTign});
            f.add("Tend", std::vector<double>(Tend, Tend + N));
            f.add("Yend", std::vector<double>(out, out + N * K));
            f.add("tign", std::vector<double>(tign, tign + N));
            f.add("stats", std::vector<double>(stats, stats + N * 8));
        }
        f.save(dumpFile);
    }
    std::printf("{\"mode\": \"%s\", \"n\": %zu, \"procs\": %d, \"seconds\": %.6f, \"per_second\": %.6g, \"checksum\": %.17g}\n",
                mode.c_str(), N, procs, t1 - t0, N / (t1 - t0), sum);
    return 0;
}
