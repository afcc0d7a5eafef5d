// TEST INFRASTRUCTURE / reference-side tool. Drives the UNMODIFIED reference (Cantera 4.0.0a2
// built by build_ref.sh) on seeded synthetic states and dumps inputs + reference outputs as a
// named-array file (ctb_arrays.h) for tests/golden/.
//
//   gen_golden rates  <yaml> <phase> <out> <N> <seed>
//   gen_golden ddx    <yaml> <phase> <out> <N> <seed>
//   gen_golden rhs    <yaml> <phase> <out> <N> <seed> [reactorType]
//   gen_golden rjac   <yaml> <phase> <out> <N> <seed>          reactor Jacobian (getJacobianElements), dense
//   gen_golden ddxdigest <yaml> <phase> <out> <N> <seed>       J r, l^T J, |J||r|, |l|^T|J| of netProductionRates_ddX
//   gen_golden derivs <yaml> <phase> <out> <N> <seed>          _ddT/_ddP/_ddC families and getDelta*
//   gen_golden ignite <yaml> <phase> <out> <N> <seed> <fuel> <nO2> <O2> <N2> <Tlo> <Thi> <Plo_atm> <Phi_atm> <tEnd> [rtol atol]
//
// States follow SURVEY.md §8(d): T ~ U(800, 2500) K, P = 101325 * U(1, 50) Pa, Y_k = u_k / sum u.
#include "ctb_arrays.h"

#include "cantera/base/Solution.h"
#include "cantera/base/global.h"
#include "cantera/kinetics/Kinetics.h"
#include "cantera/thermo/ThermoPhase.h"
#include "cantera/transport/Transport.h"
#include "cantera/zeroD/IdealGasConstPressureMoleReactor.h"
#include "cantera/zeroD/IdealGasConstPressureReactor.h"
#include "cantera/zeroD/IdealGasMoleReactor.h"
#include "cantera/zeroD/IdealGasReactor.h"
#include "cantera/zeroD/ReactorNet.h"
#include "cantera/numerics/eigen_sparse.h"

#include <cmath>
#include <functional>
#include <iostream>
#include <random>

using namespace Cantera;

struct States {
    std::vector<double> T, P, Y;
};

static States randomStates(size_t N, size_t K, uint64_t seed)
{
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    States s;
    s.T.resize(N); s.P.resize(N); s.Y.resize(N * K);
    for (size_t n = 0; n < N; n++) {
        s.T[n] = 800.0 + 1700.0 * U(rng);
        s.P[n] = OneAtm * (1.0 + 49.0 * U(rng));
        double sum = 0;
        for (size_t k = 0; k < K; k++) {
            s.Y[n * K + k] = U(rng);
            sum += s.Y[n * K + k];
        }
        for (size_t k = 0; k < K; k++) {
            s.Y[n * K + k] /= sum;
        }
    }
    // edge cases the reference defines behaviour for: a negative mass fraction is clipped
    // (Phase.cpp:342-360), exact zeros, an un-normalised vector, T exactly at a NASA Tmid.
    if (N >= 8) {
        s.Y[0 * K + 0] = -1e-3;
        for (size_t k = 1; k < K; k += 2) s.Y[1 * K + k] = 0.0;
        for (size_t k = 0; k < K; k++) s.Y[2 * K + k] *= 3.0;
        s.T[3] = 1000.0;
        for (size_t k = 0; k < K; k++) s.Y[4 * K + k] = (k == K - 1) ? 1.0 : 0.0;
    }
    return s;
}

static void densify(const Eigen::SparseMatrix<double>& J, double* out, size_t K)
{
    for (size_t i = 0; i < K * K; i++) out[i] = 0.0;
    for (int c = 0; c < J.outerSize(); c++) {
        for (Eigen::SparseMatrix<double>::InnerIterator it(J, c); it; ++it) {
            out[(size_t) it.col() * K + it.row()] = it.value();
        }
    }
}

static int runRates(shared_ptr<Solution> sol, const std::string& out, size_t N, uint64_t seed)
{
    auto gas = sol->thermo();
    auto kin = sol->kinetics();
    size_t K = gas->nSpecies(), R = kin->nReactions();
    States s = randomStates(N, K, seed);
    std::vector<double> kf(N * R), kc(N * R), ropf(N * R), ropr(N * R), ropn(N * R);
    std::vector<double> wdot(N * K), cdot(N * K), ddot(N * K);
    for (size_t n = 0; n < N; n++) {
        gas->setState_TPY(s.T[n], s.P[n], span<const double>(&s.Y[n * K], K));
        kin->getFwdRateConstants(span<double>(&kf[n * R], R));
        kin->getEquilibriumConstants(span<double>(&kc[n * R], R));
        kin->getFwdRatesOfProgress(span<double>(&ropf[n * R], R));
        kin->getRevRatesOfProgress(span<double>(&ropr[n * R], R));
        kin->getNetRatesOfProgress(span<double>(&ropn[n * R], R));
        kin->getNetProductionRates(span<double>(&wdot[n * K], K));
        kin->getCreationRates(span<double>(&cdot[n * K], K));
        kin->getDestructionRates(span<double>(&ddot[n * K], K));
    }
    ArrayFile f;
    f.add("dims", std::vector<int32_t>{(int32_t) N, (int32_t) K, (int32_t) R});
    f.add("T", s.T); f.add("P", s.P); f.add("Y", s.Y);
    f.add("kf", kf); f.add("kc", kc); f.add("ropf", ropf); f.add("ropr", ropr); f.add("ropnet", ropn);
    f.add("wdot", wdot); f.add("cdot", cdot); f.add("ddot", ddot);
    f.save(out);
    return 0;
}

static int runDdx(shared_ptr<Solution> sol, const std::string& out, size_t N, uint64_t seed)
{
    auto gas = sol->thermo();
    auto kin = sol->kinetics();
    size_t K = gas->nSpecies();
    States s = randomStates(N, K, seed);
    // variants: default, skip-falloff, skip-third-bodies, rtol-delta 1e-6, and ddCi default
    std::vector<double> jdef(N * K * K), jsf(N * K * K), jstb(N * K * K), jdel(N * K * K), jci(N * K * K);
    for (size_t n = 0; n < N; n++) {
        gas->setState_TPY(s.T[n], s.P[n], span<const double>(&s.Y[n * K], K));
        AnyMap st;
        kin->setDerivativeSettings(st); // empty map = reset to defaults
        densify(kin->netProductionRates_ddX(), &jdef[n * K * K], K);
        densify(kin->netProductionRates_ddCi(), &jci[n * K * K], K);
        st["skip-falloff"] = true;
        kin->setDerivativeSettings(st);
        densify(kin->netProductionRates_ddX(), &jsf[n * K * K], K);
        AnyMap st2;
        kin->setDerivativeSettings(st2);
        st2["skip-third-bodies"] = true;
        kin->setDerivativeSettings(st2);
        densify(kin->netProductionRates_ddX(), &jstb[n * K * K], K);
        AnyMap st3;
        kin->setDerivativeSettings(st3);
        st3["rtol-delta"] = 1e-6;
        kin->setDerivativeSettings(st3);
        densify(kin->netProductionRates_ddX(), &jdel[n * K * K], K);
        AnyMap st4;
        kin->setDerivativeSettings(st4);
    }
    ArrayFile f;
    f.add("dims", std::vector<int32_t>{(int32_t) N, (int32_t) K, (int32_t) kin->nReactions()});
    f.add("T", s.T); f.add("P", s.P); f.add("Y", s.Y);
    f.add("ddX", jdef); f.add("ddX_skip_falloff", jsf); f.add("ddX_skip_third_bodies", jstb);
    f.add("ddX_delta1e-6", jdel); f.add("ddCi", jci);
    f.save(out);
    return 0;
}

static shared_ptr<ReactorBase> makeReactor(const std::string& type, shared_ptr<Solution> sol)
{
    if (type == "IdealGasConstPressureMoleReactor") return std::make_shared<IdealGasConstPressureMoleReactor>(sol, false);
    if (type == "IdealGasMoleReactor") return std::make_shared<IdealGasMoleReactor>(sol, false);
    if (type == "IdealGasReactor") return std::make_shared<IdealGasReactor>(sol, false);
    if (type == "IdealGasConstPressureReactor") return std::make_shared<IdealGasConstPressureReactor>(sol, false);
    throw std::runtime_error("unknown reactor type " + type);
}

static int runRhs(shared_ptr<Solution> sol, const std::string& out, size_t N, uint64_t seed,
                  const std::string& type = "IdealGasConstPressureMoleReactor")
{
    auto gas = sol->thermo();
    size_t K = gas->nSpecies();
    States s = randomStates(N, K, seed);
    auto reactor = makeReactor(type, sol);
    ReactorNet net(reactor);
    net.initialize();
    const size_t neq = net.neq();
    std::vector<double> y(N * neq), ydot(N * neq);
    for (size_t n = 0; n < N; n++) {
        for (size_t k = 0; k < K; k++) s.Y[n * K + k] = std::abs(s.Y[n * K + k]);
        gas->setState_TPY(s.T[n], s.P[n], span<const double>(&s.Y[n * K], K));
        net.setInitialTime(0.0);
        net.initialize();
        net.getState(span<double>(&y[n * neq], neq));
        net.eval(0.0, span<double>(&y[n * neq], neq), span<double>(&ydot[n * neq], neq),
                 span<double>());
    }
    ArrayFile f;
    f.add("dims", std::vector<int32_t>{(int32_t) N, (int32_t) K, (int32_t) sol->kinetics()->nReactions(), (int32_t) neq});
    f.add("T", s.T); f.add("P", s.P); f.add("Y", s.Y);
    f.add("y", y); f.add("ydot", ydot);
    f.addText("reactor", type);
    f.save(out);
    return 0;
}

// IdealGasConstPressureMoleReactor::getJacobianElements (IdealGasConstPressureMoleReactor.cpp:108-174)
// through ReactorBase::jacobian(), dense column-major (K+1)^2 per state; variants: default derivative
// settings and skip-falloff (purely analytic species block).
static int runRjac(shared_ptr<Solution> sol, const std::string& out, size_t N, uint64_t seed)
{
    auto gas = sol->thermo();
    size_t K = gas->nSpecies(), n1 = K + 1;
    States s = randomStates(N, K, seed);
    auto reactor = std::make_shared<IdealGasConstPressureMoleReactor>(sol, false);
    ReactorNet net(reactor);
    std::vector<double> y(N * n1), jdef(N * n1 * n1), jsf(N * n1 * n1);
    for (size_t n = 0; n < N; n++) {
        for (size_t k = 0; k < K; k++) s.Y[n * K + k] = std::abs(s.Y[n * K + k]);
        gas->setState_TPY(s.T[n], s.P[n], span<const double>(&s.Y[n * K], K));
        net.setInitialTime(0.0);
        net.initialize();
        net.getState(span<double>(&y[n * n1], n1));
        AnyMap st;
        reactor->setDerivativeSettings(st);
        densify(reactor->jacobian(), &jdef[n * n1 * n1], n1);
        AnyMap st2;
        st2["skip-falloff"] = true;
        reactor->setDerivativeSettings(st2);
        densify(reactor->jacobian(), &jsf[n * n1 * n1], n1);
        AnyMap st3;
        st3["skip-falloff"] = false;
        reactor->setDerivativeSettings(st3);
    }
    ArrayFile f;
    f.add("dims", std::vector<int32_t>{(int32_t) N, (int32_t) K, (int32_t) sol->kinetics()->nReactions()});
    f.add("T", s.T); f.add("P", s.P); f.add("Y", s.Y);
    f.add("y", y); f.add("jac", jdef); f.add("jac_skip_falloff", jsf);
    f.save(out);
    return 0;
}

// Digest of netProductionRates_ddX for many states: J r and l^T J for fixed seeded vectors, plus
// the magnitude sums |J||r| and |l|^T|J| that scale the comparison (a single wrong element (k, j)
// of size delta moves (J r)_k by delta r_j and (l^T J)_j by delta l_k).
static int runDdxDigest(shared_ptr<Solution> sol, const std::string& out, size_t N, uint64_t seed)
{
    auto gas = sol->thermo();
    auto kin = sol->kinetics();
    size_t K = gas->nSpecies();
    States s = randomStates(N, K, seed);
    std::mt19937_64 rng(seed ^ 0x9e3779b97f4a7c15ULL);
    std::uniform_real_distribution<double> U(0.5, 1.5);
    std::vector<double> r(K), l(K);
    for (size_t k = 0; k < K; k++) { r[k] = U(rng); l[k] = U(rng); }
    const char* names[3] = {"skip_falloff", "skip_third_bodies", "default"};
    ArrayFile f;
    f.add("dims", std::vector<int32_t>{(int32_t) N, (int32_t) K, (int32_t) kin->nReactions()});
    f.add("T", s.T); f.add("P", s.P); f.add("Y", s.Y); f.add("r", r); f.add("l", l);
    for (int v = 0; v < 3; v++) {
        AnyMap st;
        kin->setDerivativeSettings(st);
        if (v == 0) st["skip-falloff"] = true;
        if (v == 1) st["skip-third-bodies"] = true;
        if (v < 2) kin->setDerivativeSettings(st);
        std::vector<double> Jr(N * K, 0.0), lJ(N * K, 0.0), aJr(N * K, 0.0), alJ(N * K, 0.0);
        for (size_t n = 0; n < N; n++) {
            gas->setState_TPY(s.T[n], s.P[n], span<const double>(&s.Y[n * K], K));
            auto J = kin->netProductionRates_ddX();
            for (int c = 0; c < J.outerSize(); c++) {
                for (Eigen::SparseMatrix<double>::InnerIterator it(J, c); it; ++it) {
                    Jr[n * K + it.row()] += it.value() * r[it.col()];
                    aJr[n * K + it.row()] += std::abs(it.value()) * r[it.col()];
                    lJ[n * K + it.col()] += it.value() * l[it.row()];
                    alJ[n * K + it.col()] += std::abs(it.value()) * l[it.row()];
                }
            }
        }
        f.add(std::string("Jr_") + names[v], Jr); f.add(std::string("lJ_") + names[v], lJ);
        f.add(std::string("aJr_") + names[v], aJr); f.add(std::string("alJ_") + names[v], alJ);
    }
    AnyMap reset;
    kin->setDerivativeSettings(reset);
    f.save(out);
    return 0;
}

// The derivative families of Kinetics.h:1040-1330 (BulkKinetics.cpp:284-385, Kinetics.cpp:459-602)
// and the reaction deltas (BulkKinetics.cpp:200-257).
static int runDerivs(shared_ptr<Solution> sol, const std::string& out, size_t N, uint64_t seed)
{
    auto gas = sol->thermo();
    auto kin = sol->kinetics();
    size_t K = gas->nSpecies(), R = kin->nReactions();
    States s = randomStates(N, K, seed);
    ArrayFile f;
    f.add("dims", std::vector<int32_t>{(int32_t) N, (int32_t) K, (int32_t) R});
    f.add("T", s.T); f.add("P", s.P); f.add("Y", s.Y);
    struct Get { const char* name; bool species; std::function<void(Kinetics&, span<double>)> fn; };
    std::vector<Get> gets = {
        {"kf_ddT", false, [](Kinetics& k, span<double> o) { k.getFwdRateConstants_ddT(o); }},
        {"kf_ddP", false, [](Kinetics& k, span<double> o) { k.getFwdRateConstants_ddP(o); }},
        {"kf_ddC", false, [](Kinetics& k, span<double> o) { k.getFwdRateConstants_ddC(o); }},
        {"ropf_ddT", false, [](Kinetics& k, span<double> o) { k.getFwdRatesOfProgress_ddT(o); }},
        {"ropf_ddP", false, [](Kinetics& k, span<double> o) { k.getFwdRatesOfProgress_ddP(o); }},
        {"ropf_ddC", false, [](Kinetics& k, span<double> o) { k.getFwdRatesOfProgress_ddC(o); }},
        {"ropr_ddT", false, [](Kinetics& k, span<double> o) { k.getRevRatesOfProgress_ddT(o); }},
        {"ropr_ddP", false, [](Kinetics& k, span<double> o) { k.getRevRatesOfProgress_ddP(o); }},
        {"ropr_ddC", false, [](Kinetics& k, span<double> o) { k.getRevRatesOfProgress_ddC(o); }},
        {"ropnet_ddT", false, [](Kinetics& k, span<double> o) { k.getNetRatesOfProgress_ddT(o); }},
        {"ropnet_ddP", false, [](Kinetics& k, span<double> o) { k.getNetRatesOfProgress_ddP(o); }},
        {"ropnet_ddC", false, [](Kinetics& k, span<double> o) { k.getNetRatesOfProgress_ddC(o); }},
        {"cdot_ddT", true, [](Kinetics& k, span<double> o) { k.getCreationRates_ddT(o); }},
        {"cdot_ddP", true, [](Kinetics& k, span<double> o) { k.getCreationRates_ddP(o); }},
        {"cdot_ddC", true, [](Kinetics& k, span<double> o) { k.getCreationRates_ddC(o); }},
        {"ddot_ddT", true, [](Kinetics& k, span<double> o) { k.getDestructionRates_ddT(o); }},
        {"ddot_ddP", true, [](Kinetics& k, span<double> o) { k.getDestructionRates_ddP(o); }},
        {"ddot_ddC", true, [](Kinetics& k, span<double> o) { k.getDestructionRates_ddC(o); }},
        {"wdot_ddT", true, [](Kinetics& k, span<double> o) { k.getNetProductionRates_ddT(o); }},
        {"wdot_ddP", true, [](Kinetics& k, span<double> o) { k.getNetProductionRates_ddP(o); }},
        {"wdot_ddC", true, [](Kinetics& k, span<double> o) { k.getNetProductionRates_ddC(o); }},
        {"deltaGibbs", false, [](Kinetics& k, span<double> o) { k.getDeltaGibbs(o); }},
        {"deltaEnthalpy", false, [](Kinetics& k, span<double> o) { k.getDeltaEnthalpy(o); }},
        {"deltaEntropy", false, [](Kinetics& k, span<double> o) { k.getDeltaEntropy(o); }},
        {"deltaSSGibbs", false, [](Kinetics& k, span<double> o) { k.getDeltaSSGibbs(o); }},
        {"deltaSSEnthalpy", false, [](Kinetics& k, span<double> o) { k.getDeltaSSEnthalpy(o); }},
        {"deltaSSEntropy", false, [](Kinetics& k, span<double> o) { k.getDeltaSSEntropy(o); }},
    };
    for (auto& g : gets) {
        const size_t w = g.species ? K : R;
        std::vector<double> v(N * w);
        for (size_t n = 0; n < N; n++) {
            gas->setState_TPY(s.T[n], s.P[n], span<const double>(&s.Y[n * K], K));
            g.fn(*kin, span<double>(&v[n * w], w));
        }
        f.add(g.name, v);
    }
    f.save(out);
    return 0;
}

// Ignition: one IdealGasConstPressureMoleReactor + ReactorNet reused across states exactly like
// samples/cxx/openmp_ignition/openmp_ignition.cpp:60-76. Composition fuel/O2/N2 at equivalence
// ratio phi ~ U(0.5, 2): X = {fuel: phi * 1, O2: nO2, N2: nO2 * 3.76} with nO2 = the
// stoichiometric O2 moles per mole of fuel.
static int runIgnite(shared_ptr<Solution> sol, const std::string& out, size_t N, uint64_t seed,
                     const std::string& fuel, double nO2, const std::string& o2, const std::string& n2,
                     double Tlo, double Thi, double Plo, double Phi, double tEnd, double dTign,
                     double rtol, double atol)
{
    auto gas = sol->thermo();
    size_t K = gas->nSpecies();
    std::mt19937_64 rng(seed);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    auto reactor = std::make_shared<IdealGasConstPressureMoleReactor>(sol, false);
    ReactorNet net(reactor);
    if (rtol > 0) net.setRelativeTolerance(rtol);
    if (atol > 0) net.setAbsoluteTolerance(atol);
    std::vector<double> T0(N), P0(N), Y0(N * K), Tend(N), Yend(N * K), tign(N), stats(N * 8);
    size_t kf = gas->speciesIndex(fuel), ko = gas->speciesIndex(o2), kn = gas->speciesIndex(n2);
    for (size_t n = 0; n < N; n++) {
        T0[n] = Tlo + (Thi - Tlo) * U(rng);
        P0[n] = OneAtm * (Plo + (Phi - Plo) * U(rng));
        double phi = 0.5 + 1.5 * U(rng);
        std::vector<double> X(K, 0.0);
        X[kf] = phi; X[ko] = nO2; X[kn] = 3.76 * nO2;
        gas->setState_TPX(T0[n], P0[n], X);
        gas->getMassFractions(span<double>(&Y0[n * K], K));
        // Re-set through TPY so the engine (which takes Y) starts from the bit-identical state
        gas->setState_TPY(T0[n], P0[n], span<const double>(&Y0[n * K], K));
        net.setInitialTime(0.0);
        tign[n] = std::nan("");
        double tPrev = 0.0, TPrev = T0[n];
        if (dTign > 0) {
            // pass 1: the openmp_ignition.cpp:70-73 step loop, crossing time by linear
            // interpolation between internal steps
            while (net.time() < tEnd && std::isnan(tign[n])) {
                double t = net.step();
                double T = reactor->temperature();
                if (T >= T0[n] + dTign) {
                    tign[n] = tPrev + (t - tPrev) * (T0[n] + dTign - TPrev) / (T - TPrev);
                }
                tPrev = t; TPrev = T;
            }
            gas->setState_TPY(T0[n], P0[n], span<const double>(&Y0[n * K], K));
            net.setInitialTime(0.0);
        }
        // pass 2: ReactorNet::advance (interpolated output at tEnd)
        net.advance(tEnd);
        Tend[n] = gas->temperature();
        gas->getMassFractions(span<double>(&Yend[n * K], K));
        AnyMap st = net.solverStats();
        const char* keys[8] = {"steps", "rhs_evals", "jac_evals", "lin_solve_setups",
                               "nonlinear_iters", "nonlinear_conv_fails", "err_test_fails",
                               "last_order"};
        for (int j = 0; j < 8; j++) {
            stats[n * 8 + j] = st.hasKey(keys[j]) ? (double) st[keys[j]].asInt() : -1.0;
        }
    }
    ArrayFile f;
    f.add("dims", std::vector<int32_t>{(int32_t) N, (int32_t) K, (int32_t) sol->kinetics()->nReactions()});
    f.add("T", T0); f.add("P", P0); f.add("Y", Y0);
    f.add("tEnd", std::vector<double>{tEnd}); f.add("dTign", std::vector<double>{dTign});
    f.add("Tend", Tend); f.add("Yend", Yend); f.add("tign", tign); f.add("stats", stats);
    f.save(out);
    return 0;
}

// Mixture-averaged transport (MixTransport): the fitted temperature polynomials the
// reference builds at start-up (public getters of GasTransport.h:107-115) and, for N seeded states, the
// viscosity, thermal conductivity and mixture-averaged diffusion coefficients it computes from them.
static int runTrans(const std::string& yaml, const std::string& phase, const std::string& out, size_t N, uint64_t seed)
{
    auto sol = newSolution(yaml, phase, "mixture-averaged");
    auto gas = sol->thermo();
    auto tr = sol->transport();
    size_t K = gas->nSpecies();
    std::vector<double> visc(K * 5), cond(K * 5), diff;
    for (size_t k = 0; k < K; k++) {
        tr->getViscosityPolynomial(k, span<double>(&visc[5 * k], 5));
        tr->getConductivityPolynomial(k, span<double>(&cond[5 * k], 5));
    }
    for (size_t i = 0; i < K; i++) {
        for (size_t j = i; j < K; j++) {
            double c[5];
            tr->getBinDiffusivityPolynomial(i, j, span<double>(c, 5));
            diff.insert(diff.end(), c, c + 5);
        }
    }
    States s = randomStates(N, K, seed);
    std::vector<double> mu(N), lambda(N), dmix(N * K);
    for (size_t n = 0; n < N; n++) {
        gas->setState_TPY(s.T[n], s.P[n], span<const double>(&s.Y[n * K], K));
        mu[n] = tr->viscosity();
        lambda[n] = tr->thermalConductivity();
        tr->getMixDiffCoeffs(span<double>(&dmix[n * K], K));
    }
    auto mw = gas->molecularWeights();
    ArrayFile f;
    f.add("dims", std::vector<int32_t>{(int32_t) N, (int32_t) K, 0});
    f.add("mol_weights", std::vector<double>(mw.begin(), mw.end()));
    f.add("visc_poly", visc); f.add("cond_poly", cond); f.add("diff_poly", diff);
    f.add("T", s.T); f.add("P", s.P); f.add("Y", s.Y);
    f.add("viscosity", mu); f.add("conductivity", lambda); f.add("mixdiff", dmix);
    f.save(out);
    return 0;
}

int main(int argc, char** argv)
{
    if (argc < 7) {
        std::cerr << "usage: gen_golden rates|ddx|rhs|rjac|ddxdigest|derivs|trans|ignite yaml phase out N seed [...]\n";
        return 2;
    }
    try {
        std::string mode = argv[1], phase = argv[3];
        if (phase == "-") phase = "";
        size_t N = std::stoul(argv[5]);
        uint64_t seed = std::stoull(argv[6]);
        if (mode == "trans") return runTrans(argv[2], phase, argv[4], N, seed);
        auto sol = newSolution(argv[2], phase, "none");
        if (mode == "rates") return runRates(sol, argv[4], N, seed);
        if (mode == "ddx") return runDdx(sol, argv[4], N, seed);
        if (mode == "rhs") return runRhs(sol, argv[4], N, seed, argc > 7 ? argv[7] : "IdealGasConstPressureMoleReactor");
        if (mode == "rjac") return runRjac(sol, argv[4], N, seed);
        if (mode == "ddxdigest") return runDdxDigest(sol, argv[4], N, seed);
        if (mode == "derivs") return runDerivs(sol, argv[4], N, seed);
        if (mode == "ignite") {
            if (argc < 17) {
                std::cerr << "ignite needs fuel nO2 O2 N2 Tlo Thi Plo Phi tEnd dTign [rtol atol]\n";
                return 2;
            }
            double rtol = argc > 17 ? std::stod(argv[17]) : -1, atol = argc > 18 ? std::stod(argv[18]) : -1;
            return runIgnite(sol, argv[4], N, seed, argv[7], std::stod(argv[8]), argv[9], argv[10],
                             std::stod(argv[11]), std::stod(argv[12]), std::stod(argv[13]),
                             std::stod(argv[14]), std::stod(argv[15]), std::stod(argv[16]), rtol, atol);
        }
        std::cerr << "unknown mode " << mode << "\n";
        return 2;
    } catch (std::exception& e) {
        std::cerr << e.what() << "\n";
        return 1;
    }
}
