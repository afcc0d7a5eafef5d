/*
 * ctb_oracle.h — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C, one-state-at-a-time CPU restatement of the reference's algorithm for the batched
 * ideal-gas kinetics path (SURVEY.md §8 rows a1–a13). Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library; the CUDA product path
 * (cantera_b200/libctb.so) never links or calls it.
 *
 * Parity status: PINNED. The restatement is checked (tests/test_oracle.py) against golden
 * vectors produced by the unmodified reference compiled from /root/reference
 * (oracle/ref_build/build_ref.sh + gen_golden.cpp -> tests/golden/), which in turn reproduces
 * the reference's own known-answer tests (test/kinetics/pdep.cpp:52-222 etc.,
 * oracle/ref_build/ref_kat.cpp).
 *
 * Every function names the reference file:line it follows. Operation ORDER is kept identical to
 * the reference wherever it is observable in the last bits.
 */
#ifndef CTB_ORACLE_H
#define CTB_ORACLE_H

#include "../include/ctb_clib.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ctbo_mech ctbo_mech;

/* Load a .ctbm table file (format: include/cantera_b200/MechTables.h). NULL on failure. */
ctbo_mech* ctbo_load(const char* path);
void ctbo_free(ctbo_mech* m);
const ctb_mech_tables* ctbo_tables(const ctbo_mech* m);
int32_t ctbo_nSpecies(const ctbo_mech* m);
int32_t ctbo_nReactions(const ctbo_mech* m);
void ctbo_setMultiplier(ctbo_mech* m, int32_t i, double f);

/* setState_TPY + BulkKinetics::updateROP + getters. Any output pointer may be NULL.
 *  kf    [R] getFwdRateConstants (BulkKinetics.cpp:153-161, default non-legacy: no [M])
 *  kc    [R] getEquilibriumConstants (BulkKinetics.cpp:163-179)
 *  ropf/ropr/ropnet [R] (Kinetics.cpp:375-397)
 *  wdot/cdot/ddot [K]  (Kinetics.cpp:420-457)
 *  concm [R] thirdBodyConcentrations (0 where the reaction has no third body)
 * Returns 0, or 1 if the mass fractions cannot be normalised (Phase.cpp:349-353). */
int32_t ctbo_rates(const ctbo_mech* m, double T, double P, const double* Y,
                   double* kf, double* kc, double* ropf, double* ropr, double* ropnet,
                   double* wdot, double* cdot, double* ddot, double* concm);

/* Same for n states: Y[n*ldY + k], outputs [n*K + k] / [n*R + i]. */
int32_t ctbo_rates_batch(const ctbo_mech* m, int64_t n, const double* T, const double* P,
                         const double* Y, int64_t ldY, double* kf, double* kc, double* ropf,
                         double* ropr, double* ropnet, double* wdot, double* cdot, double* ddot);

/* Kinetics::netProductionRates_ddX (ddCi = 0) or _ddCi (ddCi = 1), dense column-major K x K:
 * jac[j*K + k] = d wdot_k / d X_j (Kinetics.cpp:604-612, BulkKinetics.cpp:408-455,662-701).
 * analyticFalloff = 0 reproduces the reference's finite difference in [M] with rtolDelta
 * (MultiRate.h:135-165); = 1 uses the exact derivative of the falloff expressions. */
int32_t ctbo_ddX(const ctbo_mech* m, double T, double P, const double* Y, int32_t skipThirdBodies,
                 int32_t skipFalloff, double rtolDelta, int32_t analyticFalloff, int32_t ddCi,
                 double* jac);

/* IdealGasConstPressureMoleReactor::updateState + eval + ReactorNet::eval
 * (IdealGasConstPressureMoleReactor.cpp:38-106, ReactorNet.cpp:617-635):
 * y = [T, n_1..n_K] (kmol), P constant; ydot = RHS/LHS. */
int32_t ctbo_reactor_rhs(const ctbo_mech* m, double P, const double* y, double* ydot);

/* Initial reactor state from (T, P, Y) with volume 1 m^3
 * (IdealGasConstPressureMoleReactor::getState :22-30, MoleReactor::getMoles): y[K+1]. */
int32_t ctbo_reactor_initial_state(const ctbo_mech* m, double T, double P, const double* Y,
                                   double* y);

/* (T, Y) of the contents for a reactor state vector y (Phase::setMolesNoTruncate). */
int32_t ctbo_reactor_contents(const ctbo_mech* m, const double* y, double* T, double* Y);

#ifdef __cplusplus
}
#endif
#endif
