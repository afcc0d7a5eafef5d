/*
 * ctb_clib.h — C ABI of the B200-native batched ideal-gas kinetics engine.
 *
 * This is the drop-in boundary: plain C, integer handles, caller-owned buffers, no exception
 * ever crosses it. It sits beside the reference's generated CLib and follows its conventions:
 *   - handles into process-global cabinets      (interfaces/clib/src/Cabinet.h:57-206)
 *   - error return -1 / ERR=-999 / DERR=-999.999 (interfaces/clib/include/cantera_clib/clib_defs.h:27-33,
 *                                                 interfaces/clib/src/clib_utils.h:34-56)
 *   - message fetched with a getter             (ct_getCanteraError, sourcegen/headers/ct.yaml:46-49)
 *   - (len, ptr) array arguments, size-checked  (sourcegen/clib/templates.yaml "size_check")
 *
 * Each entry point names the reference interface it replaces for a BATCH of independent
 * thermochemical states. One handle = one CUDA device + one stream; calls on a handle are
 * serialised by the caller (the reference's objects are not thread-safe either:
 * samples/cxx/openmp_ignition/openmp_ignition.cpp:26-29).
 *
 * Batch layout: T[n], P[n] and Y[n*ldY + k] (one state per row, species contiguous; ldY >= K).
 * Per-state outputs use the same row layout with their own leading dimension.
 * memspace: CTB_HOST (0) = host pointers; the library copies in chunks through device staging buffers on
 *           four streams (DESIGN.md "End-to-end path"). Pinned buffers (cudaHostAlloc / cudaHostRegister)
 *           are read and written by the DMA engines directly; pageable buffers of the rate getters go
 *           through the handle's own pinned ring (packed and unpacked by host threads, about a third of
 *           the pinned throughput); pageable buffers of the other calls are staged by the driver;
 *           CTB_DEVICE (1) = device pointers on the handle's device, no copies.
 *
 * There is no CPU fallback: if no CUDA device is usable every constructor fails with -1.
 */
#ifndef CTB_CLIB_H
#define CTB_CLIB_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CTB_ERR  (-999)
#define CTB_DERR (-999.999)
#define CTB_HOST   0
#define CTB_DEVICE 1

/* Reaction rate parameterisations (ReactionRateFactory.cpp:39-110, subset on the hot path) */
#define CTB_RATE_ARRHENIUS 0 /* ArrheniusRate, elementary or three-body (Arrhenius.h:196) */
#define CTB_RATE_LINDEMANN 1 /* LindemannRate (Falloff.h:120-122)                          */
#define CTB_RATE_TROE      2 /* TroeRate (Falloff.cpp:283-300)                             */
#define CTB_RATE_SRI       3 /* SriRate (Falloff.cpp:389-403)                              */
#define CTB_RATE_TSANG     4 /* TsangRate (Falloff.cpp:480-494)                            */
#define CTB_RATE_PLOG      5 /* PlogRate (PlogRate.h:110-164)                              */
#define CTB_RATE_CHEBYSHEV 6 /* ChebyshevRate (ChebyshevRate.h:133-171)                    */

#define CTB_RXN_REVERSIBLE      1 /* Reaction::reversible                                   */
#define CTB_RXN_THIRD_BODY      2 /* Reaction::thirdBody() != nullptr                       */
#define CTB_RXN_MASS_ACTION     4 /* ThirdBody::mass_action (Reaction.h:257)                */
#define CTB_RXN_CHEM_ACTIVATED  8 /* FalloffRate::chemicallyActivated() (Falloff.h:167-171) */

/*
 * Flat mechanism tables: what BatchedGasKinetics extracts from a loaded reference Solution
 * through its public API (include/cantera_b200/MechanismExtract.h), or what a .ctbm file holds.
 * All reaction-indexed arrays are in the reference's reaction order. SI units as in the
 * reference (kmol, m, s, J, K, Pa); Ea already divided by GasConstant (ArrheniusBase::m_Ea_R).
 */
typedef struct ctb_mech_tables {
    int32_t n_species;           /* K */
    int32_t n_reactions;         /* R */
    double  p_ref;               /* SpeciesThermoInterpType::refPressure, Pa */
    const double* mol_weights;   /* [K] Phase::molecularWeights */
    const double* nasa_tmid;     /* [K] NasaPoly2::m_midT */
    const double* nasa_tmin;     /* [K] */
    const double* nasa_tmax;     /* [K] */
    const double* nasa_low;      /* [K*7] a0..a6 for T <= Tmid (NasaPoly2.h:107) */
    const double* nasa_high;     /* [K*7] */

    const int32_t* rxn_type;     /* [R] CTB_RATE_* */
    const int32_t* rxn_flags;    /* [R] CTB_RXN_* bits */
    const double* rate_A;        /* [R] ArrheniusBase::m_A (high-pressure limit for falloff) */
    const double* rate_b;        /* [R] */
    const double* rate_Ea_R;     /* [R] */
    const double* low_A;         /* [R] FalloffRate::lowRate(), 0 for other types */
    const double* low_b;         /* [R] */
    const double* low_Ea_R;      /* [R] */
    const double* falloff_coeffs;/* [R*5] Troe: a, 1/T3, 1/T1, T2 ; SRI: a,b,c,d,e ; Tsang: a,b */
    const double* multiplier;    /* [R] Kinetics::multiplier(i) (m_perturb) */
    const double* delta_n;       /* [R] BulkKinetics::m_dn (BulkKinetics.cpp:28-36) */

    const int32_t* reac_ptr;     /* [R+1] CSR over the reactant list handed to m_reactantStoich.add */
    const int32_t* reac_species; /*       (Kinetics.cpp:740-777; includes order-only species)        */
    const double*  reac_stoich;
    const double*  reac_order;
    const int32_t* prod_ptr;     /* [R+1] CSR over products (order == stoich) */
    const int32_t* prod_species;
    const double*  prod_stoich;

    const double*  tb_default;   /* [R] ThirdBody::default_efficiency (0 where no third body) */
    const int32_t* tb_ptr;       /* [R+1] CSR over explicit efficiencies, species ascending   */
    const int32_t* tb_species;   /*       (BulkKinetics.cpp:66-84)                             */
    const double*  tb_eff;       /*       raw efficiencies (the library forms eff - default)  */

    const int32_t* plog_ptr;     /* [R+1] CSR over (P, Arrhenius) entries, P ascending */
    const double*  plog_P;       /* Pa */
    const double*  plog_A;
    const double*  plog_b;
    const double*  plog_Ea_R;

    const int32_t* cheb_ptr;     /* [R+1] CSR over coefficient data (row-major nT x nP) */
    const int32_t* cheb_nT;      /* [R] */
    const int32_t* cheb_nP;      /* [R] */
    const double*  cheb_limits;  /* [R*4] Tmin, Tmax, Pmin, Pmax */
    const double*  cheb_coeffs;  /* includes the log10 unit offset (ChebyshevRate.cpp:82-83) */
} ctb_mech_tables;

/* ---- library ---------------------------------------------------------------------------- */

/* ct_getCanteraError (ct.yaml:46-49): copies the last error of this thread; returns the length
 * needed (including the terminator) when called with (0, NULL). */
int32_t ctb_getError(int32_t bufLen, char* buf);
/* Sets this thread's error message (for companion libraries that report through ctb_getError, e.g. the
 * reference bridge libctb_cantera.so). Returns 0. */
int32_t ctb_setError(const char* message);
/* Number of usable CUDA devices, or -1 with an error set. */
int32_t ctb_nDevices(void);
/* "0.1.0 sm_100a" style version string; same length convention as ctb_getError. */
int32_t ctb_version(int32_t bufLen, char* buf);
/* Measured FP64 throughput of `device` in TFLOP/s: a register-resident DFMA micro-benchmark
 * (8 independent chains per thread, every SM full), timed with CUDA events. This is the roofline
 * denominator of the FP64-bound kernels (MEASURED_PEAKS.json has no FP64 entry).
 * Returns -999.999 with an error set on failure. */
double ctb_fp64PeakTflops(int32_t device);
/* The same measurement for the FP64 tensor-core instruction (mma.sync m8n8k4, DMMA): the number that decides
 * whether a tensor-core Newton factorisation could pay. */
double ctb_fp64DmmaPeakTflops(int32_t device);

/* ---- BatchedGasKinetics: Kinetics / BulkKinetics for N states ---------------------------- */

/* newKinetics (KineticsFactory.cpp:65-97) for a mechanism already reduced to flat tables.
 * Returns a handle >= 0. */
int32_t ctb_kin_newFromTables(const ctb_mech_tables* tables, int32_t device);
/* (ctb_kin_new(yaml, phase): see include/ctb_cantera.h, the optional reference bridge.) */
/* Same, from a .ctbm table file written by the extractor (oracle/ref_build/ctb_extract). */
int32_t ctb_kin_newFromFile(const char* ctbmPath, int32_t device);
int32_t ctb_kin_del(int32_t handle);
/* Run subsequent device-memory calls on an existing cudaStream_t; NULL is the CUDA default stream
 * (what torch.cuda.current_stream() is unless changed). Until this is called the handle uses a
 * private non-blocking stream. */
int32_t ctb_kin_setStream(int32_t handle, void* cudaStream);
/* Block until all work queued on the handle's stream has finished. */
int32_t ctb_kin_synchronize(int32_t handle);

int32_t ctb_kin_nSpecies(int32_t handle);              /* kin_nTotalSpecies (ctkin.yaml) */
int32_t ctb_kin_nReactions(int32_t handle);            /* kin_nReactions */
int32_t ctb_kin_setMultiplier(int32_t handle, int32_t i, double f); /* Kinetics::setMultiplier */
double  ctb_kin_multiplier(int32_t handle, int32_t i);              /* Kinetics::multiplier */
/* Kinetics::setDerivativeSettings (Kinetics.h:676-685, BulkKinetics.cpp:259-282):
 * skip-third-bodies, skip-falloff, rtol-delta. */
int32_t ctb_kin_setDerivativeSettings(int32_t handle, int32_t skipThirdBodies,
                                      int32_t skipFalloff, double rtolDelta);

/* setState_TPY + Kinetics::getNetProductionRates (Kinetics.cpp:447-457): wdot[n*ldW + k],
 * kmol/m^3/s. */
int32_t ctb_kin_getNetProductionRates(int32_t handle, int64_t nStates, const double* T,
                                      const double* P, const double* Y, int64_t ldY,
                                      double* wdot, int64_t ldW, int32_t memspace);
/* Kinetics::getCreationRates / getDestructionRates (Kinetics.cpp:420-445) */
int32_t ctb_kin_getCreationRates(int32_t handle, int64_t nStates, const double* T,
                                 const double* P, const double* Y, int64_t ldY,
                                 double* cdot, int64_t ldW, int32_t memspace);
int32_t ctb_kin_getDestructionRates(int32_t handle, int64_t nStates, const double* T,
                                    const double* P, const double* Y, int64_t ldY,
                                    double* ddot, int64_t ldW, int32_t memspace);
/* Kinetics::getFwdRatesOfProgress / getRevRatesOfProgress / getNetRatesOfProgress
 * (Kinetics.cpp:375-397): rop[n*ldR + i]. */
int32_t ctb_kin_getFwdRatesOfProgress(int32_t handle, int64_t nStates, const double* T,
                                      const double* P, const double* Y, int64_t ldY,
                                      double* rop, int64_t ldR, int32_t memspace);
int32_t ctb_kin_getRevRatesOfProgress(int32_t handle, int64_t nStates, const double* T,
                                      const double* P, const double* Y, int64_t ldY,
                                      double* rop, int64_t ldR, int32_t memspace);
int32_t ctb_kin_getNetRatesOfProgress(int32_t handle, int64_t nStates, const double* T,
                                      const double* P, const double* Y, int64_t ldY,
                                      double* rop, int64_t ldR, int32_t memspace);
/* BulkKinetics::getFwdRateConstants (BulkKinetics.cpp:153-161; includes [M] of mass-action
 * third bodies and the multiplier) and getEquilibriumConstants (:141-151). */
int32_t ctb_kin_getFwdRateConstants(int32_t handle, int64_t nStates, const double* T,
                                    const double* P, const double* Y, int64_t ldY,
                                    double* kf, int64_t ldR, int32_t memspace);
int32_t ctb_kin_getEquilibriumConstants(int32_t handle, int64_t nStates, const double* T,
                                        const double* P, const double* Y, int64_t ldY,
                                        double* kc, int64_t ldR, int32_t memspace);
/* Kinetics::netProductionRates_ddX (Kinetics.cpp:604-607), dense column-major K x K per state
 * exactly as kin_getNetProductionRates_ddX returns it (ctkin.yaml:65-90):
 * jac[n*K*K + j*K + k] = d wdot_k / d X_j. bufLen >= nStates*K*K. */
int32_t ctb_kin_getNetProductionRates_ddX(int32_t handle, int64_t nStates, const double* T,
                                          const double* P, const double* Y, int64_t ldY,
                                          double* jac, int64_t bufLen, int32_t memspace);
/* Kinetics::netProductionRates_ddCi (BulkKinetics.cpp:443-455 contracted with the
 * stoichiometric matrix), same layout. */
int32_t ctb_kin_getNetProductionRates_ddCi(int32_t handle, int64_t nStates, const double* T,
                                           const double* P, const double* Y, int64_t ldY,
                                           double* jac, int64_t bufLen, int32_t memspace);

/* NO COUNTERPART IN THE REFERENCE: added for callers that may not
 * allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc). Same result as ctb_kin_getNetProductionRates_ddX with memspace = CTB_DEVICE (device
 * pointers only; queued on the handle's stream; derivative settings honoured), but the call makes NO
 * cudaMalloc / cudaFree / allocating library call: the Jacobian tile kernel's global scratch is a
 * file-scope static __device__ array of CTB_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds
 * until exit, and the grid is sized to the tiles that array holds (fewer tiles in flight than the
 * allocating entry point for large mechanisms: same numbers, lower throughput).
 * Limits, rejected with -1 and a message: nStates > CTB_NOALLOC_MAX_STATES; a mechanism whose scratch for
 * ONE tile of 16 states, (3 R + 1 + nThirdBody) * 128 bytes, exceeds the array; a mechanism that does not
 * fit the tile kernel's shared memory.
 * Every call on a device shares the one array (also across handles): calls that overlap on the GPU
 * (different streams, no dependency) RACE; serialise them. Nothing has to be cleared between calls: each
 * launch rewrites all it reads, including the one row that must be zero. The first call on a device
 * resolves the array's address (cudaGetSymbolAddress) and loads the kernel: make it before a capture. */
#define CTB_NOALLOC_SCRATCH_BYTES (32 * 1024 * 1024)
#define CTB_NOALLOC_MAX_STATES (1 << 22)
int32_t ctb_kin_getNetProductionRates_ddX_noAlloc(int32_t handle, int64_t nStates, const double* T,
                                                  const double* P, const double* Y, int64_t ldY,
                                                  double* jac, int64_t bufLen);

/* The remaining Kinetics getters, one entry point: out[n*ld + i] with row width R (reaction quantities)
 * or K (species quantities). quantity = CTB_Q_*:
 *   getFwdRateConstants_ddT/_ddP/_ddC, get{Fwd,Rev,Net}RatesOfProgress_ddT/_ddP/_ddC (BulkKinetics.cpp:284-385),
 *   get{Creation,Destruction,NetProduction}Rates_ddT/_ddP/_ddC (Kinetics.cpp:459-602),
 *   getDeltaGibbs / Enthalpy / Entropy and getDeltaSSGibbs / SSEnthalpy / SSEntropy (BulkKinetics.cpp:200-257).
 * Derivative semantics are the reference's (Kinetics.h:600-700): _ddT at constant pressure-independent
 * concentrations, _ddC with respect to the total molar concentration at constant T and X; finite
 * differences with rtol-delta where the reference differences (non-Arrhenius rates, 1/Kc). */
#define CTB_Q_KF_DDT      0
#define CTB_Q_KF_DDP      1
#define CTB_Q_KF_DDC      2
#define CTB_Q_ROPF_DDT    3
#define CTB_Q_ROPF_DDP    4
#define CTB_Q_ROPF_DDC    5
#define CTB_Q_ROPR_DDT    6
#define CTB_Q_ROPR_DDP    7
#define CTB_Q_ROPR_DDC    8
#define CTB_Q_ROPNET_DDT  9
#define CTB_Q_ROPNET_DDP  10
#define CTB_Q_ROPNET_DDC  11
#define CTB_Q_CDOT_DDT    12
#define CTB_Q_CDOT_DDP    13
#define CTB_Q_CDOT_DDC    14
#define CTB_Q_DDOT_DDT    15
#define CTB_Q_DDOT_DDP    16
#define CTB_Q_DDOT_DDC    17
#define CTB_Q_WDOT_DDT    18
#define CTB_Q_WDOT_DDP    19
#define CTB_Q_WDOT_DDC    20
#define CTB_Q_DELTA_GIBBS       21
#define CTB_Q_DELTA_ENTHALPY    22
#define CTB_Q_DELTA_ENTROPY     23
#define CTB_Q_DELTA_SS_GIBBS    24
#define CTB_Q_DELTA_SS_ENTHALPY 25
#define CTB_Q_DELTA_SS_ENTROPY  26
#define CTB_Q_COUNT             27
int32_t ctb_kin_getQuantity(int32_t handle, int32_t quantity, int64_t nStates, const double* T, const double* P,
                            const double* Y, int64_t ldY, double* out, int64_t ld, int32_t memspace);

/* ---- BatchedMixTransport: mixture-averaged transport for N states -------- */

/* newTransport("mixture-averaged") (TransportFactory.cpp, MixTransport::init) for tables: the fitted
 * temperature polynomials of the reference's GasTransport (non-CK mode, 5 coefficients in ln T), read
 * through its public getters getViscosityPolynomial / getConductivityPolynomial /
 * getBinDiffusivityPolynomial (GasTransport.h:107-115): viscPoly, condPoly [K*5]; diffPoly [K(K+1)/2*5],
 * pairs i <= j row by row (GasTransport.cpp:778-795); molWeights [K]. Returns a handle or -1. */
int32_t ctb_trans_newFromArrays(int32_t nSpecies, const double* molWeights, const double* viscPoly,
                                const double* condPoly, const double* diffPoly, int32_t device);
int32_t ctb_trans_del(int32_t handle);
int32_t ctb_trans_nSpecies(int32_t handle);
/* Stream of the device-memory calls, as ctb_kin_setStream. */
int32_t ctb_trans_setStream(int32_t handle, void* cudaStream);
int32_t ctb_trans_synchronize(int32_t handle);
/* trans_viscosity, trans_thermalConductivity and trans_getMixDiffCoeffs (sourcegen/headers/cttrans.yaml:12-16;
 * GasTransport::viscosity GasTransport.cpp:60-82, MixTransport::thermalConductivity MixTransport.cpp:33-50,
 * GasTransport::getMixDiffCoeffs GasTransport.cpp:170-201) after setState_TPY, for every state of the batch:
 * viscosity[n] (Pa s), conductivity[n] (W/m/K), mixDiff[n*ldD + k] (m^2/s). Any of the three output
 * pointers may be NULL. */
int32_t ctb_trans_getMixtureTransport(int32_t handle, int64_t nStates, const double* T, const double* P,
                                      const double* Y, int64_t ldY, double* viscosity, double* conductivity,
                                      double* mixDiff, int64_t ldD, int32_t memspace);

/* ---- BatchReactorNet: N x { IdealGasConstPressureMoleReactor + ReactorNet } --------------- */

/* ReactorNet(reactor) for nReactors independent reactors sharing the mechanism of a kinetics
 * handle (ReactorNet.cpp:37-170). Each reactor starts with volume 1 m^3 (ReactorBase::m_vol). */
int32_t ctb_reactornet_new(int32_t kinHandle, int64_t nReactors);
/* The same for the other ideal-gas reactor models (reactor_new + ReactorNet, ctreactor.yaml / ReactorFactory):
 * solution vectors y are  [T, n_k] | [T, V, n_k] | [m, V, T, Y_k] | [m, T, Y_k]  (kmol, m^3, kg). The models
 * other than the first integrate with a forward-difference Newton matrix (n + 1 right-hand sides per setup). */
#define CTB_REACTOR_CONST_PRESSURE_MOLE 0 /* IdealGasConstPressureMoleReactor (IdealGasConstPressureMoleReactor.cpp) */
#define CTB_REACTOR_MOLE                1 /* IdealGasMoleReactor, constant volume (IdealGasMoleReactor.cpp:97-172)   */
#define CTB_REACTOR_IDEAL_GAS           2 /* IdealGasReactor, constant volume (IdealGasReactor.cpp:38-117)           */
#define CTB_REACTOR_CONST_PRESSURE      3 /* IdealGasConstPressureReactor (IdealGasConstPressureReactor.cpp:36-106)  */
int32_t ctb_reactornet_newTyped(int32_t kinHandle, int64_t nReactors, int32_t reactorType);
/* ReactorNet::neq per reactor: K + 1, K + 2, K + 3, K + 2 for the four models. */
int32_t ctb_reactornet_nEquations(int32_t handle);
int32_t ctb_reactornet_del(int32_t handle);
/* Solution::setState_TPY on every reactor's contents + ReactorNet::setInitialTime(0)
 * (ReactorNet.cpp:172-177): the integrator restarts from these states. */
int32_t ctb_reactornet_setState(int32_t handle, const double* T, const double* P,
                                const double* Y, int64_t ldY, int32_t memspace);
int32_t ctb_reactornet_setRelativeTolerance(int32_t handle, double rtol); /* ReactorNet.cpp:190-198 */
int32_t ctb_reactornet_setAbsoluteTolerance(int32_t handle, double atol); /* ReactorNet.cpp:200-209 */
int32_t ctb_reactornet_setMaxSteps(int32_t handle, int32_t maxSteps);     /* Integrator::setMaxSteps */
/* The tolerances the integrator actually applies (ReactorNet::rtol / atol, ReactorNet.h:101-108):
 * out[0] = rtol, out[1] = absolute tolerance of the temperature row, out[2] = absolute tolerance of
 * the species rows, out[3] = 1 if the absolute tolerance was set by the caller, else 0. A caller-set
 * value is applied unchanged to every row (ReactorNet.cpp:200-209). The default (1e-15, times the
 * reactor's total moles for the species rows: MoleReactor.cpp:64-76, so out[2] is then per kmol) is
 * applied with a safety factor 0.1 — see DESIGN.md "Tolerance mapping". len >= 4. */
int32_t ctb_reactornet_getTolerances(int32_t handle, int32_t len, double* out);
int32_t ctb_reactornet_setInitialTime(int32_t handle, double t0);
/* ReactorNet::advance(t) (ReactorNet.cpp:351-357) for every reactor. Numerical failure of a
 * reactor never fails the call: it is reported through ctb_reactornet_status. */
int32_t ctb_reactornet_advance(int32_t handle, double t);
double  ctb_reactornet_time(int32_t handle);                              /* ReactorNet::time */
/* Contents after the last advance: T[n], Y[n*ldY + k] (ThermoPhase::temperature, massFractions) */
int32_t ctb_reactornet_getState(int32_t handle, double* T, double* Y, int64_t ldY,
                                int32_t memspace);
/* ReactorNet::getState (ReactorNet.cpp:688-695): the solution vectors y[n*(K+1) + i] = [T, n_1..n_K]
 * (kmol) of every reactor (layout by reactor model). bufLen >= nReactors*nEquations. */
int32_t ctb_reactornet_getStateVector(int32_t handle, double* y, int64_t bufLen, int32_t memspace);
/* ReactorNet::eval (ReactorNet.cpp:617-635; IdealGasConstPressureMoleReactor::eval
 * IdealGasConstPressureMoleReactor.cpp:52-106) at the current solution vectors: ydot, same layout. */
int32_t ctb_reactornet_eval(int32_t handle, double* ydot, int64_t bufLen, int32_t memspace);
/* Reactor::jacobian / IdealGasConstPressureMoleReactor::getJacobianElements
 * (IdealGasConstPressureMoleReactor.cpp:108-174) at the current solution vectors, dense column-major
 * (K+1) x (K+1) per reactor: jac[n*(K+1)^2 + j*(K+1) + i] = d ydot_i / d y_j, y = [T, n_1..n_K].
 * This is the matrix the batched Newton iteration uses. It differs from the reference's semi-analytic
 * preconditioner template in two documented ways: the species block carries the exact
 * volume-dilution terms (d(V wdot_k)/d n_j = ddCi_kj + wdot_k / c_tot - sum_l ddCi_kl X_l; the reference
 * keeps ddCi_kj only, :114-123) and the falloff [M]-derivative is analytic instead of a finite
 * difference (MultiRate.h:135-165). T column: one-sided difference with dT = T sqrt(eps) as :126-152;
 * T row as :153-173 on this species block. bufLen >= nReactors*(K+1)^2. */
int32_t ctb_reactornet_getJacobian(int32_t handle, double* jac, int64_t bufLen, int32_t memspace);
/* Per-reactor integrator status: 0 ok, >0 failure code (CTB_STATUS_*). */
int32_t ctb_reactornet_status(int32_t handle, int64_t len, int32_t* status);
/* Per-reactor counters, stats[n*CTB_NSTATS + j] (ReactorNet::solverStats,
 * CVodesIntegrator.cpp:572-621): steps, rhs_evals, jac_evals, lin_setups, nonlin_iters,
 * nonlin_conv_fails, err_test_fails, last_order. */
#define CTB_NSTATS 8
int32_t ctb_reactornet_solverStats(int32_t handle, int64_t len, int64_t* stats);
/* Number of control/rhs/setup kernel rounds the batch-synchronous integrator needed in the last
 * advance (0 when the persistent one-kernel integrator ran). */
int32_t ctb_reactornet_lastRounds(int32_t handle);
/* First time each reactor's temperature crossed T0 + dT (ignition delay as in
 * samples/cxx/openmp_ignition/openmp_ignition.cpp:71); NaN where it has not. Set dT before
 * advancing; dT <= 0 disables the monitor. */
int32_t ctb_reactornet_setIgnitionMonitor(int32_t handle, double dT);
int32_t ctb_reactornet_ignitionDelay(int32_t handle, int64_t len, double* tIgn);

#define CTB_STATUS_OK              0
#define CTB_STATUS_TOO_MANY_STEPS  1 /* CV_TOO_MUCH_WORK (m_maxsteps, CVodesIntegrator.h:131) */
#define CTB_STATUS_ERR_TEST_FAIL   2 /* CV_ERR_FAILURE  */
#define CTB_STATUS_CONV_FAIL       3 /* CV_CONV_FAILURE */
#define CTB_STATUS_STEP_TOO_SMALL  4 /* CV_TOO_MUCH_ACC / h underflow */
#define CTB_STATUS_NOT_FINITE      5 /* ReactorNet::checkFinite (ReactorNet.cpp:634) */

#ifdef __cplusplus
}
#endif
#endif
