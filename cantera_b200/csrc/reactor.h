// reactor.h — BatchReactorNet: N independent { IdealGasConstPressureMoleReactor + ReactorNet }
// pairs sharing one mechanism, integrated on the device by a per-reactor adaptive BDF/Newton
// solver (reactor_kernels.cu) that replaces CVODES on this path.
#pragma once

#include "engine.h"
#include "reactor_types.h"

#include <memory>

namespace ctb {

//! Counters kept per reactor, in the order of ReactorNet::solverStats
//! (CVodesIntegrator.cpp:572-621) as exported through ctb_reactornet_solverStats.
enum StatSlot {
    STAT_STEPS = 0, STAT_RHS = 1, STAT_JAC = 2, STAT_SETUPS = 3, STAT_NLITERS = 4,
    STAT_NCFAILS = 5, STAT_ETFAILS = 6, STAT_ORDER = 7
};

//! Device-side description of one advance() call; see reactor_kernels.cu.
struct ReactorBatch {
    long long N;
    int neq;               // equations per reactor (reactorNeq)
    int type, tIdx, sIdx;  // reactor model, temperature row, first species row
    double* y;             // [N * neq] current solution (T, n_1..n_K), row per reactor
    double* P;             // [N] constant pressure of each reactor (ReactorBase.cpp:44)
    double* T0;            // [N] temperature at setState (ignition monitor reference)
    double* hist;          // [N * histStride] saved integrator memory (Nordsieck array, step history)
    long long histStride;  // doubles per reactor
    double* tIgn;          // [N] first crossing of T0 + dT, NaN if none
    int32_t* status;       // [N]
    long long* stats;      // [N * CTB_NSTATS]
    unsigned long long* counter; // work-queue head
    double t0;             // time the saved state belongs to (all reactors are output at the same t)
    double tout;
    double rtol, atol;
    int atolUser;          // 1: atol applies to every component (ReactorNet.cpp:200-209);
                           // 0: MoleReactor::updateDefaultTolerances (MoleReactor.cpp:64-76)
    int maxSteps;
    int restart;           // 1: integrator is (re)initialised from y (ReactorNet::setInitialTime)
    double ignDT;          // <= 0: monitor off
    double crateFloor;     // lower bound of the Newton convergence-rate estimate (see bdf.cuh)
};

void launchAdvance(const DeviceMech& m, const ReactorBatch& b, cudaStream_t stream);

// ---- batch-synchronous integrator (reactor_batch.cu)
//! Byte offsets of the slot-interleaved arrays inside the workspace.
struct BatchLayout {
    size_t slots, pc, setupList, counters, jacScratch;
};
//! Upper bound of useful slots on this device (0: the mechanism does not fit the thread kernels).
long long batchSlotCap(const DeviceMech& m, int neq);
size_t batchWorkspaceBytes(int neq, long long S, BatchLayout* lay, size_t jacScratchBytesFor);
//! Bytes of global scratch the reactor Jacobian tile kernel may use (part of the workspace)
size_t reactorJacScratchBytes(const DeviceMech& m);
//! Integrates every reactor of b to b.tout with S slots; synchronous (returns when all reactors are
//! done). Returns the number of control/rhs/setup rounds.
long long launchAdvanceBatch(const DeviceMech& m, const ReactorBatch& b, void* workspace, long long S,
                             cudaStream_t stream);
//! y <- getState of a reactor holding (T, P, Y) in 1 m^3 (IdealGasConstPressureMoleReactor.cpp:22-30)
void launchReactorInit(const DeviceMech& m, int type, long long N, const double* T, const double* P, const double* Y,
                       long long ldY, double* y, double* Pout, double* T0, cudaStream_t stream);
//! (T, Y) of the contents (Phase::setMolesNoTruncate -> massFractions)
void launchReactorContents(const DeviceMech& m, int type, long long N, const double* y, double* T, double* Y,
                           long long ldY, cudaStream_t stream);
//! ydot = ReactorNet::eval at the current solution vectors
void launchReactorEval(const DeviceMech& m, long long N, const double* y, const double* P, double* ydot,
                       cudaStream_t stream);
//! The same through the tile kernel (reactor_batch.cu); false if the mechanism does not fit it
bool launchReactorEvalTile(const DeviceMech& m, int type, long long N, const double* y, const double* P, double* ydot,
                           cudaStream_t stream);
//! J = d(ydot)/dy of every reactor through the Jacobian tile kernel (reactor_batch.cu); scratch:
//! reactorJacScratchBytes() of device memory. false if the mechanism does not fit the kernel.
bool launchReactorJacobianTile(const DeviceMech& m, long long N, const double* y, const double* P, double* jac,
                               void* scratch, cudaStream_t stream);
//! doubles of integrator memory per reactor
long long reactorHistDoubles(int neq);

class BatchReactorNet {
public:
    //! ReactorNet(reactor) x N (ReactorNet.cpp:37-170)
    BatchReactorNet(std::shared_ptr<BatchedGasKinetics> kin, long long nReactors, int reactorType = 0);
    int reactorType() const { return m_type; }
    int nEquations() const { return m_neq; }
    ~BatchReactorNet();
    BatchReactorNet(const BatchReactorNet&) = delete;

    long long nReactors() const { return m_N; }
    //! Solution::setState_TPY on every reactor + ReactorNet::setInitialTime(0)
    void setState(const double* T, const double* P, const double* Y, long long ldY, int memspace);
    void setRelativeTolerance(double rtol); // ReactorNet.cpp:190-198
    void setAbsoluteTolerance(double atol); // ReactorNet.cpp:200-209
    void setMaxSteps(int n);                // Integrator::setMaxSteps (CVodesIntegrator.h:131)
    //! {rtol, atol of the T row, atol of the species rows, 1 if atol was set by the user (else the
    //! species value is per kmol of reactor contents)} as the integrator applies them
    void getTolerances(double* out4) const;
    void setInitialTime(double t0);         // ReactorNet.cpp:172-177
    void advance(double t);                 // ReactorNet.cpp:351-357
    double time() const { return m_time; }
    long long lastRounds() const { return m_lastRounds; }
    void getState(double* T, double* Y, long long ldY, int memspace);
    //! ReactorNet::getState (ReactorNet.cpp:688-695): y[n*(K+1) + i] = [T, n_1..n_K] of every reactor
    void getStateVector(double* y, long long bufLen, int memspace);
    //! ReactorNet::eval (ReactorNet.cpp:617-635) at the current state: ydot[n*(K+1) + i]
    void eval(double* ydot, long long bufLen, int memspace);
    //! The Jacobian d(ydot)/dy of ReactorNet::eval at the current state, dense column-major
    //! (K+1) x (K+1) per reactor: jac[n*(K+1)^2 + j*(K+1) + i] = d ydot_i / d y_j — the matrix the Newton
    //! iteration uses (IdealGasConstPressureMoleReactor::getJacobianElements,
    //! IdealGasConstPressureMoleReactor.cpp:108-174, plus the exact volume-dilution terms).
    void jacobian(double* jac, long long bufLen, int memspace);
    void status(long long len, int32_t* out);
    void solverStats(long long len, long long* out);
    void setIgnitionMonitor(double dT);
    void ignitionDelay(long long len, double* out);

private:
    double effectiveAtol() const;
    std::shared_ptr<BatchedGasKinetics> m_kin;
    long long m_N;
    int m_type = 0;
    int m_neq;
    double m_time = 0.0;
    double m_rtol = 1e-9, m_atol = 1e-15; // ReactorNet.h:483-485
    bool m_atolUser = false;
    int m_maxSteps = 20000;               // CVodesIntegrator.h:131
    bool m_haveState = false;
    bool m_restart = true;
    double m_ignDT = 0.0;
    DevBuf m_y, m_P, m_T0, m_hist, m_tIgn, m_status, m_stats, m_counter, m_io, m_ws;
    long long m_slots = 0;
    long long m_lastRounds = 0;
};

} // namespace ctb
