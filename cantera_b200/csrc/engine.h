// engine.h — host-side C++ core behind the C ABI: BatchedGasKinetics (the batched counterpart of
// Kinetics/BulkKinetics for one mechanism on one CUDA device). BatchReactorNet lives in reactor.h.
#pragma once

#include "../../include/cantera_b200/MechTables.h"
#include "kin_kernels.h"

#include <cuda_runtime.h>

#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace ctb {

//! Error type translated to return code -1 at the C boundary (the reference's CanteraError,
//! interfaces/clib/src/clib_utils.h:34-56).
struct CtbError : public std::runtime_error {
    CtbError(const std::string& where, const std::string& what)
        : std::runtime_error(where + ": " + what) {}
};

void cudaCheck(cudaError_t e, const char* where);

//! RAII device / pinned buffers
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    void reserve(size_t n);
    ~DevBuf();
    template <class T> T* as() { return static_cast<T*>(p); }
};

//! Pinned host memory (cudaHostAlloc): the staging ring of host-memory calls with pageable buffers.
struct PinBuf {
    void* p = nullptr;
    size_t bytes = 0;
    void reserve(size_t n);
    ~PinBuf();
    template <class T> T* as() { return static_cast<T*>(p); }
};

struct RateOutputs {
    double* wdot = nullptr;
    double* cdot = nullptr;
    double* ddot = nullptr;
    long long ldW = 0;
    double* ropf = nullptr;
    double* ropr = nullptr;
    double* ropnet = nullptr;
    double* kf = nullptr;
    double* kc = nullptr;
    long long ldR = 0;
};

class BatchedGasKinetics {
public:
    //! newKinetics (KineticsFactory.cpp:65-97) for a mechanism reduced to flat tables.
    BatchedGasKinetics(const ctb_mech_tables& tables, int device);
    ~BatchedGasKinetics();
    BatchedGasKinetics(const BatchedGasKinetics&) = delete;

    int nSpecies() const { return m_tables.n_species; }     // Kinetics::nTotalSpecies
    int nReactions() const { return m_tables.n_reactions; } // Kinetics::nReactions
    void setMultiplier(int i, double f);                    // Kinetics::setMultiplier
    double multiplier(int i) const;                         // Kinetics::multiplier
    //! Kinetics::setDerivativeSettings (BulkKinetics.cpp:259-282)
    void setDerivativeSettings(bool skipThirdBodies, bool skipFalloff, double rtolDelta);

    void setStream(cudaStream_t s) { m_userStream = s; m_hasUserStream = true; }
    cudaStream_t stream() const { return m_hasUserStream ? m_userStream : m_stream; }
    void synchronize();

    //! Any subset of the rate getters for N states (setState_TPY + updateROP + getters).
    void rates(long long N, const double* T, const double* P, const double* Y, long long ldY,
               const RateOutputs& out, int memspace);
    //! netProductionRates_ddX (ddCi = false) or _ddCi, dense column-major K x K per state.
    void jacobian(long long N, const double* T, const double* P, const double* Y, long long ldY,
                  double* jac, long long bufLen, bool ddCi, int memspace);

    //! netProductionRates_ddX for device pointers without any device allocation (static scratch array of
    //! kin_kernels.cu; ctb_kin_getNetProductionRates_ddX_noAlloc).
    void jacobianNoAlloc(long long N, const double* T, const double* P, const double* Y, long long ldY,
                         double* jac, long long bufLen);

    //! One of the remaining getters (CTB_Q_* of include/ctb_clib.h: the _ddT/_ddP/_ddC families and the
    //! reaction deltas) for N states: out[n*ld + i], width K (species quantities) or R (reaction quantities).
    void quantity(int what, long long N, const double* T, const double* P, const double* Y, long long ldY,
                  double* out, long long ld, int memspace);

    const DeviceMech& mech() const { return m_dev; }
    const cantera_b200::MechTables& tables() const { return m_tables; }
    int device() const { return m_dev.device; }

private:
    void uploadMultipliers();
    void ratesPageable(long long N, const double* T, const double* P, const double* Y, long long ldY,
                       const RateOutputs& out, bool full);
    cantera_b200::MechTables m_tables;
    PackedMech m_packed;
    DeviceMech m_dev;
    cudaStream_t m_stream = nullptr;     // own stream
    static constexpr int NSTAGE = 4; // chunks in flight: H2D of one overlaps the kernel and D2H of others
    cudaStream_t m_copyStream[NSTAGE] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t m_evt[NSTAGE] = {nullptr, nullptr, nullptr, nullptr};
    cudaStream_t m_userStream = nullptr;
    bool m_hasUserStream = false;
    bool m_skip3b = false, m_skipFalloff = false;
    double m_rtolDelta = 1e-8;
    DevBuf m_stage[NSTAGE]; // device staging for host-memory calls
    PinBuf m_pin[NSTAGE];   // pinned ring for pageable host buffers (rates)
    DevBuf m_jacScratch[NSTAGE + 1]; // per-stream scratch of the Jacobian tile kernel (last: the handle's own stream)
};

} // namespace ctb
