// ctb_clib.cu — the extern "C" boundary (include/ctb_clib.h). Handle cabinets, error stack and
// exception translation follow the reference's CLib conventions
// (interfaces/clib/src/Cabinet.h:57-206, interfaces/clib/src/clib_utils.h:34-56).
#include <algorithm>
#include "../../include/ctb_clib.h"
#include "engine.h"
#include "reactor.h"
#include "transport.h"

#include <cmath>
#include <cstring>
#include <memory>
#include <mutex>
#include <vector>

using namespace ctb;

namespace {

thread_local std::string t_lastError; // per-thread error stack (src/base/application.h:121-146)

//! Cabinet<T>: process-global vector of shared_ptr; handles are indices; deleted slots stay
//! null so that use-after-delete raises and double-delete is harmless (Cabinet.h:101-139).
template <class T>
class Cabinet {
public:
    static int add(std::shared_ptr<T> obj) {
        std::lock_guard<std::mutex> lock(mutex());
        store().push_back(std::move(obj));
        return (int) store().size() - 1;
    }
    static std::shared_ptr<T> at(int h) {
        std::lock_guard<std::mutex> lock(mutex());
        if (h < 0 || h >= (int) store().size()) {
            throw CtbError("Cabinet::at", "Index " + std::to_string(h) + " out of range.");
        }
        if (!store()[h]) {
            throw CtbError("Cabinet::at", "Object with index " + std::to_string(h) + " has been deleted.");
        }
        return store()[h];
    }
    static void del(int h) {
        std::lock_guard<std::mutex> lock(mutex());
        if (h < 0 || h >= (int) store().size()) {
            throw CtbError("Cabinet::del", "Index " + std::to_string(h) + " out of range.");
        }
        store()[h].reset();
    }
private:
    static std::vector<std::shared_ptr<T>>& store() { static std::vector<std::shared_ptr<T>> s; return s; }
    static std::mutex& mutex() { static std::mutex m; return m; }
};

using KinCabinet = Cabinet<BatchedGasKinetics>;
using NetCabinet = Cabinet<BatchReactorNet>;
using TransCabinet = Cabinet<BatchedMixTransport>;

//! handleAllExceptions (clib_utils.h:34-56): -1 for library errors, ERR for anything else.
template <class R, class F>
R guarded(R ctErr, R otherErr, F&& f)
{
    try {
        return f();
    } catch (CtbError& e) {
        t_lastError = e.what();
        return ctErr;
    } catch (std::exception& e) {
        t_lastError = e.what();
        return otherErr;
    } catch (...) {
        t_lastError = "unknown exception";
        return otherErr;
    }
}

template <class F>
int32_t guard(F&& f) { return guarded<int32_t>(-1, CTB_ERR, std::forward<F>(f)); }

int32_t copyString(const std::string& s, int32_t bufLen, char* buf)
{
    // string getters return the required length when called with (0, NULL)
    // (test/clib/test_ctkin.cpp:119-123)
    int32_t need = (int32_t) s.size() + 1;
    if (bufLen > 0 && buf) {
        int32_t n = std::min(need, bufLen);
        std::memcpy(buf, s.c_str(), n - 1);
        buf[n - 1] = '\0';
    }
    return need;
}

} // namespace

namespace {
// 8 independent DFMA chains per thread, fully unrolled: the FP64 pipe is the only limiter.
__global__ void __launch_bounds__(256) dfmaPeakKernel(double* out, int iters, double a, double b)
{
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
            x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
        }
    }
    const double r = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (r == 123.456) out[0] = r; // never true: keeps the chains alive
}
// The same question for the FP64 tensor-core path: mma.sync m8n8k4 (DMMA), 8 independent accumulator
// fragments per warp. north_star: "unless ncu shows FP64 DMMA pays for Newton factorisation".
__global__ void __launch_bounds__(256) dmmaPeakKernel(double* out, int iters, double a, double b)
{
    double c[8][2];
#pragma unroll
    for (int u = 0; u < 8; u++) { c[u][0] = threadIdx.x + u; c[u][1] = u; }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int u = 0; u < 8; u++) {
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                             : "+d"(c[u][0]), "+d"(c[u][1]) : "d"(a), "d"(b));
            }
        }
    }
    double r = 0.0;
#pragma unroll
    for (int u = 0; u < 8; u++) r += c[u][0] + c[u][1];
    if (r == 123.456) out[0] = r;
}
} // namespace

extern "C" {

double ctb_fp64DmmaPeakTflops(int32_t device)
{
    return guarded<double>(CTB_DERR, CTB_DERR, [&] {
        cudaCheck(cudaSetDevice(device), "cudaSetDevice");
        int sms = 0;
        cudaCheck(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device), "cudaDeviceGetAttribute");
        double* d = nullptr;
        cudaCheck(cudaMalloc(&d, 64), "cudaMalloc");
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        const int iters = 2048, grid = sms * 8, threads = 256;
        double best = 0.0;
        for (int rep = 0; rep < 4; rep++) {
            cudaEventRecord(e0);
            dmmaPeakKernel<<<grid, threads>>>(d, iters, 1e-3, 1e-3);
            cudaEventRecord(e1);
            cudaCheck(cudaEventSynchronize(e1), "dmmaPeakKernel");
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            // one m8n8k4 = 8 * 8 * 4 multiply-adds per warp
            const double flop = 2.0 * 256.0 * 8 * 4 * (double) iters * (double) grid * (threads / 32);
            if (rep > 0) best = std::max(best, flop / (ms * 1e-3) / 1e12);
        }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        cudaFree(d);
        return best;
    });
}

double ctb_fp64PeakTflops(int32_t device)
{
    return guarded<double>(CTB_DERR, CTB_DERR, [&] {
        cudaCheck(cudaSetDevice(device), "cudaSetDevice");
        int sms = 0;
        cudaCheck(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device), "cudaDeviceGetAttribute");
        double* d = nullptr;
        cudaCheck(cudaMalloc(&d, 64), "cudaMalloc");
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        const int iters = 4096, grid = sms * 8, threads = 256;
        double best = 0.0;
        for (int rep = 0; rep < 4; rep++) {
            cudaEventRecord(e0);
            dfmaPeakKernel<<<grid, threads>>>(d, iters, 0.999999, 1e-9);
            cudaEventRecord(e1);
            cudaCheck(cudaEventSynchronize(e1), "dfmaPeakKernel");
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            const double flop = 2.0 * 8 * 16 * (double) iters * (double) grid * threads;
            if (rep > 0) best = std::max(best, flop / (ms * 1e-3) / 1e12);
        }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        cudaFree(d);
        return best;
    });
}

int32_t ctb_getError(int32_t bufLen, char* buf) { return copyString(t_lastError, bufLen, buf); }

int32_t ctb_nDevices(void)
{
    return guard([&] {
        int n = 0;
        cudaError_t e = cudaGetDeviceCount(&n);
        if (e != cudaSuccess) throw CtbError("ctb_nDevices", cudaGetErrorString(e));
        return (int32_t) n;
    });
}

int32_t ctb_version(int32_t bufLen, char* buf) { return copyString("0.1.0 sm_100a", bufLen, buf); }

int32_t ctb_setError(const char* message)
{
    t_lastError = message ? message : "";
    return 0;
}

// ---- BatchedGasKinetics --------------------------------------------------------------------

int32_t ctb_kin_newFromTables(const ctb_mech_tables* tables, int32_t device)
{
    return guard([&] {
        if (!tables) throw CtbError("ctb_kin_newFromTables", "null tables");
        return (int32_t) KinCabinet::add(std::make_shared<BatchedGasKinetics>(*tables, device));
    });
}

int32_t ctb_kin_newFromFile(const char* ctbmPath, int32_t device)
{
    return guard([&] {
        if (!ctbmPath) throw CtbError("ctb_kin_newFromFile", "null path");
        cantera_b200::MechTables m;
        try {
            m = cantera_b200::MechTables::load(ctbmPath);
        } catch (std::exception& e) {
            throw CtbError("ctb_kin_newFromFile", e.what());
        }
        return (int32_t) KinCabinet::add(std::make_shared<BatchedGasKinetics>(m.view(), device));
    });
}

int32_t ctb_kin_del(int32_t handle)
{
    return guard([&] { KinCabinet::del(handle); return 0; });
}

int32_t ctb_kin_setStream(int32_t handle, void* cudaStream)
{
    return guard([&] { KinCabinet::at(handle)->setStream((cudaStream_t) cudaStream); return 0; });
}

int32_t ctb_kin_synchronize(int32_t handle)
{
    return guard([&] { KinCabinet::at(handle)->synchronize(); return 0; });
}

int32_t ctb_kin_nSpecies(int32_t handle)
{
    return guard([&] { return (int32_t) KinCabinet::at(handle)->nSpecies(); });
}

int32_t ctb_kin_nReactions(int32_t handle)
{
    return guard([&] { return (int32_t) KinCabinet::at(handle)->nReactions(); });
}

int32_t ctb_kin_setMultiplier(int32_t handle, int32_t i, double f)
{
    return guard([&] { KinCabinet::at(handle)->setMultiplier(i, f); return 0; });
}

double ctb_kin_multiplier(int32_t handle, int32_t i)
{
    return guarded<double>(CTB_DERR, CTB_DERR, [&] { return KinCabinet::at(handle)->multiplier(i); });
}

int32_t ctb_kin_setDerivativeSettings(int32_t handle, int32_t skipThirdBodies, int32_t skipFalloff,
                                      double rtolDelta)
{
    return guard([&] {
        KinCabinet::at(handle)->setDerivativeSettings(skipThirdBodies != 0, skipFalloff != 0, rtolDelta);
        return 0;
    });
}

#define CTB_SPECIES_GETTER(NAME, FIELD)                                                          \
    int32_t NAME(int32_t handle, int64_t nStates, const double* T, const double* P,              \
                 const double* Y, int64_t ldY, double* out, int64_t ldW, int32_t memspace)       \
    {                                                                                            \
        return guard([&] {                                                                       \
            if (!out) throw CtbError(#NAME, "null output pointer");                              \
            RateOutputs o;                                                                       \
            o.FIELD = out;                                                                       \
            o.ldW = ldW;                                                                         \
            KinCabinet::at(handle)->rates(nStates, T, P, Y, ldY, o, memspace);                   \
            return 0;                                                                            \
        });                                                                                      \
    }

#define CTB_REACTION_GETTER(NAME, FIELD)                                                         \
    int32_t NAME(int32_t handle, int64_t nStates, const double* T, const double* P,              \
                 const double* Y, int64_t ldY, double* out, int64_t ldR, int32_t memspace)       \
    {                                                                                            \
        return guard([&] {                                                                       \
            if (!out) throw CtbError(#NAME, "null output pointer");                              \
            RateOutputs o;                                                                       \
            o.FIELD = out;                                                                       \
            o.ldR = ldR;                                                                         \
            KinCabinet::at(handle)->rates(nStates, T, P, Y, ldY, o, memspace);                   \
            return 0;                                                                            \
        });                                                                                      \
    }

CTB_SPECIES_GETTER(ctb_kin_getNetProductionRates, wdot)
CTB_SPECIES_GETTER(ctb_kin_getCreationRates, cdot)
CTB_SPECIES_GETTER(ctb_kin_getDestructionRates, ddot)
CTB_REACTION_GETTER(ctb_kin_getFwdRatesOfProgress, ropf)
CTB_REACTION_GETTER(ctb_kin_getRevRatesOfProgress, ropr)
CTB_REACTION_GETTER(ctb_kin_getNetRatesOfProgress, ropnet)
CTB_REACTION_GETTER(ctb_kin_getFwdRateConstants, kf)
CTB_REACTION_GETTER(ctb_kin_getEquilibriumConstants, kc)

int32_t ctb_kin_getQuantity(int32_t handle, int32_t quantity, int64_t nStates, const double* T, const double* P,
                            const double* Y, int64_t ldY, double* out, int64_t ld, int32_t memspace)
{
    return guard([&] {
        KinCabinet::at(handle)->quantity(quantity, nStates, T, P, Y, ldY, out, ld, memspace);
        return 0;
    });
}

int32_t ctb_kin_getNetProductionRates_ddX(int32_t handle, int64_t nStates, const double* T,
                                          const double* P, const double* Y, int64_t ldY, double* jac,
                                          int64_t bufLen, int32_t memspace)
{
    return guard([&] {
        KinCabinet::at(handle)->jacobian(nStates, T, P, Y, ldY, jac, bufLen, false, memspace);
        return 0;
    });
}

int32_t ctb_kin_getNetProductionRates_ddCi(int32_t handle, int64_t nStates, const double* T,
                                           const double* P, const double* Y, int64_t ldY, double* jac,
                                           int64_t bufLen, int32_t memspace)
{
    return guard([&] {
        KinCabinet::at(handle)->jacobian(nStates, T, P, Y, ldY, jac, bufLen, true, memspace);
        return 0;
    });
}

int32_t ctb_kin_getNetProductionRates_ddX_noAlloc(int32_t handle, int64_t nStates, const double* T,
                                                  const double* P, const double* Y, int64_t ldY, double* jac,
                                                  int64_t bufLen)
{
    return guard([&] {
        KinCabinet::at(handle)->jacobianNoAlloc(nStates, T, P, Y, ldY, jac, bufLen);
        return 0;
    });
}

// ---- BatchedMixTransport -------------------------------------------------------------------

int32_t ctb_trans_newFromArrays(int32_t nSpecies, const double* molWeights, const double* viscPoly,
                                const double* condPoly, const double* diffPoly, int32_t device)
{
    return guard([&] {
        return (int32_t) TransCabinet::add(std::make_shared<BatchedMixTransport>(nSpecies, molWeights, viscPoly, condPoly,
                                                                                 diffPoly, device));
    });
}

int32_t ctb_trans_del(int32_t handle)
{
    return guard([&] { TransCabinet::del(handle); return 0; });
}

int32_t ctb_trans_nSpecies(int32_t handle)
{
    return guard([&] { return (int32_t) TransCabinet::at(handle)->nSpecies(); });
}

int32_t ctb_trans_setStream(int32_t handle, void* cudaStream)
{
    return guard([&] { TransCabinet::at(handle)->setStream(static_cast<cudaStream_t>(cudaStream)); return 0; });
}

int32_t ctb_trans_synchronize(int32_t handle)
{
    return guard([&] { TransCabinet::at(handle)->synchronize(); return 0; });
}

int32_t ctb_trans_getMixtureTransport(int32_t handle, int64_t nStates, const double* T, const double* P,
                                      const double* Y, int64_t ldY, double* viscosity, double* conductivity,
                                      double* mixDiff, int64_t ldD, int32_t memspace)
{
    return guard([&] {
        TransCabinet::at(handle)->compute(nStates, T, P, Y, ldY, viscosity, conductivity, mixDiff, ldD, memspace);
        return 0;
    });
}

// ---- BatchReactorNet -----------------------------------------------------------------------

int32_t ctb_reactornet_new(int32_t kinHandle, int64_t nReactors)
{
    return guard([&] {
        return (int32_t) NetCabinet::add(std::make_shared<BatchReactorNet>(KinCabinet::at(kinHandle), nReactors));
    });
}

int32_t ctb_reactornet_newTyped(int32_t kinHandle, int64_t nReactors, int32_t reactorType)
{
    return guard([&] {
        return (int32_t) NetCabinet::add(std::make_shared<BatchReactorNet>(KinCabinet::at(kinHandle), nReactors, reactorType));
    });
}

int32_t ctb_reactornet_nEquations(int32_t handle)
{
    return guard([&] { return (int32_t) NetCabinet::at(handle)->nEquations(); });
}

int32_t ctb_reactornet_del(int32_t handle)
{
    return guard([&] { NetCabinet::del(handle); return 0; });
}

int32_t ctb_reactornet_setState(int32_t handle, const double* T, const double* P, const double* Y,
                                int64_t ldY, int32_t memspace)
{
    return guard([&] { NetCabinet::at(handle)->setState(T, P, Y, ldY, memspace); return 0; });
}

int32_t ctb_reactornet_setRelativeTolerance(int32_t handle, double rtol)
{
    return guard([&] { NetCabinet::at(handle)->setRelativeTolerance(rtol); return 0; });
}

int32_t ctb_reactornet_setAbsoluteTolerance(int32_t handle, double atol)
{
    return guard([&] { NetCabinet::at(handle)->setAbsoluteTolerance(atol); return 0; });
}

int32_t ctb_reactornet_setMaxSteps(int32_t handle, int32_t maxSteps)
{
    return guard([&] { NetCabinet::at(handle)->setMaxSteps(maxSteps); return 0; });
}

int32_t ctb_reactornet_getTolerances(int32_t handle, int32_t len, double* out)
{
    return guard([&] {
        if (len < 4 || !out) throw CtbError("ctb_reactornet_getTolerances", "array size (" + std::to_string(len) + ") is too small; needs 4");
        NetCabinet::at(handle)->getTolerances(out);
        return 0;
    });
}

int32_t ctb_reactornet_setInitialTime(int32_t handle, double t0)
{
    return guard([&] { NetCabinet::at(handle)->setInitialTime(t0); return 0; });
}

int32_t ctb_reactornet_advance(int32_t handle, double t)
{
    return guard([&] { NetCabinet::at(handle)->advance(t); return 0; });
}

int32_t ctb_reactornet_lastRounds(int32_t handle)
{
    return guard([&] { return (int32_t) NetCabinet::at(handle)->lastRounds(); });
}

double ctb_reactornet_time(int32_t handle)
{
    return guarded<double>(CTB_DERR, CTB_DERR, [&] { return NetCabinet::at(handle)->time(); });
}

int32_t ctb_reactornet_getState(int32_t handle, double* T, double* Y, int64_t ldY, int32_t memspace)
{
    return guard([&] { NetCabinet::at(handle)->getState(T, Y, ldY, memspace); return 0; });
}

int32_t ctb_reactornet_getStateVector(int32_t handle, double* y, int64_t bufLen, int32_t memspace)
{
    return guard([&] { NetCabinet::at(handle)->getStateVector(y, bufLen, memspace); return 0; });
}

int32_t ctb_reactornet_eval(int32_t handle, double* ydot, int64_t bufLen, int32_t memspace)
{
    return guard([&] { NetCabinet::at(handle)->eval(ydot, bufLen, memspace); return 0; });
}

int32_t ctb_reactornet_getJacobian(int32_t handle, double* jac, int64_t bufLen, int32_t memspace)
{
    return guard([&] { NetCabinet::at(handle)->jacobian(jac, bufLen, memspace); return 0; });
}

int32_t ctb_reactornet_status(int32_t handle, int64_t len, int32_t* status)
{
    return guard([&] { NetCabinet::at(handle)->status(len, status); return 0; });
}

int32_t ctb_reactornet_solverStats(int32_t handle, int64_t len, int64_t* stats)
{
    return guard([&] { NetCabinet::at(handle)->solverStats(len, (long long*) stats); return 0; });
}

int32_t ctb_reactornet_setIgnitionMonitor(int32_t handle, double dT)
{
    return guard([&] { NetCabinet::at(handle)->setIgnitionMonitor(dT); return 0; });
}

int32_t ctb_reactornet_ignitionDelay(int32_t handle, int64_t len, double* tIgn)
{
    return guard([&] { NetCabinet::at(handle)->ignitionDelay(len, tIgn); return 0; });
}

} // extern "C"
