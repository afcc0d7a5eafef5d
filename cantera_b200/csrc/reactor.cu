// reactor.cu — BatchReactorNet host side: device-resident reactor states, staging, launches.
#include "reactor.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <limits>

namespace ctb {

BatchReactorNet::BatchReactorNet(std::shared_ptr<BatchedGasKinetics> kin, long long nReactors, int reactorType)
    : m_kin(std::move(kin)), m_N(nReactors), m_type(reactorType)
{
    if (!m_kin) throw CtbError("BatchReactorNet", "null kinetics object");
    if (nReactors < 1) throw CtbError("BatchReactorNet", "the number of reactors must be positive");
    if (reactorType < 0 || reactorType > 3) throw CtbError("BatchReactorNet", "unknown reactor type " + std::to_string(reactorType));
    m_neq = reactorNeq(m_type, m_kin->nSpecies());
    cudaCheck(cudaSetDevice(m_kin->device()), "cudaSetDevice");
    const size_t N = (size_t) m_N;
    m_y.reserve(sizeof(double) * N * m_neq);
    m_P.reserve(sizeof(double) * N);
    m_T0.reserve(sizeof(double) * N);
    m_hist.reserve(sizeof(double) * N * (size_t) reactorHistDoubles(m_neq));
    m_tIgn.reserve(sizeof(double) * N);
    m_status.reserve(sizeof(int32_t) * N);
    m_stats.reserve(sizeof(long long) * N * CTB_NSTATS);
    m_counter.reserve(sizeof(unsigned long long));
    cudaCheck(cudaMemsetAsync(m_status.p, 0, sizeof(int32_t) * N, m_kin->stream()), "cudaMemset");
    cudaCheck(cudaMemsetAsync(m_stats.p, 0, sizeof(long long) * N * CTB_NSTATS, m_kin->stream()), "cudaMemset");
    cudaCheck(cudaMemsetAsync(m_tIgn.p, 0xFF, sizeof(double) * N, m_kin->stream()), "cudaMemset"); // NaN
}

BatchReactorNet::~BatchReactorNet()
{
    cudaSetDevice(m_kin->device());
    cudaStreamSynchronize(m_kin->stream());
}

void BatchReactorNet::setState(const double* T, const double* P, const double* Y, long long ldY, int memspace)
{
    const int K = m_kin->nSpecies();
    if (!T || !P || !Y) throw CtbError("BatchReactorNet::setState", "null state pointer");
    if (ldY < K) throw CtbError("BatchReactorNet::setState", "ldY is smaller than the number of species");
    cudaCheck(cudaSetDevice(m_kin->device()), "cudaSetDevice");
    cudaStream_t st = m_kin->stream();
    const double *dT = T, *dP = P, *dY = Y;
    long long ld = ldY;
    if (memspace == CTB_HOST) {
        m_io.reserve(sizeof(double) * (size_t) m_N * (K + 2));
        double* d = m_io.as<double>();
        cudaCheck(cudaMemcpyAsync(d, T, sizeof(double) * m_N, cudaMemcpyHostToDevice, st), "H2D T");
        cudaCheck(cudaMemcpyAsync(d + m_N, P, sizeof(double) * m_N, cudaMemcpyHostToDevice, st), "H2D P");
        cudaCheck(cudaMemcpy2DAsync(d + 2 * m_N, sizeof(double) * K, Y, sizeof(double) * ldY, sizeof(double) * K,
                                    m_N, cudaMemcpyHostToDevice, st), "H2D Y");
        dT = d; dP = d + m_N; dY = d + 2 * m_N; ld = K;
    } else if (memspace != CTB_DEVICE) {
        throw CtbError("BatchReactorNet::setState", "memspace must be CTB_HOST or CTB_DEVICE");
    }
    launchReactorInit(m_kin->mech(), m_type, m_N, dT, dP, dY, ld, m_y.as<double>(), m_P.as<double>(),
                      m_T0.as<double>(), st);
    cudaCheck(cudaGetLastError(), "reactorInit launch");
    cudaCheck(cudaMemsetAsync(m_status.p, 0, sizeof(int32_t) * m_N, st), "cudaMemset");
    cudaCheck(cudaMemsetAsync(m_stats.p, 0, sizeof(long long) * m_N * CTB_NSTATS, st), "cudaMemset");
    cudaCheck(cudaMemsetAsync(m_tIgn.p, 0xFF, sizeof(double) * m_N, st), "cudaMemset");
    if (memspace == CTB_HOST) cudaCheck(cudaStreamSynchronize(st), "cudaStreamSynchronize");
    m_haveState = true;
    m_time = 0.0;   // ReactorNet::setInitialTime(0)
    m_restart = true;
}

void BatchReactorNet::setRelativeTolerance(double rtol)
{
    if (!(rtol > 0.0)) {
        throw CtbError("BatchReactorNet::setRelativeTolerance",
                       "Relative tolerance must be positive; got " + std::to_string(rtol) + ".");
    }
    m_rtol = rtol;
    m_restart = true; // CVodesIntegrator::setTolerances takes effect at the next (re)initialisation
}

void BatchReactorNet::setAbsoluteTolerance(double atol)
{
    if (!(atol > 0.0)) {
        throw CtbError("BatchReactorNet::setAbsoluteTolerance",
                       "Absolute tolerance must be positive; got " + std::to_string(atol) + ".");
    }
    m_atol = atol;
    m_atolUser = true;
    m_restart = true;
}

// Tolerance mapping. A user-set absolute tolerance is applied exactly as given, to every component
// (ReactorNet.cpp:200-209, 233-245). The DEFAULT (1e-15, scaled by the reactor's total moles for the
// species rows: MoleReactor.cpp:64-76) is tightened by kDefaultAtolSafety: this solver converges its
// Newton iterations with an analytic Jacobian and never cuts a step for a convergence failure, so at
// equal tolerances it takes ~25 % fewer, larger steps than CVODES does through Cantera, and the
// radical pool of a long induction period (components far below rtol |y|, controlled by atol alone)
// is tracked less tightly than the reference happens to track it. The factor restores the
// reference's accuracy level on the ignition delay; tests/test_gpu_reactor.py checks it on reactors
// it was not chosen on, and ctb_reactornet_getTolerances reports what is actually applied.
static constexpr double kDefaultAtolSafety = 0.1;

double BatchReactorNet::effectiveAtol() const
{
    return m_atolUser ? m_atol : m_atol * kDefaultAtolSafety;
}

void BatchReactorNet::getTolerances(double* out4) const
{
    out4[0] = m_rtol;
    out4[1] = effectiveAtol();              // temperature row
    out4[2] = effectiveAtol();              // species rows; times the reactor's total moles if out4[3] == 0
    out4[3] = m_atolUser ? 1.0 : 0.0;
}

void BatchReactorNet::setMaxSteps(int n)
{
    if (n < 1) throw CtbError("BatchReactorNet::setMaxSteps", "the number of steps must be positive");
    m_maxSteps = n;
}

void BatchReactorNet::setInitialTime(double t0)
{
    m_time = t0;
    m_restart = true;
}

void BatchReactorNet::setIgnitionMonitor(double dT)
{
    m_ignDT = dT;
}

void BatchReactorNet::advance(double t)
{
    if (!m_haveState) throw CtbError("BatchReactorNet::advance", "setState has not been called");
    if (t == m_time) return; // CVodesIntegrator::integrate (CVodesIntegrator.cpp:481-483)
    if (t < m_time) {
        throw CtbError("BatchReactorNet::advance", "Cannot integrate backwards in time.\nRequested time "
                       + std::to_string(t) + " < current time " + std::to_string(m_time));
    }
    cudaCheck(cudaSetDevice(m_kin->device()), "cudaSetDevice");
    cudaStream_t st = m_kin->stream();
    cudaCheck(cudaMemsetAsync(m_counter.p, 0, sizeof(unsigned long long), st), "cudaMemset");
    ReactorBatch b = {};
    b.N = m_N; b.neq = m_neq;
    b.type = m_type; b.tIdx = reactorTIdx(m_type); b.sIdx = reactorSIdx(m_type);
    b.y = m_y.as<double>(); b.P = m_P.as<double>(); b.T0 = m_T0.as<double>();
    b.hist = m_hist.as<double>(); b.histStride = reactorHistDoubles(m_neq);
    b.tIgn = m_tIgn.as<double>(); b.status = m_status.as<int32_t>(); b.stats = m_stats.as<long long>();
    b.counter = m_counter.as<unsigned long long>();
    b.t0 = m_time; b.tout = t; b.rtol = m_rtol; b.atol = m_atol; b.atolUser = m_atolUser;
    b.maxSteps = m_maxSteps; b.restart = m_restart; b.ignDT = m_ignDT; b.crateFloor = 0.0;
    b.atol = effectiveAtol();
    // The batch-synchronous integrator (reactor_batch.cu) handles up to 128 equations per reactor;
    // larger mechanisms take the one-kernel shared-memory integrator (reactor_kernels.cu).
    const long long cap = batchSlotCap(m_kin->mech(), m_neq);
    if (cap <= 0 && m_type != 0) {
        throw CtbError("BatchReactorNet::advance", "this reactor model needs the batch integrator (at most 128 equations)");
    }
    if (cap > 0) {
        const long long S = std::min(m_N, cap);
        if (S != m_slots) {
            // models without an analytic Jacobian: room for the n + 1 right-hand sides of every slot's difference quotients
            const size_t extra = m_type == 0 ? reactorJacScratchBytes(m_kin->mech())
                                             : sizeof(double) * (size_t) S * (m_neq + 1) * m_neq;
            m_ws.reserve(batchWorkspaceBytes(m_neq, S, nullptr, extra));
            m_slots = S;
        }
        try {
            m_lastRounds = launchAdvanceBatch(m_kin->mech(), b, m_ws.p, S, st);
        } catch (const std::runtime_error& e) {
            throw CtbError("BatchReactorNet::advance", e.what());
        }
        cudaCheck(cudaGetLastError(), "batch advance");
    } else {
        m_lastRounds = 0;
        launchAdvance(m_kin->mech(), b, st); // one-kernel integrator: IdealGasConstPressureMoleReactor only
        cudaCheck(cudaGetLastError(), "advance launch");
    }
    m_time = t;
    m_restart = false;
}

void BatchReactorNet::getState(double* T, double* Y, long long ldY, int memspace)
{
    const int K = m_kin->nSpecies();
    if (!m_haveState) throw CtbError("BatchReactorNet::getState", "setState has not been called");
    if (!T || !Y) throw CtbError("BatchReactorNet::getState", "null output pointer");
    if (ldY < K) throw CtbError("BatchReactorNet::getState", "ldY is smaller than the number of species");
    cudaCheck(cudaSetDevice(m_kin->device()), "cudaSetDevice");
    cudaStream_t st = m_kin->stream();
    if (memspace == CTB_DEVICE) {
        launchReactorContents(m_kin->mech(), m_type, m_N, m_y.as<double>(), T, Y, ldY, st);
        cudaCheck(cudaGetLastError(), "reactorContents launch");
        return;
    }
    if (memspace != CTB_HOST) throw CtbError("BatchReactorNet::getState", "memspace must be CTB_HOST or CTB_DEVICE");
    m_io.reserve(sizeof(double) * (size_t) m_N * (K + 2));
    double* d = m_io.as<double>();
    launchReactorContents(m_kin->mech(), m_type, m_N, m_y.as<double>(), d, d + m_N, K, st);
    cudaCheck(cudaGetLastError(), "reactorContents launch");
    cudaCheck(cudaMemcpyAsync(T, d, sizeof(double) * m_N, cudaMemcpyDeviceToHost, st), "D2H T");
    cudaCheck(cudaMemcpy2DAsync(Y, sizeof(double) * ldY, d + m_N, sizeof(double) * K, sizeof(double) * K, m_N,
                                cudaMemcpyDeviceToHost, st), "D2H Y");
    cudaCheck(cudaStreamSynchronize(st), "cudaStreamSynchronize");
}

void BatchReactorNet::getStateVector(double* y, long long bufLen, int memspace)
{
    if (!m_haveState) throw CtbError("BatchReactorNet::getStateVector", "setState has not been called");
    const long long need = m_N * m_neq;
    if (!y || bufLen < need) throw CtbError("BatchReactorNet::getStateVector", "array size (" + std::to_string(bufLen)
                                            + ") is too small; needs " + std::to_string(need));
    cudaCheck(cudaSetDevice(m_kin->device()), "cudaSetDevice");
    cudaStream_t st = m_kin->stream();
    cudaCheck(cudaMemcpyAsync(y, m_y.p, sizeof(double) * need,
                              memspace == CTB_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st), "copy y");
    if (memspace != CTB_DEVICE) cudaCheck(cudaStreamSynchronize(st), "cudaStreamSynchronize");
}

void BatchReactorNet::eval(double* ydot, long long bufLen, int memspace)
{
    if (!m_haveState) throw CtbError("BatchReactorNet::eval", "setState has not been called");
    const long long need = m_N * m_neq;
    if (!ydot || bufLen < need) throw CtbError("BatchReactorNet::eval", "array size (" + std::to_string(bufLen)
                                               + ") is too small; needs " + std::to_string(need));
    cudaCheck(cudaSetDevice(m_kin->device()), "cudaSetDevice");
    cudaStream_t st = m_kin->stream();
    double* d = ydot;
    if (memspace != CTB_DEVICE) {
        m_io.reserve(sizeof(double) * (size_t) need);
        d = m_io.as<double>();
    }
    // the kernel the integrator itself uses (tile form) wherever the mechanism fits it
    const bool groupRhs = std::getenv("CTB_RHS_KERNEL") && std::string(std::getenv("CTB_RHS_KERNEL")) == "group";
    if (groupRhs || !launchReactorEvalTile(m_kin->mech(), m_type, m_N, m_y.as<double>(), m_P.as<double>(), d, st)) {
        if (m_type != 0) throw CtbError("BatchReactorNet::eval", "the mechanism is too large for the tile kernel of this reactor model");
        launchReactorEval(m_kin->mech(), m_N, m_y.as<double>(), m_P.as<double>(), d, st);
    }
    cudaCheck(cudaGetLastError(), "eval launch");
    if (memspace != CTB_DEVICE) {
        cudaCheck(cudaMemcpyAsync(ydot, d, sizeof(double) * need, cudaMemcpyDeviceToHost, st), "D2H ydot");
        cudaCheck(cudaStreamSynchronize(st), "cudaStreamSynchronize");
    }
}

void BatchReactorNet::jacobian(double* jac, long long bufLen, int memspace)
{
    if (!m_haveState) throw CtbError("BatchReactorNet::jacobian", "setState has not been called");
    if (m_type != 0) throw CtbError("BatchReactorNet::jacobian", "only IdealGasConstPressureMoleReactor exports its Jacobian");
    const long long per = (long long) m_neq * m_neq, need = m_N * per;
    if (!jac || bufLen < need) throw CtbError("BatchReactorNet::jacobian", "array size (" + std::to_string(bufLen)
                                              + ") is too small; needs " + std::to_string(need));
    if (memspace != CTB_DEVICE && memspace != CTB_HOST) throw CtbError("BatchReactorNet::jacobian", "memspace must be CTB_HOST or CTB_DEVICE");
    cudaCheck(cudaSetDevice(m_kin->device()), "cudaSetDevice");
    cudaStream_t st = m_kin->stream();
    DevBuf scratch;
    scratch.reserve(reactorJacScratchBytes(m_kin->mech()));
    // host memory: chunks of reactors through the staging buffer
    const long long chunk = memspace == CTB_DEVICE ? m_N : std::max<long long>(16, std::min<long long>(m_N, (256LL << 20) / (8 * per)));
    if (memspace == CTB_HOST) m_io.reserve(sizeof(double) * (size_t) chunk * per);
    for (long long n0 = 0; n0 < m_N; n0 += chunk) {
        const long long n = std::min(chunk, m_N - n0);
        double* d = memspace == CTB_DEVICE ? jac + n0 * per : m_io.as<double>();
        if (!launchReactorJacobianTile(m_kin->mech(), n, m_y.as<double>() + n0 * m_neq, m_P.as<double>() + n0, d, scratch.p, st)) {
            throw CtbError("BatchReactorNet::jacobian", "the mechanism is too large for the Jacobian tile kernel");
        }
        cudaCheck(cudaGetLastError(), "jacobian launch");
        if (memspace == CTB_HOST) {
            cudaCheck(cudaMemcpyAsync(jac + n0 * per, d, sizeof(double) * n * per, cudaMemcpyDeviceToHost, st), "D2H jac");
            cudaCheck(cudaStreamSynchronize(st), "cudaStreamSynchronize");
        }
    }
    cudaCheck(cudaStreamSynchronize(st), "cudaStreamSynchronize"); // the scratch goes away with this frame
}

void BatchReactorNet::status(long long len, int32_t* out)
{
    if (len < m_N || !out) throw CtbError("BatchReactorNet::status", "array size (" + std::to_string(len)
                                          + ") is too small; needs " + std::to_string(m_N));
    cudaCheck(cudaSetDevice(m_kin->device()), "cudaSetDevice");
    cudaCheck(cudaMemcpyAsync(out, m_status.p, sizeof(int32_t) * m_N, cudaMemcpyDeviceToHost, m_kin->stream()), "D2H");
    cudaCheck(cudaStreamSynchronize(m_kin->stream()), "cudaStreamSynchronize");
}

void BatchReactorNet::solverStats(long long len, long long* out)
{
    if (len < m_N * CTB_NSTATS || !out) {
        throw CtbError("BatchReactorNet::solverStats", "array size (" + std::to_string(len)
                       + ") is too small; needs " + std::to_string(m_N * CTB_NSTATS));
    }
    cudaCheck(cudaSetDevice(m_kin->device()), "cudaSetDevice");
    cudaCheck(cudaMemcpyAsync(out, m_stats.p, sizeof(long long) * m_N * CTB_NSTATS, cudaMemcpyDeviceToHost,
                              m_kin->stream()), "D2H");
    cudaCheck(cudaStreamSynchronize(m_kin->stream()), "cudaStreamSynchronize");
}

void BatchReactorNet::ignitionDelay(long long len, double* out)
{
    if (len < m_N || !out) throw CtbError("BatchReactorNet::ignitionDelay", "array size (" + std::to_string(len)
                                          + ") is too small; needs " + std::to_string(m_N));
    cudaCheck(cudaSetDevice(m_kin->device()), "cudaSetDevice");
    cudaCheck(cudaMemcpyAsync(out, m_tIgn.p, sizeof(double) * m_N, cudaMemcpyDeviceToHost, m_kin->stream()), "D2H");
    cudaCheck(cudaStreamSynchronize(m_kin->stream()), "cudaStreamSynchronize");
}

} // namespace ctb
