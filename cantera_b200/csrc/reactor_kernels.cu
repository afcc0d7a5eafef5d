// reactor_kernels.cu — kernels of BatchReactorNet.
//
//   (advanceKernel<NT>, the persistent single-kernel integrator, lives in reactor_advance.cu)
//   reactorInitKernel   getState of a fresh reactor from (T, P, Y)
//   reactorContentsKernel (T, Y) of the contents from the solution vector
#include "bdf.cuh"

#include <algorithm>
#include <cstdlib>
#include <stdexcept>

namespace ctb {

long long reactorHistDoubles(int neq) { return 6LL * neq + BDF_HIST_SCALARS; }

// reactor_advance.cu, one translation unit per group width
bool launchAdvance32(const DeviceMech& m, const ReactorBatch& b, int maxThreads, cudaStream_t stream);
bool launchAdvance64(const DeviceMech& m, const ReactorBatch& b, int maxThreads, cudaStream_t stream);
bool launchAdvance128(const DeviceMech& m, const ReactorBatch& b, int maxThreads, cudaStream_t stream);
bool launchAdvance256(const DeviceMech& m, const ReactorBatch& b, int maxThreads, cudaStream_t stream);

namespace {

// ReactorNet::eval (ReactorNet.cpp:617-635) for every reactor at its current solution vector:
// a warp per reactor.
__global__ void __launch_bounds__(256) evalKernel(const MechView v, const unsigned char* __restrict__ gmech,
                                                  long long N, const double* __restrict__ y,
                                                  const double* __restrict__ P, double* __restrict__ ydot,
                                                  const int groupDoubles)
{
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + v.hotBytes);
    tmaStage(smem, gmech, v.hotBytes, bar);
    using G = NamedGroup<32>;
    const int gInCta = threadIdx.x / 32;
    double* base = reinterpret_cast<double*>(smem + v.hotBytes + 16) + (size_t) gInCta * groupDoubles;
    const G g(base, 1 + gInCta);
    const int K = v.K, n = K + 1;
    ReactorWork wk = {};
    wk.n = n;
    wk.y = base; wk.f = wk.y + n; wk.c = wk.f + n; wk.grt = wk.c + K; wk.hrt = wk.grt + K; wk.cpr = wk.hrt + K;
    wk.q = wk.cpr + K;
    const long long stride = (long long) gridDim.x * (blockDim.x / 32);
    for (long long r = (long long) blockIdx.x * (blockDim.x / 32) + gInCta; r < N; r += stride) {
        for (int i = g.lane(); i < n; i += 32) wk.y[i] = y[r * n + i];
        g.sync();
        RhsAux aux;
        reactorRhs<G, false>(g, v, smem, gmech, wk, P[r], wk.y, wk.f, aux);
        for (int i = g.lane(); i < n; i += 32) ydot[r * n + i] = wk.f[i];
        g.sync();
    }
}

// One thread per reactor: Phase::setMassFractions (Phase.cpp:342-360) + setState_TP, then getState of
// the reactor model with m_vol = 1 m^3 (layouts: reactor.h).
__global__ void reactorInitKernel(const MechView v, const unsigned char* __restrict__ gmech, const int type, long long N,
                                  const double* __restrict__ T, const double* __restrict__ P,
                                  const double* __restrict__ Y, long long ldY, double* __restrict__ y,
                                  double* __restrict__ Pout, double* __restrict__ T0)
{
    const long long r = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= N) return;
    const int K = v.K;
    const double* rmolwt = CTB_TAB(double, gmech, v.o_rmolwt);
    const double* Yr = Y + r * ldY;
    double sum = 0.0;
    for (int k = 0; k < K; k++) sum += fmax(Yr[k], 0.0);
    const double invSum = 1.0 / sum;
    double sym = 0.0;
    for (int k = 0; k < K; k++) sym += (fmax(Yr[k], 0.0) * invSum) * rmolwt[k];
    const double mmw = 1.0 / sym;
    const double dens = P[r] * mmw / (T[r] * CTB_GAS_CONSTANT);
    const double mass = dens * 1.0;
    const int sIdx = reactorSIdx(type);
    double* yr = y + r * (K + sIdx);
    yr[reactorTIdx(type)] = T[r];
    if (type == 1) yr[1] = 1.0;                       // volume
    if (type == 2) { yr[0] = mass; yr[1] = 1.0; }
    if (type == 3) yr[0] = mass;
    for (int k = 0; k < K; k++) {
        const double yk = fmax(Yr[k], 0.0) * invSum;
        yr[sIdx + k] = reactorIsMole(type) ? mass * rmolwt[k] * yk : yk;
    }
    Pout[r] = P[r];
    T0[r] = T[r];
}

__global__ void reactorContentsKernel(const MechView v, const unsigned char* __restrict__ gmech, const int type, long long N,
                                      const double* __restrict__ y, double* __restrict__ T,
                                      double* __restrict__ Y, long long ldY)
{
    const long long r = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= N) return;
    const int K = v.K, sIdx = reactorSIdx(type);
    const double* molwt = CTB_TAB(double, gmech, v.o_molwt);
    const double* yr = y + r * (K + sIdx);
    if (reactorIsMole(type)) {
        double totalMass = 0.0;
        for (int k = 0; k < K; k++) totalMass += yr[sIdx + k] * molwt[k];
        const double f = 1.0 / totalMass;
        for (int k = 0; k < K; k++) Y[r * ldY + k] = (yr[sIdx + k] * molwt[k]) * f;
    } else {
        // ThermoPhase::massFractions after setMassFractions_NoNorm: the values as integrated
        for (int k = 0; k < K; k++) Y[r * ldY + k] = yr[sIdx + k];
    }
    T[r] = yr[reactorTIdx(type)];
}

} // namespace

void launchAdvance(const DeviceMech& m, const ReactorBatch& b, cudaStream_t stream)
{
    if (b.N <= 0) return;
    const MechView& v = m.view;
    // lanes per reactor: the number of reactors resident on an SM is bounded by shared memory
    // (one Newton matrix each), so the group is widened until ~512 threads are busy
    const size_t groupBytes = (16 + reactorWorkDoubles(v.K, v.R)) * 8;
    const size_t budget = (size_t) m.smemOptIn - 1024 - v.hotBytes - 16;
    const int fit = (int) std::max<size_t>(1, budget / groupBytes);
    int nt = 32;
    while (nt < 256 && fit * nt * 2 <= 512) nt *= 2;
    if (const char* e = std::getenv("CTB_BDF_NT")) nt = std::atoi(e);
    bool ok;
    switch (nt) {
    case 32: ok = launchAdvance32(m, b, 512, stream); break;
    case 64: ok = launchAdvance64(m, b, 512, stream); break;
    case 128: ok = launchAdvance128(m, b, 512, stream); break;
    case 256: ok = launchAdvance256(m, b, 512, stream); break;
    default: throw std::runtime_error("CTB_BDF_NT must be 32, 64, 128 or 256");
    }
    if (!ok) throw std::runtime_error("mechanism too large for the shared-memory reactor kernel");
}

void launchReactorEval(const DeviceMech& m, long long N, const double* y, const double* P, double* ydot,
                       cudaStream_t stream)
{
    if (N <= 0) return;
    const MechView& v = m.view;
    const int groupDoubles = (2 * (v.K + 1) + 4 * v.K + v.R + 1) & ~1;
    const int groups = 8;
    const size_t smemBytes = v.hotBytes + 16 + (size_t) groups * groupDoubles * 8;
    if (smemBytes > (size_t) m.smemOptIn - 1024) throw std::runtime_error("mechanism too large for the eval kernel");
    cudaFuncSetAttribute(evalKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smemBytes);
    const int grid = (int) std::max<long long>(1, std::min<long long>((N + groups - 1) / groups, 2LL * m.numSMs));
    evalKernel<<<grid, groups * 32, smemBytes, stream>>>(v, m.dBlob, N, y, P, ydot, groupDoubles);
}

void launchReactorInit(const DeviceMech& m, int type, long long N, const double* T, const double* P, const double* Y,
                       long long ldY, double* y, double* Pout, double* T0, cudaStream_t stream)
{
    if (N <= 0) return;
    const int threads = 128;
    reactorInitKernel<<<(unsigned) ((N + threads - 1) / threads), threads, 0, stream>>>(m.view, m.dBlob, type, N, T, P, Y,
                                                                                      ldY, y, Pout, T0);
}

void launchReactorContents(const DeviceMech& m, int type, long long N, const double* y, double* T, double* Y, long long ldY,
                           cudaStream_t stream)
{
    if (N <= 0) return;
    const int threads = 128;
    reactorContentsKernel<<<(unsigned) ((N + threads - 1) / threads), threads, 0, stream>>>(m.view, m.dBlob, type, N, y, T,
                                                                                          Y, ldY);
}

} // namespace ctb
