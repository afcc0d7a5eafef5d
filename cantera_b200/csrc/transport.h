// transport.h — BatchedMixTransport: the reference's mixture-averaged transport model (MixTransport,
// the step on the other side of the chemistry substep in a reacting-flow solver) for N
// independent states per call. The temperature fits (GasTransport's polynomials) are inputs: they come from
// the reference's public getters (GasTransport.h:107-115), exactly as the mechanism tables do.
#pragma once

#include "engine.h"

namespace ctb {

struct TransportArgs {
    long long N;
    const double* T;
    const double* P;
    const double* Y;
    long long ldY;
    double* visc;  // [n] or null     Transport::viscosity
    double* cond;  // [n] or null     Transport::thermalConductivity
    double* dmix;  // [n*ldD + k] or null   Transport::getMixDiffCoeffs
    long long ldD;
};

class BatchedMixTransport {
public:
    //! mw[K]; viscPoly, condPoly [K*5]; diffPoly [K(K+1)/2 * 5] in the order of GasTransport::m_diffcoeffs
    //! (pairs i <= j, row by row: GasTransport.cpp:778-795).
    BatchedMixTransport(int K, const double* mw, const double* viscPoly, const double* condPoly, const double* diffPoly,
                        int device);
    ~BatchedMixTransport();
    BatchedMixTransport(const BatchedMixTransport&) = delete;

    int nSpecies() const { return m_K; }
    void setStream(cudaStream_t s) { m_userStream = s; m_hasUserStream = true; }
    cudaStream_t stream() const { return m_hasUserStream ? m_userStream : m_stream; }
    void synchronize();
    //! any subset of the three outputs (null = not wanted)
    void compute(long long N, const double* T, const double* P, const double* Y, long long ldY, double* visc, double* cond,
                 double* dmix, long long ldD, int memspace);

private:
    int m_K, m_device, m_numSMs = 0;
    double* m_tab = nullptr; // device: [rmw K][mw K][visc 5K][cond 5K][diff 5 npair][wq K*K][wk1 K*K]
    cudaStream_t m_stream = nullptr, m_userStream = nullptr;
    bool m_hasUserStream = false;
    DevBuf m_stage;
};

} // namespace ctb
