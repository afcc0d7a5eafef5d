// derivs_kernels.h — launcher of the derivative-family / reaction-delta kernel (derivs_kernels.cu).
#pragma once

#include "kin_kernels.h"

namespace ctb {

//! Device pointers; any output may be null. Index [0..2] = _ddT, _ddP, _ddC.
//! delta: deltaGibbs, deltaEnthalpy, deltaEntropy, deltaSSGibbs, deltaSSEnthalpy, deltaSSEntropy.
struct DerivArgs {
    long long N;
    const double* T;
    const double* P;
    const double* Y;
    long long ldY;
    double* kf[3];     // [n*ldR + i], reference reaction order
    double* ropf[3];
    double* ropr[3];
    double* ropnet[3];
    double* delta[6];
    long long ldR;
    double* cdot[3];   // [n*ldW + k]
    double* ddot[3];
    double* wdot[3];
    long long ldW;
    double rtolDelta;
    int skipThirdBodies, skipFalloff;
};

void launchDerivs(const DeviceMech& m, const DerivArgs& a, cudaStream_t stream);

} // namespace ctb
