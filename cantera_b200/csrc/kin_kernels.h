// kin_kernels.h — host-visible launchers of the batched kinetics kernels.
#pragma once

#include "mech_pack.h"

#include <cuda_runtime.h>

namespace ctb {

//! A mechanism resident on one device.
struct DeviceMech {
    MechView view;
    TileView tile;                  // tables of the state-tile kernels (rates_tile.cuh)
    unsigned char* dBlob = nullptr; // device copy of PackedMech::blob
    int device = 0;
    int numSMs = 0;
    int smemOptIn = 0; // cudaDevAttrMaxSharedMemoryPerBlockOptin
};

//! Batch layout of the C ABI (include/ctb_clib.h): T[n], P[n], Y[n*ldY + k]; device pointers.
struct RatesArgs {
    long long N;
    const double* T;
    const double* P;
    const double* Y;
    long long ldY;
    double* wdot; // [n*ldW + k], may be null in FULL mode
    double* cdot;
    double* ddot;
    long long ldW;
    double* ropf; // [n*ldR + i] in the reference's reaction order
    double* ropr;
    double* ropnet;
    double* kf;
    double* kc;
    long long ldR;
};

struct JacArgs {
    long long N;
    const double* T;
    const double* P;
    const double* Y;
    long long ldY;
    double* jac; // [n*K*K + j*K + k] = d wdot_k / d X_j (or d c_j)
    int ddCi;    // 0: netProductionRates_ddX, 1: netProductionRates_ddCi
    int skipThirdBodies;
    int skipFalloff;
};

//! full = false: wdot only (fast path); full = true: any subset of the outputs.
void launchRates(const DeviceMech& m, const RatesArgs& a, bool full, cudaStream_t stream);
void launchJac(const DeviceMech& m, const JacArgs& a, cudaStream_t stream);

} // namespace ctb
