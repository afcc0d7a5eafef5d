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
//! Launch description of the Jacobian tile kernels; tiles == 0: the mechanism does not fit them.
struct JacTilePlan {
    int tiles = 0, grid = 0;
    size_t smemBytes = 0, tileDoubles = 0, tableBytes = 0, scratchBytes = 0;
};
JacTilePlan planJacTile(const DeviceMech& m, long long nStates);
//! Bytes of global scratch a Jacobian launch may need (per concurrent launch / stream).
size_t jacScratchBytes(const DeviceMech& m);
//! scratch: jacScratchBytes() of device memory owned by this launch until it completes (nullptr:
//! the lane-group kernel of rates.cuh is used).
void launchJac(const DeviceMech& m, const JacArgs& a, cudaStream_t stream, void* scratch);
//! The Jacobian tile kernel on the library's static scratch array (CTB_NOALLOC_SCRATCH_BYTES per GPU): no
//! allocation; m.device must be current. Throws when the mechanism does not fit.
void launchJacStatic(const DeviceMech& m, const JacArgs& a, cudaStream_t stream);
size_t jacStaticScratchBytes();

} // namespace ctb
