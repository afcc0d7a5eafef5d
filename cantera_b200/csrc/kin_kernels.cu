// kin_kernels.cu — batched BulkKinetics getters as sm_100a kernels.
//
//   ratesKernel<NT, FULL>  setState_TPY + updateROP + getNetProductionRates (+ every other getter
//                          when FULL) for N states; a sub-warp group of NT lanes per state.
//   jacKernel<G>           netProductionRates_ddX / _ddCi, dense column-major K x K per state.
//
// Shared memory: [hot mechanism tables (TMA bulk copy)] [mbarrier] [per-group work areas].
// The kinetics kernels keep the library exp(): expHot (rates.cuh) shortens the dependent chain,
// which pays in the latency-bound reactor kernels (RHS kernel -13 %) but not here, where 13 warps
// per SM already cover it and the extra clamp costs issue slots (151 vs 153 M states/s measured).
#define CTB_EXP_LIB
#include "kin_kernels.h"
#include "rates_thread.cuh"
#include "jac_tile.cuh"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <type_traits>
#include <stdexcept>
#include <string>

namespace ctb {

namespace {

template <class G> __device__ __forceinline__ G makeGroup(double* scratch);
template <> __device__ __forceinline__ SubWarp<32> makeGroup<SubWarp<32>>(double*) { return SubWarp<32>(); }
template <> __device__ __forceinline__ Cta<128> makeGroup<Cta<128>>(double* s) { return Cta<128>(s); }
template <> __device__ __forceinline__ Cta<256> makeGroup<Cta<256>>(double* s) { return Cta<256>(s); }

// ---------------------------------------------------------------------------------------- rates

template <int NT, bool FULL>
__global__ void __launch_bounds__(512) ratesKernel(const MechView v, const unsigned char* __restrict__ gmech,
                                                   const RatesArgs a, const int workDoubles)
{
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + v.hotBytes);
    tmaStage(smem, gmech, v.hotBytes, bar);
    const unsigned char* sm = smem;

    using G = SubWarp<NT>;
    const G g;
    const int groupsPerCta = blockDim.x / NT;
    const int gInCta = threadIdx.x / NT;
    double* work = reinterpret_cast<double*>(smem + v.hotBytes + 16) + (size_t) gInCta * workDoubles;
    const int K = v.K, R = v.R;
    double* c = work;
    double* grt = c + K;
    double* ropf = grt + K;       // FULL: ropf, ropr, rfn ; else: q only
    double* ropr = ropf + R;
    double* rfn = ropr + R;

    const long long totalGroups = (long long) gridDim.x * groupsPerCta;
    const long long gid = (long long) blockIdx.x * groupsPerCta + gInCta;
    const int32_t* perm = CTB_TAB(int32_t, gmech, v.o_perm);

    for (long long base = 0; base < a.N; base += totalGroups) {
        long long n = base + gid;
        const bool valid = n < a.N;
        if (!valid) n = a.N - 1; // idle groups shadow the last state so that the warp stays converged
        GasScalars s;
        stateFromTPY(g, v, sm, a.T[n], a.P[n], a.Y + n * a.ldY, c, s);
        thermoPhase<G, false>(g, v, sm, s, grt, nullptr, nullptr);
        g.sync();
        RopOut o = {};
        if (FULL) {
            o.ropf = ropf; o.ropr = ropr; o.rfn = rfn;
        } else {
            o.q = ropf;
        }
        ropPhase(g, v, sm, s, c, grt, o);
        g.sync();
        if (valid) {
            if (!FULL) {
                double* out = a.wdot + n * a.ldW;
                if (NT == 8) {
                    // balanced deal of the species to the 8 lanes (mech_pack.h o_gl8*)
                    const int32_t* lp = CTB_TAB(int32_t, sm, v.o_gl8Ptr);
                    const uint16_t* ls = CTB_TAB(uint16_t, sm, v.o_gl8Sp);
                    for (int i = lp[g.lane()]; i < lp[g.lane() + 1]; i++) {
                        const int k = ls[i];
                        out[k] = gatherSpecies<0>(v, sm, gmech, k, ropf, nullptr);
                    }
                } else {
                    for (int k = g.lane(); k < K; k += NT) {
                        out[k] = gatherSpecies<0>(v, sm, gmech, k, ropf, nullptr);
                    }
                }
            } else {
                for (int e = g.lane(); e < R; e += NT) {
                    const int i = perm[e];
                    const double f = ropf[e], r = ropr[e];
                    if (a.ropf) a.ropf[n * a.ldR + i] = f;
                    if (a.ropr) a.ropr[n * a.ldR + i] = r;
                    if (a.ropnet) a.ropnet[n * a.ldR + i] = f - r;
                    if (a.kf) a.kf[n * a.ldR + i] = rfn[e];
                    if (a.kc) {
                        // getEquilibriumConstants (BulkKinetics.cpp:163-179): every reaction,
                        // exp(-dG0/RT + dn log c0)
                        union { uint4 u; RxnRec r; } rr;
                        rr.u = CTB_TAB(uint4, sm, v.o_rec)[e];
                        double dg = 0.0 + sideSum(v, sm, rr.r.np, rr.r.p, grt);
                        dg -= sideSum(v, sm, rr.r.nr, rr.r.r, grt);
                        a.kc[n * a.ldR + i] = exp(-dg * s.rrt + CTB_TAB(double, sm, v.o_dn)[e] * s.logStandConc);
                    }
                }
                for (int k = g.lane(); k < K; k += NT) {
                    if (a.cdot) a.cdot[n * a.ldW + k] = gatherSpecies<1>(v, sm, gmech, k, ropf, ropr);
                    if (a.ddot) a.ddot[n * a.ldW + k] = gatherSpecies<2>(v, sm, gmech, k, ropf, ropr);
                }
                if (a.wdot) {
                    // form q in place (ropf <- ropf - ropr) after the creation/destruction gathers
                    g.sync();
                    for (int e = g.lane(); e < R; e += NT) ropf[e] = ropf[e] - ropr[e];
                    g.sync();
                    for (int k = g.lane(); k < K; k += NT) {
                        a.wdot[n * a.ldW + k] = gatherSpecies<0>(v, sm, gmech, k, ropf, nullptr);
                    }
                }
            }
        }
        g.sync();
    }
}


// ------------------------------------------------------------------------------- rates, tiles

//! setState_TPY + updateROP + getNetProductionRates for tiles of 16 states (rates_tile.cuh): W warps
//! per tile, several tiles per CTA, persistent grid. Shared memory:
//!   [tile tables (TMA bulk copy)] [mbarrier] [per tile: c/wdot rows, u rows, v rows, q rows, scratch]
template <int W, int MAXT, int ILP>
__global__ void __launch_bounds__(MAXT) ratesTileKernel(const TileView tv, const unsigned char* __restrict__ gblob,
                                                        const RatesArgs a, const int tileDoubles, const int tableBytes)
{
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + tableBytes);
    tmaStage(smem, gblob + tv.begin, tv.bytes, bar);
    constexpr int TT = W * 32, NL = 2 * W;
    const int ti = threadIdx.x / TT, tid = threadIdx.x % TT;
    const int tilesPerCta = blockDim.x / TT;
    const int s = tid & (CTB_TS - 1), j = tid / CTB_TS;
    const int K = tv.K, R = tv.R;
    double* c = reinterpret_cast<double*>(smem + tableBytes + 128) + (size_t) ti * tileDoubles;
    double* u = c + K * CTB_LDC;
    double* v = u + K * CTB_LDC;
    double* q = v + K * CTB_LDC;
    double* red = q; // [2][NL][16] scratch of the normalisation sums: dead before the q rows are written
    Tile t;
    t.tp = smem;
    t.cS = c + s; t.uS = u + s; t.vS = v + s; t.qS = q + s;
    if (tid < CTB_TS) q[R * CTB_TS + tid] = 0.0;
    volatile int* safeFlag = reinterpret_cast<volatile int*>(smem + tableBytes + 16) + ti;
    if (tid == 0) *safeFlag = 0;
    const int barId = 1 + ti;
    const long long nTiles = (a.N + CTB_TS - 1) / CTB_TS;
    for (long long tile = (long long) blockIdx.x * tilesPerCta + ti; tile < nTiles; tile += (long long) gridDim.x * tilesPerCta) {
        const long long n0 = tile * CTB_TS;
        const int nvalid = (int) min((long long) CTB_TS, a.N - n0);
        // ---- coalesced transposed load of the tile's mass fractions into the c rows
        for (int i = tid; i < CTB_TS * K; i += TT) {
            const int si = (int) __umulhi((unsigned) i, tv.magicK);
            const int k = i - si * K;
            const int sn = min(si, nvalid - 1);
            c[k * CTB_LDC + si] = a.Y[(n0 + sn) * a.ldY + k];
        }
        const long long n = n0 + min(s, nvalid - 1);
        const double Tn = a.T[n], Pn = a.P[n];
        tileSync(barId, TT);
        TileScalars sc;
        double dens;
        kinStateTile<NL, TT>(tv, t, red, barId, s, j, Tn, Pn, sc, dens);
        // ---- concentrations, u = exp(-g'), v = c exp(g')
        double unused = 0.0;
        if (speciesPhaseTile<NL, false>(tv, t, sc, 1.0, dens, j, unused)) *safeFlag = 1;
        tileSync(barId, TT);
        const bool safe = *safeFlag != 0;
        if (safe) {
            speciesPhaseSafeTile<NL>(tv, t, sc, j);
            tileSync(barId, TT);
        }
        // the scratch may have covered the zero row of q (mechanisms with fewer than 2 NL reactions)
        if (2 * NL > R && tid < CTB_TS) q[R * CTB_TS + tid] = 0.0;
        ropPhaseTile<NL, false, ILP>(tv, t, sc, j, safe);
        tileSync(barId, TT);
        if (tid == 0) *safeFlag = 0;
        gatherPhaseTile<NL>(tv, t, j, [&](int k, double w) { t.cS[k * CTB_LDC] = w; });
        tileSync(barId, TT);
        // ---- coalesced transposed store of the wdot rows
        for (int i = tid; i < CTB_TS * K; i += TT) {
            const int si = (int) __umulhi((unsigned) i, tv.magicK);
            const int k = i - si * K;
            if (si < nvalid) a.wdot[(n0 + si) * a.ldW + k] = c[k * CTB_LDC + si];
        }
        tileSync(barId, TT);
    }
}

// ---------------------------------------------------------------------------- jacobian, tiles

//! netProductionRates_ddX / _ddCi for tiles of 16 states (jac_tile.cuh), dense column-major K x K per
//! state. W = 4 warps per tile (8 reaction lanes). Shared memory:
//!   [tile + Jacobian tables (TMA)] [mbarrier, flags] [per tile: c, u, w rows | q rows / 8 column buffers]
template <int W, int MAXT>
__global__ void __launch_bounds__(MAXT) jacTileKernel(const TileView tv, const unsigned char* __restrict__ gblob,
                                                      const JacArgs a, double* __restrict__ scratch, const int tileDoubles,
                                                      const int tableBytes)
{
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + tableBytes);
    tmaStage(smem, gblob + tv.begin, tv.bytesJac, bar);
    constexpr int TT = W * 32, NL = 2 * W;
    const int ti = threadIdx.x / TT, tid = threadIdx.x % TT;
    const int tilesPerCta = blockDim.x / TT;
    const int s = tid & (CTB_TS - 1), j = tid / CTB_TS;
    const int K = tv.K, R = tv.R;
    double* c = reinterpret_cast<double*>(smem + tableBytes + 128) + (size_t) ti * tileDoubles;
    double* u = c + K * CTB_LDC;
    double* w = u + K * CTB_LDC;
    double* q = w + K * CTB_LDC;          // q rows, later the NL private column buffers [K][17]
    double* red = q;
    Tile t;
    t.tp = smem;
    t.cS = c + s; t.uS = u + s; t.vS = w + s; t.qS = q + s;
    double* colBuf = q + (size_t) j * K * CTB_LDC;
    double* colS = colBuf + s;
    JacRows jr;
    {
        double* g = scratch + ((size_t) blockIdx.x * tilesPerCta + ti) * jacScratchDoubles(R, tv.nTb) + s;
        jr.kfM = g; jr.krM = g + (size_t) R * CTB_TS; jr.h2 = jr.krM + (size_t) R * CTB_TS;
        jr.we = jr.h2 + (size_t) (R + 1) * CTB_TS;
        if (j == 0) jr.h2[R * CTB_TS] = 0.0; // the all-zero row of the padded gather entries
    }
    volatile int* safeFlag = reinterpret_cast<volatile int*>(smem + tableBytes + 16) + ti;
    if (tid == 0) *safeFlag = 0;
    const int barId = 1 + ti;
    const bool skip3b = a.skipThirdBodies != 0, skipFalloff = a.skipFalloff != 0;
    const unsigned halfMask = 0xFFFFu << (16 * ((threadIdx.x >> 4) & 1));
    const int l16 = threadIdx.x & 15;
    const long long nTiles = (a.N + CTB_TS - 1) / CTB_TS;
    for (long long tile = (long long) blockIdx.x * tilesPerCta + ti; tile < nTiles; tile += (long long) gridDim.x * tilesPerCta) {
        const long long n0 = tile * CTB_TS;
        const int nvalid = (int) min((long long) CTB_TS, a.N - n0);
        for (int i = tid; i < CTB_TS * K; i += TT) {
            const int si = (int) __umulhi((unsigned) i, tv.magicK);
            const int k = i - si * K;
            const int sn = min(si, nvalid - 1);
            c[k * CTB_LDC + si] = a.Y[(n0 + sn) * a.ldY + k];
        }
        const long long n = n0 + min(s, nvalid - 1);
        const double Tn = a.T[n], Pn = a.P[n];
        tileSync(barId, TT);
        TileScalars sc;
        double dens;
        kinStateTile<NL, TT>(tv, t, red, barId, s, j, Tn, Pn, sc, dens);
        double unused = 0.0;
        if (speciesPhaseTile<NL, false, true>(tv, t, sc, 1.0, dens, j, unused)) *safeFlag = 1;
        tileSync(barId, TT);
        const bool safe = *safeFlag != 0;
        if (safe) {
            speciesPhaseSafeTile<NL>(tv, t, sc, j);
            tileSync(barId, TT);
        }
        if (tid < CTB_TS) q[R * CTB_TS + tid] = 0.0; // the column buffers of the previous tile covered it
        for (int e = j; e < R; e += NL) ropJacTile<false>(tv, t, sc, e, safe, skip3b, skipFalloff, 0.0, jr);
        tileSync(barId, TT);
        if (tid == 0) *safeFlag = 0;
        // dense default-efficiency part D_k = sum_e nu_ke we_e default_e into the (dead) w rows
        gatherPhase2Tile<NL>(tv, t, jr.h2, j, [&](int k, double wdot, double D) { t.vS[k * CTB_LDC] = D; });
        tileSync(barId, TT); // q rows are dead: the column buffers take their place
        const double scale = a.ddCi ? 1.0 : sc.ctot; // BulkKinetics.cpp:671-676
        jacColumnsTile<NL>(tv, t, jr, colS, j, skip3b, [&](int jc) {
            for (int k = 0; k < K; k++) colS[k * CTB_LDC] = (colS[k * CTB_LDC] + t.vS[k * CTB_LDC]) * scale;
            __syncwarp(halfMask);
            // transposed write: for every state of the tile, lanes over the rows k of column jc
            for (int ss = 0; ss < nvalid; ss++) {
                double* dst = a.jac + (n0 + ss) * (long long) K * K + (long long) jc * K;
                for (int k = l16; k < K; k += 16) dst[k] = colBuf[k * CTB_LDC + ss];
            }
            __syncwarp(halfMask);
        });
        tileSync(barId, TT);
    }
}

// --------------------------------------------------------------------------- rates, thread/state

//! setState_TPY + updateROP + getNetProductionRates, one THREAD per state (rates_thread.cuh).
//! Shared memory: [hot mechanism tables][mbarrier][c, g, w : 3 x K x st doubles], st = blockDim+1.
template <int U>
__global__ void __launch_bounds__(256, 1) ratesThreadKernel(const MechView v, const unsigned char* __restrict__ gmech,
                                                            const RatesArgs a, const int st)
{
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + v.hotBytes);
    tmaStage(smem, gmech, v.hotBytes, bar);
    const unsigned char* sm = smem;
    const int K = v.K;
    double* cAll = reinterpret_cast<double*>(smem + v.hotBytes + 16);
    double* gAll = cAll + (size_t) K * st;
    double* wAll = gAll + (size_t) K * st;
    TArr t;
    t.c = cAll + threadIdx.x; t.g = gAll + threadIdx.x; t.w = wAll + threadIdx.x; t.st = st;
    const int NTH = blockDim.x;
    for (long long base = (long long) blockIdx.x * NTH; base < a.N; base += (long long) gridDim.x * NTH) {
        const int nb = (int) min((long long) NTH, a.N - base);
        loadTransposed(cAll, st, a.Y + base * a.ldY, a.ldY, nb, K);
        const bool valid = (int) threadIdx.x < nb;
        double T = 1000.0, P = 101325.0;
        if (valid) {
            T = a.T[base + threadIdx.x];
            P = a.P[base + threadIdx.x];
        }
        __syncthreads();
        if (!valid) {
            for (int k = 0; k < K; k++) t.c[k * st] = 1.0; // idle threads work on a benign state
        }
        GasScalars s;
        double dens;
        stateFromTPY_T(v, sm, T, P, t, s, dens);
        thermoT(v, sm, s, t, dens);
        reactionsT<false, U>(v, sm, s, t);
        __syncthreads();
        storeTransposed(a.wdot + base * a.ldW, a.ldW, wAll, st, nb, K);
        __syncthreads();
    }
}


// ------------------------------------------------------------------------------------- jacobian

//! One group (a warp, or the whole CTA when CTA_GROUP) per state.
template <int NT, bool CTA_GROUP>
__global__ void __launch_bounds__(CTA_GROUP ? NT : 256) jacKernel(const MechView v,
                                                                 const unsigned char* __restrict__ gmech,
                                                                 const JacArgs a, const int workDoubles)
{
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + v.hotBytes);
    tmaStage(smem, gmech, v.hotBytes, bar);
    const unsigned char* sm = smem;
    double* scratch = reinterpret_cast<double*>(smem + v.hotBytes + 16); // 16 doubles of Cta scratch

    using G = typename std::conditional<CTA_GROUP, Cta<NT>, SubWarp<32>>::type;
    const int groupsPerCta = CTA_GROUP ? 1 : blockDim.x / 32;
    const int gInCta = CTA_GROUP ? 0 : threadIdx.x / 32;
    G g = makeGroup<G>(scratch);
    double* work = scratch + 16 + (size_t) gInCta * workDoubles;
    const int K = v.K, R = v.R;
    double* c = work;
    double* grt = c + K;
    double* rfn = grt + K;
    double* rkc = rfn + R;
    double* concm = rkc + R;
    double* dk = concm + R;
    double* w = dk + R;
    double* D = w + R;       // [K + R]
    double* J = D + K + R;   // [K * K]

    const long long totalGroups = (long long) gridDim.x * groupsPerCta;
    const long long gid = (long long) blockIdx.x * groupsPerCta + gInCta;
    for (long long base = 0; base < a.N; base += totalGroups) {
        long long n = base + gid;
        const bool valid = n < a.N;
        if (!valid) n = a.N - 1;
        GasScalars s;
        stateFromTPY(g, v, sm, a.T[n], a.P[n], a.Y + n * a.ldY, c, s);
        thermoPhase<G, false>(g, v, sm, s, grt, nullptr, nullptr);
        g.sync();
        RopOut o = {};
        o.rfn = rfn; o.rkc = rkc; o.concm = concm; o.dlnkdM = dk;
        ropPhase(g, v, sm, s, c, grt, o);
        g.sync();
        jacobianCi(g, v, sm, gmech, c, rfn, rkc, concm, dk, w, D, J, K, a.ddCi ? 1.0 : s.ctot,
                   a.skipThirdBodies != 0, a.skipFalloff != 0);
        if (valid) {
            double* out = a.jac + n * (long long) K * K;
            for (int i = g.lane(); i < K * K; i += G::NT) out[i] = J[i];
        }
        g.sync();
    }
}

} // namespace

// ------------------------------------------------------------------------------------ launchers

template <int NT, bool FULL>
static void launchRatesT(const DeviceMech& m, const RatesArgs& a, cudaStream_t stream)
{
    const MechView& v = m.view;
    const int workDoubles = 2 * v.K + (FULL ? 3 : 1) * v.R + 2; // +2: keep groups 16-byte aligned
    const int workAligned = (workDoubles + 1) & ~1;
    const size_t fixed = v.hotBytes + 16;
    const size_t budget = (size_t) m.smemOptIn - 1024;
    if (fixed + (size_t) workAligned * 8 > budget) {
        throw std::runtime_error("mechanism too large for the shared-memory rates kernel");
    }
    int groups = (int) ((budget - fixed) / ((size_t) workAligned * 8));
    int threads = std::min(512, groups * NT);
    threads = (threads / 32) * 32;
    if (threads < 32) {
        throw std::runtime_error("mechanism too large for the shared-memory rates kernel");
    }
    groups = threads / NT;
    const size_t smemBytes = fixed + (size_t) groups * workAligned * 8;
    auto kern = ratesKernel<NT, FULL>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smemBytes);
    int ctasPerSm = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctasPerSm, kern, threads, smemBytes);
    if (ctasPerSm < 1) ctasPerSm = 1;
    long long want = (a.N + groups - 1) / groups;
    int grid = (int) std::min<long long>(want, (long long) m.numSMs * ctasPerSm);
    if (grid < 1) grid = 1;
    kern<<<grid, threads, smemBytes, stream>>>(v, m.dBlob, a, workAligned);
}

//! Bytes of shared memory one tile needs with NL reaction lanes.
static size_t tileDoublesFor(const TileView& tv, int NL)
{
    size_t d = 3 * (size_t) tv.K * CTB_LDC + std::max((size_t) (tv.R + 1) * CTB_TS, 2 * (size_t) NL * CTB_TS);
    return (d + 15) & ~(size_t) 15;
}

template <int W, int MAXT, int ILP>
static bool launchRatesTileT(const DeviceMech& m, const RatesArgs& a, cudaStream_t stream)
{
    const TileView& tv = m.tile;
    const size_t tableBytes = ((size_t) tv.bytes + 127) & ~(size_t) 127;
    const size_t tileDoubles = tileDoublesFor(tv, 2 * W);
    const size_t fixed = tableBytes + 128;
    const size_t budget = (size_t) m.smemOptIn; // no static shared memory in this kernel
    if (fixed + tileDoubles * 8 > budget) return false;
    int tiles = (int) ((budget - fixed) / (tileDoubles * 8));
    tiles = std::min(tiles, std::min(15, MAXT / (W * 32)));
    if (tiles < 1) return false;
    const long long nTiles = (a.N + CTB_TS - 1) / CTB_TS;
    tiles = (int) std::min<long long>(tiles, std::max<long long>(1, (nTiles + m.numSMs - 1) / m.numSMs));
    const size_t smemBytes = fixed + (size_t) tiles * tileDoubles * 8;
    auto kern = ratesTileKernel<W, MAXT, ILP>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smemBytes);
    const int grid = (int) std::max<long long>(1, std::min<long long>((nTiles + tiles - 1) / tiles, m.numSMs));
    kern<<<grid, tiles * W * 32, smemBytes, stream>>>(tv, m.dBlob, a, (int) tileDoubles, (int) tableBytes);
    return true;
}

//! The wdot-only path: tiles of 16 states, 8 warps (16 reaction lanes) per tile. Measured on gri30
//! (profiles/r02_wdot_tile_sweep.txt): 4 warps per tile 252 M states/s, 8 warps 266 M; two reactions in
//! flight per lane change nothing at 8 warps.
static bool launchRatesTile(const DeviceMech& m, const RatesArgs& a, cudaStream_t stream)
{
    return launchRatesTileT<8, 768, 1>(m, a, stream);
}

//! Threads per CTA of the thread-per-state kernels for this mechanism (0: does not fit).
int threadKernelThreads(const DeviceMech& m, int arraysPerThread)
{
    const MechView& v = m.view;
    const long long avail = (long long) m.smemOptIn - 1024 - v.hotBytes - 16;
    long long nth = avail / ((long long) arraysPerThread * v.K * 8) - 1;
    nth = std::min<long long>(256, (nth / 32) * 32);
    return nth >= 64 ? (int) nth : 0;
}

template <int U>
static void launchRatesThreadT(const DeviceMech& m, const RatesArgs& a, int nth, cudaStream_t stream)
{
    const MechView& v = m.view;
    const int st = nth + 1;
    const size_t smemBytes = v.hotBytes + 16 + (size_t) 3 * v.K * st * 8;
    auto kern = ratesThreadKernel<U>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smemBytes);
    const long long want = (a.N + nth - 1) / nth;
    const int grid = (int) std::max<long long>(1, std::min<long long>(want, m.numSMs));
    kern<<<grid, nth, smemBytes, stream>>>(v, m.dBlob, a, st);
}

void launchRates(const DeviceMech& m, const RatesArgs& a, bool full, cudaStream_t stream)
{
    if (a.N <= 0) return;
    // Default: a sub-warp group of 8 lanes per state. CTB_RATES_KERNEL=thread selects the
    // thread-per-state kernel (rates_thread.cuh) for wdot-only calls when the mechanism leaves room
    // for >= 64 states per SM in shared memory.
    if (!full) {
        const char* force = std::getenv("CTB_RATES_KERNEL");
        if (!(force && std::strcmp(force, "group") == 0) && !(force && std::strcmp(force, "thread") == 0)
            && launchRatesTile(m, a, stream)) {
            return;
        }
        const int nth = threadKernelThreads(m, 3);
        if (nth > 0 && force && std::strcmp(force, "thread") == 0) {
            launchRatesThreadT<4>(m, a, nth, stream);
            return;
        }
    }
    // lanes per state: enough states in flight per SM vs. idle lanes in the species loops
    if (full) {
        launchRatesT<8, true>(m, a, stream);
    } else {
        launchRatesT<8, false>(m, a, stream);
    }
}

template <int NT, bool CTA_GROUP>
static bool launchJacT(const DeviceMech& m, const JacArgs& a, cudaStream_t stream)
{
    const MechView& v = m.view;
    int workDoubles = 3 * v.K + 6 * v.R + v.K * v.K + 2;
    workDoubles = (workDoubles + 1) & ~1;
    const size_t fixed = v.hotBytes + 16 + 16 * 8;
    const size_t budget = (size_t) m.smemOptIn - 1024;
    if (fixed + (size_t) workDoubles * 8 > budget) return false;
    int groups = CTA_GROUP ? 1 : (int) std::min<size_t>(8, (budget - fixed) / ((size_t) workDoubles * 8));
    const int threads = CTA_GROUP ? NT : groups * 32;
    const size_t smemBytes = fixed + (size_t) groups * workDoubles * 8;
    auto kern = jacKernel<NT, CTA_GROUP>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smemBytes);
    int ctasPerSm = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctasPerSm, kern, threads, smemBytes);
    if (ctasPerSm < 1) ctasPerSm = 1;
    long long want = (a.N + groups - 1) / groups;
    int grid = (int) std::max<long long>(1, std::min<long long>(want, (long long) m.numSMs * ctasPerSm));
    kern<<<grid, threads, smemBytes, stream>>>(v, m.dBlob, a, workDoubles);
    return true;
}

constexpr int JAC_TILE_W = 4, JAC_TILE_MAXT = 512;

JacTilePlan planJacTile(const DeviceMech& m, long long nStates)
{
    JacTilePlan p;
    const TileView& tv = m.tile;
    constexpr int NL = 2 * JAC_TILE_W;
    p.tableBytes = ((size_t) tv.bytesJac + 127) & ~(size_t) 127;
    const size_t span = std::max(std::max((size_t) (tv.R + 1) * CTB_TS, (size_t) NL * tv.K * CTB_LDC), 2 * (size_t) NL * CTB_TS);
    p.tileDoubles = (3 * (size_t) tv.K * CTB_LDC + span + 15) & ~(size_t) 15;
    const size_t fixed = p.tableBytes + 128;
    const size_t budget = (size_t) m.smemOptIn;
    if (fixed + p.tileDoubles * 8 > budget) return p;
    int tiles = (int) ((budget - fixed) / (p.tileDoubles * 8));
    tiles = std::min(tiles, std::min(15, JAC_TILE_MAXT / (JAC_TILE_W * 32)));
    const long long nTiles = (nStates + CTB_TS - 1) / CTB_TS;
    tiles = (int) std::min<long long>(tiles, std::max<long long>(1, (nTiles + m.numSMs - 1) / m.numSMs));
    p.tiles = tiles;
    p.smemBytes = fixed + (size_t) tiles * p.tileDoubles * 8;
    p.grid = (int) std::max<long long>(1, std::min<long long>((nTiles + tiles - 1) / tiles, m.numSMs));
    p.scratchBytes = (size_t) p.grid * tiles * jacScratchDoubles(tv.R, tv.nTb) * 8;
    return p;
}

size_t jacScratchBytes(const DeviceMech& m)
{
    return (size_t) m.numSMs * 4 * jacScratchDoubles(m.tile.R, m.tile.nTb) * 8;
}

void launchJac(const DeviceMech& m, const JacArgs& a, cudaStream_t stream, void* scratch)
{
    if (a.N <= 0) return;
    const MechView& v = m.view;
    static const bool group = [] { const char* e = std::getenv("CTB_JAC_KERNEL"); return e && std::strcmp(e, "group") == 0; }();
    if (scratch && !group) {
        const JacTilePlan tp = planJacTile(m, a.N);
        if (tp.tiles > 0 && tp.scratchBytes <= jacScratchBytes(m)) {
            auto kern = jacTileKernel<JAC_TILE_W, JAC_TILE_MAXT>;
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) tp.smemBytes);
            kern<<<tp.grid, tp.tiles * JAC_TILE_W * 32, tp.smemBytes, stream>>>(m.tile, m.dBlob, a, static_cast<double*>(scratch),
                                                                               (int) tp.tileDoubles, (int) tp.tableBytes);
            return;
        }
    }
    // small Jacobians: a warp per state, several per CTA; large ones: the CTA per state
    const size_t perWarp = ((size_t) 3 * v.K + 6 * v.R + (size_t) v.K * v.K + 2) * 8;
    const size_t budget = (size_t) m.smemOptIn - 1024 - v.hotBytes - 160;
    bool ok;
    if (perWarp * 3 <= budget) {
        ok = launchJacT<32, false>(m, a, stream);
    } else {
        ok = launchJacT<256, true>(m, a, stream);
    }
    if (!ok) throw std::runtime_error("mechanism too large for the shared-memory Jacobian kernel");
}

// ---------------------------------------------------------------- jacobian, allocation-free launch

// Scratch of launchJacStatic: one file-scope array per GPU that loads the library, held until exit
// (include/ctb_clib.h, ctb_kin_getNetProductionRates_ddX_noAlloc). In host code the NAME below is only a
// stand-in for the symbol; the device address comes from cudaGetSymbolAddress (jacStaticScratch()).
static __device__ __align__(256) unsigned char g_jacStaticScratch[CTB_NOALLOC_SCRATCH_BYTES];

size_t jacStaticScratchBytes()
{
    return sizeof(g_jacStaticScratch);
}

//! Device address of g_jacStaticScratch on `device`, which must be the current device; resolved once
//! per device.
static double* jacStaticScratch(int device)
{
    constexpr int MAXDEV = 64;
    static std::mutex mu;
    static void* addr[MAXDEV] = {};
    if (device < 0 || device >= MAXDEV) throw std::runtime_error("device index beyond the static-scratch table");
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[device]) {
        int cur = -1;
        cudaError_t e = cudaGetDevice(&cur);
        if (e != cudaSuccess || cur != device) {
            throw std::runtime_error("static scratch set-up: device " + std::to_string(device) + " is not current");
        }
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_jacStaticScratch);
        if (e != cudaSuccess || !p) {
            throw std::runtime_error(std::string("cudaGetSymbolAddress(static Jacobian scratch): ") + cudaGetErrorString(e));
        }
        addr[device] = p;
    }
    return static_cast<double*>(addr[device]);
}

void launchJacStatic(const DeviceMech& m, const JacArgs& a, cudaStream_t stream)
{
    if (a.N <= 0) return;
    const TileView& tv = m.tile;
    const JacTilePlan full = planJacTile(m, a.N);
    if (full.tiles == 0) throw std::runtime_error("the mechanism does not fit the Jacobian tile kernel");
    const size_t perSlot = jacScratchDoubles(tv.R, tv.nTb) * 8; // one tile of 16 states in flight
    const size_t slots = sizeof(g_jacStaticScratch) / perSlot;
    if (slots == 0) {
        throw std::runtime_error("the mechanism needs " + std::to_string(perSlot) + " bytes of scratch per tile; the static "
                                 "array holds " + std::to_string(sizeof(g_jacStaticScratch)));
    }
    // as many tiles in flight as the array holds: the kernel strides over the tiles, so a smaller grid
    // only lowers the parallelism
    const long long nTiles = (a.N + CTB_TS - 1) / CTB_TS;
    int tiles = 1, grid = 1;
    long long best = 0;
    for (int t = 1; t <= full.tiles; t++) {
        const long long g = std::min<long long>(std::min<long long>((nTiles + t - 1) / t, m.numSMs), (long long) (slots / t));
        if (g >= 1 && g * t > best) { best = g * t; tiles = t; grid = (int) g; }
    }
    const size_t smemBytes = full.tableBytes + 128 + (size_t) tiles * full.tileDoubles * 8;
    double* scratch = jacStaticScratch(m.device);
    auto kern = jacTileKernel<JAC_TILE_W, JAC_TILE_MAXT>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smemBytes);
    kern<<<grid, tiles * JAC_TILE_W * 32, smemBytes, stream>>>(m.tile, m.dBlob, a, scratch, (int) full.tileDoubles,
                                                               (int) full.tableBytes);
}

} // namespace ctb
