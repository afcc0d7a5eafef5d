// reactor_batch.cu — the batch-synchronous integrator of BatchReactorNet (SURVEY.md §8 rows
// a12-a14): rounds of four kernels over a pool of SLOTS,
//
//   batchControlKernel  one warp per slot: the resumable BDF/Newton controller of bdf_batch.cuh;
//                       [controller | vectors] of the slot come in and go out by TMA bulk copies,
//                       the Newton solve is a matrix-vector product with the slot's explicit
//                       inverse streamed from HBM/L2 (coalesced, every element used once)
//   batchRhsKernel      a sub-warp group of 8 lanes per slot: reactor right-hand side at the
//                       requested state (reactorRhs of bdf.cuh on the rates device code)
//   batchJacKernel      one cooperative group per FLAGGED slot: analytic J = d(ydot)/dy assembled in
//                       shared memory, written to the slot's matrix region
//   batchInvertKernel   one CTA per FLAGGED slot: (I - gamma J)^-1 by Gauss-Jordan elimination with
//                       partial pivoting, every thread holding (a segment of) one matrix ROW IN
//                       REGISTERS; only the pivot column and the pivot row travel through shared memory
//
// with the per-slot integrator memory in HBM. Tens of thousands of reactors are in flight per
// round, so the chemistry runs at the throughput of the rates kernels.
// The control kernel works on the slot's controller block where the TMA copy put it (shared memory) instead of
// copying all of BatchCtl into registers and back on every visit: no spills, ~120 fewer loads / stores per visit
// (control 2171 -> 2020 ms per 131 072 gri30 reactors). Every lane executes the same scalar updates in lockstep.
#define CTB_CTL_IN_SMEM 1
#include "bdf_batch.cuh"
#include "jac_tile.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>

namespace ctb {

namespace {

__global__ void batchInitKernel(SlotMem mem)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < mem.S) {
        mem.ctl(s)->pc = PC_NEW;
        mem.ctl(s)->sing = 0;
        mem.pc[s] = PC_NEW;
    }
    if (s == 0) {
        *mem.nActive = mem.S;
        mem.setupCount[0] = 0;
        mem.setupCount[1] = 0;
    }
}

//! After an abnormal end of the round loop (safety net; every reactor is already bounded by
//! maxSteps): reactors still in a slot or in the queue get CTB_STATUS_TOO_MANY_STEPS and keep their
//! last accepted state, so that a numerical failure never fails the call.
__global__ void batchAbortKernel(ReactorBatch b, SlotMem mem)
{
    const long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < mem.S && mem.pc[i] != PC_IDLE) {
        const BatchCtl* c = mem.ctl((int) i);
        if (mem.pc[i] != PC_NEW) {
            const long long r = c->r;
            const double* zn0 = mem.vec((int) i);
            for (int k = 0; k < b.neq; k++) b.y[r * b.neq + k] = zn0[k];
            b.status[r] = CTB_STATUS_TOO_MANY_STEPS;
        }
    }
    const long long head = (long long) *b.counter;
    const long long r = head + i;
    if (r < b.N) b.status[r] = CTB_STATUS_TOO_MANY_STEPS;
}

// ---- warp-level TMA helpers (bulkLoad and barWait live in bdf_batch.cuh: the Newton solve uses them too)
__device__ __forceinline__ void bulkStore(void* dstGlobal, const void* srcSmem, uint32_t bytes)
{
    const uint32_t srcAddr = (uint32_t) __cvta_generic_to_shared(srcSmem);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dstGlobal), "r"(srcAddr), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
__device__ __forceinline__ void prefetchL2(const void* srcGlobal, uint32_t bytes)
{
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(srcGlobal), "r"(bytes) : "memory");
}

#ifndef CTB_CTL_WARPS
#define CTB_CTL_WARPS 16
#endif

//! Persistent CTAs, one warp per slot at a time. The warp brings the slot's [controller | vectors]
//! prefix (4.8 KB for gri30) into its own shared-memory buffer with ONE TMA bulk copy, asks the L2
//! for the slot's Newton-matrix inverse at the same time (bulk prefetch), runs the controller
//! (BatchBdf::run: vector arithmetic on shared memory, the Newton solve as a streamed matrix-vector
//! product) and writes the prefix back with a bulk store. The other warps of the CTA cover the
//! latencies; the matrix never occupies shared memory, so 16 warps per SM are resident.
template <int NPL>
__global__ void __launch_bounds__(CTB_CTL_WARPS * 32, 1) batchControlKernel(const ReactorBatch b, const SlotMem mem, const int ringCols)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, W = blockDim.x >> 5;
    const uint32_t prefix = mem.oM; // ctl + vec
    unsigned char* buf = smem + (size_t) warp * prefix;
    // [W prefixes][W x (1 + NBUF) mbarriers][W matrix rings]
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + (size_t) W * prefix) + (size_t) warp * (1 + MatRing::NBUF);
    MatRing ring;
    ring.chunkCols = ringCols;
    ring.chunkBytes = (uint32_t) ringCols * (uint32_t) mem.n * 8u;
    ring.nChunks = ringCols > 0 ? (mem.n + ringCols - 1) / ringCols : 0;
    ring.lastBytes = ringCols > 0 ? (((uint32_t) (mem.n - (ring.nChunks - 1) * ringCols) * (uint32_t) mem.n * 8u + 15u) & ~15u) : 0u;
    ring.bar = bar + 1;
    ring.buf = reinterpret_cast<double*>(smem + (((size_t) W * prefix + (size_t) W * (1 + MatRing::NBUF) * 8 + 127) & ~(size_t) 127)
                                         + (size_t) warp * MatRing::NBUF * ring.chunkBytes);
    ring.parity = 0;
    ring.src = nullptr;
    if (lane == 0) {
        for (int i = 0; i < 1 + MatRing::NBUF; i++) {
            const uint32_t barAddr = (uint32_t) __cvta_generic_to_shared(bar + i);
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(barAddr));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) mem.setupCount[mem.parity ^ 1] = 0; // next round's list
    __syncwarp();
    const SubWarp<32> g;
    uint32_t parity = 0;
    const uint32_t matBytes = ((uint32_t) (mem.n * mem.n) * 8u + 15u) & ~15u;
    for (int s = blockIdx.x * W + warp; s < mem.S; s += gridDim.x * W) {
        const int pc = mem.pc[s];
        if (pc == PC_IDLE) continue;
        unsigned char* gblk = mem.block(s);
        const long long t0 = mem.prof ? clock64() : 0;
        // the visit resumes in the Newton solve: its matrix-vector product needs the slot's inverse
        const bool useRing = ringCols > 0 && pc == PC_NEWTON;
        if (lane == 0) {
            bulkLoad(buf, gblk, prefix, bar);
            if (useRing) ring.prime(gblk + mem.oM);
            else if (pc == PC_NEWTON) prefetchL2(gblk + mem.oM, matBytes);
        }
        barWait(bar, parity);
        parity ^= 1;
        const long long t1 = mem.prof ? clock64() : 0;
        {
            BatchBdf<SubWarp<32>, NPL> bdf(g, b, mem, s, buf, mem.M(s), nullptr, useRing ? &ring : nullptr);
            bdf.run();
        }
        const long long t2 = mem.prof ? clock64() : 0;
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) bulkStore(gblk, buf, prefix);
        __syncwarp();
        if (mem.prof && lane == 0) {
            const long long t3 = clock64();
            atomicAdd(reinterpret_cast<unsigned long long*>(mem.prof + 8), (unsigned long long) (t1 - t0));
            atomicAdd(reinterpret_cast<unsigned long long*>(mem.prof + 9), (unsigned long long) (t2 - t1));
            atomicAdd(reinterpret_cast<unsigned long long*>(mem.prof + 10), (unsigned long long) (t3 - t2));
            atomicAdd(reinterpret_cast<unsigned long long*>(mem.prof + 11), 1ULL);
        }
    }
}

//! Reactor right-hand side at the requested state of every active slot: a sub-warp group of 8
//! lanes per slot (reactorRhs of bdf.cuh on the rates device code), work areas in shared memory.
__global__ void __launch_bounds__(512) batchRhsKernel(const MechView v, const unsigned char* __restrict__ gmech,
                                                      const SlotMem mem, const int groupDoubles)
{
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + v.hotBytes);
    tmaStage(smem, gmech, v.hotBytes, bar);
    using G = SubWarp<8>;
    const G g;
    const int groupsPerCta = blockDim.x / 8;
    const int gInCta = threadIdx.x / 8;
    double* base = reinterpret_cast<double*>(smem + v.hotBytes + 16) + (size_t) gInCta * groupDoubles;
    const int K = v.K, n = K + 1;
    ReactorWork wk = {};
    wk.n = n;
    wk.y = base; wk.f = wk.y + n; wk.c = wk.f + n; wk.grt = wk.c + K; wk.hrt = wk.grt + K; wk.cpr = wk.hrt + K;
    wk.q = wk.cpr + K;
    const int total = gridDim.x * groupsPerCta;
    for (int base0 = blockIdx.x * groupsPerCta; base0 < mem.S; base0 += total) {
        const int s0 = base0 + gInCta;
        const int s = min(s0, mem.S - 1);
        const bool valid = s0 < mem.S && mem.pc[s] != PC_IDLE;
        // the four groups of a warp shuffle with a full mask: they skip or work together
        if (!__any_sync(0xffffffffu, valid)) continue;
        const double* ys = mem.vec(s) + 8 * n;
        for (int i = g.lane(); i < n; i += 8) wk.y[i] = valid ? ys[i] : 1.0;
        const double P = valid ? mem.ctl(s)->P : 101325.0;
        g.sync();
        RhsAux aux;
        reactorRhs<G, false>(g, v, smem, gmech, wk, P, wk.y, wk.f, aux);
        if (valid) {
            double* fs = mem.vec(s) + 9 * n;
            for (int i = g.lane(); i < n; i += 8) fs[i] = wk.f[i];
        }
        g.sync();
    }
}

//! The same right-hand side for TILES of 16 slots (rates_tile.cuh): W warps per tile, a half-warp is the
//! 16 slots of the tile working on one reaction / species. IdealGasConstPressureMoleReactor::updateState
//! + eval (IdealGasConstPressureMoleReactor.cpp:38-106) as in reactorRhs of bdf.cuh.
//! Shared memory: [tile tables (TMA)] [mbarrier, flags] [per tile: c rows, u rows, v rows, q rows, scratch].
//! Increment of column `col` in the forward-difference Jacobian of the reactor models without an analytic
//! one. The right-hand side is a low-degree polynomial in every mole number / mass fraction apart from
//! the third-body and mean-molecular-weight factors, so a LARGE increment is accurate (relative error
//! ~ increment / total) and keeps the rounding error of the quotient small where a species is absent:
//! sqrt(eps) |y| or 1e-7 of the row's natural scale (total moles; 1 for mass fractions), whichever is larger.
__device__ __forceinline__ double fdIncrement(int type, int n, const double* ys, int col)
{
    const int sIdx = reactorSIdx(type);
    double scale;
    if (col < sIdx) {
        scale = fabs(ys[col]);                         // T, V, m rows
    } else if (reactorIsMole(type)) {
        scale = 0.0;
        for (int k = sIdx; k < n; k++) scale += fabs(ys[k]);
    } else {
        scale = 1.0;
    }
    return fmax(1.4901161193847656e-08 * fabs(ys[col]), 1e-7 * scale);
}

//! Where the tile kernel finds its states: the integrator's slots ...
struct SlotIO {
    SlotMem mem;
    __device__ __forceinline__ long long count() const { return mem.S; }
    __device__ __forceinline__ int type() const { return mem.rtype; }
    __device__ __forceinline__ bool active(long long i) const { return mem.pc[i] != PC_IDLE; }
    __device__ __forceinline__ double yv(long long i, int e) const { return (mem.vec((int) i) + 8 * mem.n)[e]; }
    __device__ __forceinline__ double* f(long long i) const { return mem.vec((int) i) + 9 * mem.n; }
    __device__ __forceinline__ double P(long long i) const { return mem.ctl((int) i)->P; }
};
//! ... or plain rows y[N][neq] (ReactorNet::eval, ctb_reactornet_eval) ...
struct RowIO {
    long long N;
    int n, rtype;
    const double* yRows;
    const double* Pv;
    double* ydot;
    __device__ __forceinline__ long long count() const { return N; }
    __device__ __forceinline__ int type() const { return rtype; }
    __device__ __forceinline__ bool active(long long) const { return true; }
    __device__ __forceinline__ double yv(long long i, int e) const { return yRows[i * n + e]; }
    __device__ __forceinline__ double* f(long long i) const { return ydot + i * n; }
    __device__ __forceinline__ double P(long long i) const { return Pv[i]; }
};
//! ... or the n + 1 states of a forward-difference Jacobian of every flagged slot: state i = (flagged
//! slot i / (n + 1), column i % (n + 1)); column n is the unperturbed state. Right-hand sides go to
//! Fbuf[i][n]; fdJacFinishKernel forms the difference quotients.
struct FdIO {
    SlotMem mem;
    double* Fbuf;
    __device__ __forceinline__ long long count() const { return (long long) mem.setupCount[mem.parity] * (mem.n + 1); }
    __device__ __forceinline__ int type() const { return mem.rtype; }
    __device__ __forceinline__ bool active(long long) const { return true; }
    __device__ __forceinline__ double yv(long long i, int e) const {
        const int slot = mem.setupList[i / (mem.n + 1)], col = (int) (i % (mem.n + 1));
        const double* ys = mem.vec(slot) + 8 * mem.n;
        const double val = ys[e];
        return e == col ? val + fdIncrement(mem.rtype, mem.n, ys, col) : val;
    }
    __device__ __forceinline__ double* f(long long i) const { return Fbuf + i * mem.n; }
    __device__ __forceinline__ double P(long long i) const { return mem.ctl(mem.setupList[i / (mem.n + 1)])->P; }
};

//! Carves one tile's shared-memory areas (layout of planRhsTile).
template <int NL>
__device__ __forceinline__ ReactorTileBufs carveReactorTile(const TileView& tv, double* base, unsigned char* smem, int tableBytes,
                                                            int ti)
{
    ReactorTileBufs b;
    b.c = base;
    b.u = b.c + tv.K * CTB_LDC;
    b.v = b.u + tv.K * CTB_LDC;
    b.q = b.v + tv.K * CTB_LDC;
    // red2 aliases the u rows (dead once the rates of progress are formed), the tile's temperatures sit
    // in a q row that the scratch of the mixture sums does not reach (they are consumed long before
    // the rates of progress are written); small mechanisms get separate areas after the q rows
    const int qSpan = max((tv.R + 1) * CTB_TS, 3 * NL * CTB_TS);
    const bool uFits = tv.K * CTB_LDC >= NL * CTB_TS, qFits = tv.R - 3 >= 3 * NL;
    b.red2 = uFits ? b.u : b.q + qSpan;
    b.Trow = qFits ? b.q + (tv.R - 3) * CTB_TS : b.q + qSpan + NL * CTB_TS;
    b.safeFlag = reinterpret_cast<volatile int*>(smem + tableBytes + 16) + ti;
    b.barId = 1 + ti;
    return b;
}

template <int W, int MAXT, class IO>
__global__ void __launch_bounds__(MAXT) batchRhsTileKernel(const TileView tv, const unsigned char* __restrict__ gblob,
                                                           const IO io, const int tileDoubles, const int tableBytes)
{
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + tableBytes);
    tmaStage(smem, gblob + tv.begin, tv.bytes, bar);
    constexpr int TT = W * 32, NL = 2 * W;
    const int ti = threadIdx.x / TT, tid = threadIdx.x % TT;
    const int tilesPerCta = blockDim.x / TT;
    const int s = tid & (CTB_TS - 1), j = tid / CTB_TS;
    const int K = tv.K, R = tv.R;
    const int type = io.type(), sIdx = reactorSIdx(type), tIdx = reactorTIdx(type), n = K + sIdx;
    const unsigned magicNeq = (unsigned) ((0x100000000ULL + (unsigned long long) n - 1) / (unsigned long long) n);
    const ReactorTileBufs b = carveReactorTile<NL>(tv, reinterpret_cast<double*>(smem + tableBytes + 128) + (size_t) ti * tileDoubles,
                                                   smem, tableBytes, ti);
    Tile t;
    t.tp = smem;
    t.cS = b.c + s; t.uS = b.u + s; t.vS = b.v + s; t.qS = b.q + s;
    if (tid == 0) *b.safeFlag = 0;
    if (tid < CTB_TS) b.q[R * CTB_TS + tid] = 0.0; // the all-zero row the padded gather entries point to
    // rows of the solution vector that are not species: which of [T, V, m] sits where (reactor.h)
    const int vIdx = type == 1 || type == 2 ? 1 : -1, mIdx = type >= 2 ? 0 : -1;
    const long long S = io.count();
    const long long nTiles = (S + CTB_TS - 1) / CTB_TS;
    for (long long tile = (long long) blockIdx.x * tilesPerCta + ti; tile < nTiles; tile += (long long) gridDim.x * tilesPerCta) {
        const long long sl0 = tile * CTB_TS;
        const long long sl = min(sl0 + s, S - 1);
        const bool valid = sl0 + s < S && io.active(sl);
        // every warp of the tile holds all 16 slots: the vote is the same in all of them
        if (!__any_sync(0xffffffffu, valid)) continue;
        // ---- transposed load of the requested solution vectors; idle slots get a benign state
        for (int i = tid; i < CTB_TS * n; i += TT) {
            const int si = (int) __umulhi((unsigned) i, magicNeq);
            const int e = i - si * n;
            const long long ss = min(sl0 + si, S - 1);
            const bool ok = sl0 + si < S && io.active(ss);
            const double val = ok ? io.yv(ss, e) : (e == tIdx ? 1000.0 : 1.0);
            if (e >= sIdx) b.c[(e - sIdx) * CTB_LDC + si] = val;
            else b.Trow[(e == tIdx ? 0 : e == vIdx ? 1 : 2) * CTB_TS + si] = val;
        }
        const double P = valid ? io.P(sl) : 101325.0;
        tileSync(b.barId, TT);
        TileScalars sc;
        ReactorTileAux aux;
        reactorStateTile<NL, TT>(tv, t, b, s, j, type, b.Trow[s], P, vIdx >= 0 ? b.Trow[CTB_TS + s] : 1.0,
                                 mIdx >= 0 ? b.Trow[2 * CTB_TS + s] : 1.0, sc, aux);
        const double Tdot = reactorRhsTile<NL, TT>(tv, t, b, tid, s, j, sc, aux, t.cS);
        if (j == 0 && valid) {
            double* fs = io.f(sl);
            fs[tIdx] = Tdot;
            if (vIdx >= 0) fs[vIdx] = 0.0; // closed reactor: no wall motion, no mass flow
            if (mIdx >= 0) fs[mIdx] = 0.0;
        }
        for (int i = tid; i < CTB_TS * K; i += TT) {
            const int si = (int) __umulhi((unsigned) i, tv.magicK);
            const int k = i - si * K;
            const long long ss = sl0 + si;
            if (ss < S && io.active(ss)) io.f(ss)[sIdx + k] = b.c[k * CTB_LDC + si];
        }
        tileSync(b.barId, TT);
    }
}

//! Difference quotients of the forward-difference Jacobian: M[slot][col][row] = (F_col - F_base) / increment.
__global__ void fdJacFinishKernel(const SlotMem mem, const double* __restrict__ Fbuf)
{
    const int n = mem.n;
    const int count = mem.setupCount[mem.parity];
    for (int e = blockIdx.x; e < count; e += gridDim.x) {
        const int slot = mem.setupList[e];
        const double* ys = mem.vec(slot) + 8 * n;
        const double* F = Fbuf + (size_t) e * (n + 1) * n;
        double* M = mem.M(slot);
        for (int i = threadIdx.x; i < n * n; i += blockDim.x) {
            const int col = i / n, row = i - col * n;
            M[i] = (F[(size_t) col * n + row] - F[(size_t) n * n + row]) / fdIncrement(mem.rtype, n, ys, col);
        }
    }
}

//! J = d(ydot)/dy at the slot's requested state, for the slots the controller flagged in this round.
//! One NamedGroup of NT threads per slot; J goes to the slot's matrix region (column-major n x n),
//! where batchInvertKernel turns it into (I - gamma J)^-1.
template <int NT>
__global__ void __launch_bounds__(512) batchJacKernel(const MechView v, const unsigned char* __restrict__ gmech,
                                                      const SlotMem mem, const int groupDoubles)
{
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + v.hotBytes);
    tmaStage(smem, gmech, v.hotBytes, bar);
    const unsigned char* sm = smem;
    using G = NamedGroup<NT>;
    const int gInCta = threadIdx.x / NT;
    const int groups = blockDim.x / NT;
    double* base = reinterpret_cast<double*>(smem + v.hotBytes + 16) + (size_t) gInCta * groupDoubles;
    const G g(base, 1 + gInCta);
    const ReactorWork wk = carveWork(base + 16, v.K, v.R);
    const int n = wk.n, ldm = wk.ldm;
    const int count = mem.setupCount[mem.parity];
    for (int e = blockIdx.x * groups + gInCta; e < count; e += gridDim.x * groups) {
        const int s = mem.setupList[e];
        const double P = mem.ctl(s)->P;
        const double gamma = mem.ctl(s)->gamma;
        const double* ys = mem.vec(s) + 8 * n;
        for (int i = g.lane(); i < n; i += NT) wk.y[i] = ys[i];
        g.sync();
        buildNewtonMatrix(g, v, sm, gmech, wk, P, gamma, mem.prof, false); // J only
        long long tk = mem.prof ? clock64() : 0;
        double* dstM = mem.M(s);
        for (int e2 = g.lane(); e2 < n * n; e2 += NT) {
            const int col = e2 / n, row = e2 - col * n;
            dstM[e2] = wk.M[(size_t) col * ldm + row];
        }
        g.sync();
        if (mem.prof && g.lane() == 0) {
            atomicAdd(reinterpret_cast<unsigned long long*>(mem.prof + 6), (unsigned long long) (clock64() - tk));
            atomicAdd(reinterpret_cast<unsigned long long*>(mem.prof + 7), 1ULL);
        }
    }
}

//! The same Jacobian for TILES of 16 flagged slots (jac_tile.cuh): J = d(ydot)/dy of the const-pressure
//! mole reactor as buildNewtonMatrix of bdf.cuh forms it —
//!   species block  d(V wdot_k)/d n_j = ddCi_kj + (V wdot_k / (V c_tot) - sum_l ddCi_kl c_l / c_tot)
//!   T row          -(sum_k h_k J_kj + Tdot cp_j) / (m cp)          (IdealGasConstPressureMoleReactor.cpp:153-173)
//!   T column       one-sided finite difference, dT = T sqrt(eps)                                 (:126-152)
//! with the dilution sum taken from the homogeneity of the rates of progress (jac_tile.cuh). W = 4 warps
//! per tile; shared memory per tile: c, u, w rows | q rows + f(T+dT) rows, later 8 column buffers.
//! Where the Jacobian tile kernel finds its states and puts its matrices: the flagged slots ...
struct SlotJacIO {
    SlotMem mem;
    __device__ __forceinline__ long long count() const { return mem.setupCount[mem.parity]; }
    __device__ __forceinline__ const double* y(long long i) const { return mem.vec(mem.setupList[i]) + 8 * mem.n; }
    __device__ __forceinline__ double P(long long i) const { return mem.ctl(mem.setupList[i])->P; }
    __device__ __forceinline__ double* M(long long i) const { return mem.M(mem.setupList[i]); }
};
//! ... or plain rows (ctb_reactornet_getJacobian)
struct RowJacIO {
    long long N;
    int n;
    const double* yRows;
    const double* Pv;
    double* jac;
    __device__ __forceinline__ long long count() const { return N; }
    __device__ __forceinline__ const double* y(long long i) const { return yRows + i * n; }
    __device__ __forceinline__ double P(long long i) const { return Pv[i]; }
    __device__ __forceinline__ double* M(long long i) const { return jac + i * (long long) n * n; }
};

template <int W, int MAXT, class IO>
__global__ void __launch_bounds__(MAXT) batchJacTileKernel(const TileView tv, const unsigned char* __restrict__ gblob,
                                                           const IO io, double* __restrict__ scratch,
                                                           const int tileDoubles, const int tableBytes)
{
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + tableBytes);
    tmaStage(smem, gblob + tv.begin, tv.bytesJac, bar);
    constexpr int TT = W * 32, NL = 2 * W;
    const int ti = threadIdx.x / TT, tid = threadIdx.x % TT;
    const int tilesPerCta = blockDim.x / TT;
    const int s = tid & (CTB_TS - 1), j = tid / CTB_TS;
    const int K = tv.K, R = tv.R, n = K + 1;
    double* base = reinterpret_cast<double*>(smem + tableBytes + 128) + (size_t) ti * tileDoubles;
    ReactorTileBufs b;
    b.c = base;
    b.u = b.c + K * CTB_LDC;
    b.v = b.u + K * CTB_LDC;
    b.q = b.v + K * CTB_LDC;
    const int qSpan = max((R + 1) * CTB_TS, 3 * NL * CTB_TS);
    double* fT = b.q + qSpan;  // f(T + dT) rows, flushed before the column buffers take over
    const int span = max(qSpan + K * CTB_LDC, NL * n * CTB_LDC);
    b.red2 = b.q + span;
    b.Trow = b.red2 + NL * CTB_TS;
    b.safeFlag = reinterpret_cast<volatile int*>(smem + tableBytes + 16) + ti;
    b.barId = 1 + ti;
    Tile t;
    t.tp = smem;
    t.cS = b.c + s; t.uS = b.u + s; t.vS = b.v + s; t.qS = b.q + s;
    double* fTS = fT + s;
    double* colBuf = b.q + (size_t) j * n * CTB_LDC;
    double* colS = colBuf + s;
    JacRows jr;
    {
        double* g = scratch + ((size_t) blockIdx.x * tilesPerCta + ti) * jacScratchDoubles(R, tv.nTb) + s;
        jr.kfM = g; jr.krM = g + (size_t) R * CTB_TS; jr.h2 = jr.krM + (size_t) R * CTB_TS;
        jr.we = jr.h2 + (size_t) (R + 1) * CTB_TS;
        if (j == 0) jr.h2[R * CTB_TS] = 0.0;
    }
    if (tid == 0) *b.safeFlag = 0;
    const unsigned halfMask = 0xFFFFu << (16 * ((threadIdx.x >> 4) & 1));
    const int l16 = threadIdx.x & 15;
    const unsigned char* spTab = smem + tv.o_sp;
    const long long count = io.count();
    const long long nTiles = (count + CTB_TS - 1) / CTB_TS;
    for (long long tile = (long long) blockIdx.x * tilesPerCta + ti; tile < nTiles; tile += (long long) gridDim.x * tilesPerCta) {
        const long long e0 = tile * CTB_TS;
        const int nvalid = (int) min((long long) CTB_TS, count - e0);
        auto loadStates = [&]() {
            for (int i = tid; i < CTB_TS * n; i += TT) {
                const int si = (int) __umulhi((unsigned) i, tv.magicN);
                const int e = i - si * n;
                const double val = io.y(e0 + min(si, nvalid - 1))[e];
                if (e == 0) b.Trow[si] = val; else b.c[(e - 1) * CTB_LDC + si] = val;
            }
        };
        loadStates();
        const double P = io.P(e0 + min(s, nvalid - 1));
        if (tid < CTB_TS) b.q[R * CTB_TS + tid] = 0.0; // the column buffers of the previous tile covered it
        tileSync(b.barId, TT);
        const double T0 = b.Trow[s];
        const double dT = T0 * 1.4901161193847656e-08; // sqrt(DBL_EPSILON)
        TileScalars sc;
        ReactorTileAux aux;
        // ---- f(T + dT) into the fT rows
        reactorStateTile<NL, TT>(tv, t, b, s, j, 0, T0 + dT, P, 1.0, 1.0, sc, aux);
        const double Tdot1 = reactorRhsTile<NL, TT>(tv, t, b, tid, s, j, sc, aux, fTS);
        // ---- base state: the Jacobian pass
        loadStates();
        tileSync(b.barId, TT);
        reactorStateTile<NL, TT>(tv, t, b, s, j, 0, T0, P, 1.0, 1.0, sc, aux);
        double* red = b.q;
        double sumq = 0.0;
        if (speciesPhaseTile<NL, true, true>(tv, t, sc, aux.f2, aux.dens, j, sumq)) *b.safeFlag = 1;
        red[(2 * NL + j) * CTB_TS + s] = sumq;
        tileSync(b.barId, TT);
        const bool safe = *b.safeFlag != 0;
        if (safe) speciesPhaseSafeTile<NL>(tv, t, sc, j);
        sumq = 0.0;
#pragma unroll
        for (int l = 0; l < NL; l++) sumq += red[(2 * NL + l) * CTB_TS + s];
        tileSync(b.barId, TT);
        if (3 * NL > R && tid < CTB_TS) b.q[R * CTB_TS + tid] = 0.0;
        const double rctot = 1.0 / aux.ctot;
        for (int e = j; e < R; e += NL) ropJacTile<true>(tv, t, sc, e, safe, false, false, rctot, jr);
        tileSync(b.barId, TT);
        if (tid == 0) *b.safeFlag = 0;
        const double cp_mole = CTB_GAS_CONSTANT * (aux.mmw * sumq);
        aux.totalCp = aux.totalMass * (cp_mole / aux.mmw);
        const double rN = 1.0 / (aux.vol * aux.ctot), rdT = 1.0 / dT, vol = aux.vol, RT = aux.RT;
        double mcpdTdt = 0.0;
        gatherPhase2Tile<NL>(tv, t, jr.h2, j, [&](int k, double wdot, double hsum) {
            double cp, h, sr, rmw;
            nasaTile(spTab, k, sc.T, sc.logT, sc.recipT, cp, h, sr, rmw);
            const double f0 = wdot * vol;
            mcpdTdt -= wdot * (h * RT) * vol;
            fTS[k * CTB_LDC] = (fTS[k * CTB_LDC] - f0) * rdT; // T column, species rows
            t.uS[k * CTB_LDC] = h;                            // the (dead) u rows now hold h_k / RT
            t.vS[k * CTB_LDC] = f0 * rN + hsum;               // the (dead) w rows: what every column gets added
        });
        b.red2[j * CTB_TS + s] = mcpdTdt;
        tileSync(b.barId, TT);
        mcpdTdt = 0.0;
#pragma unroll
        for (int l = 0; l < NL; l++) mcpdTdt += b.red2[l * CTB_TS + s];
        const double Tdot0 = mcpdTdt / aux.totalCp;
        if (j == 0) b.Trow[s] = (Tdot1 - Tdot0) * rdT;
        tileSync(b.barId, TT);
        // ---- T column of J (column 0 of M): transposed, coalesced
        for (int i = tid; i < nvalid * n; i += TT) {
            const int si = (int) __umulhi((unsigned) i, tv.magicN);
            const int e = i - si * n;
            io.M(e0 + si)[e] = e == 0 ? b.Trow[si] : fT[(e - 1) * CTB_LDC + si];
        }
        tileSync(b.barId, TT); // the q and fT rows are dead: the column buffers take their place
        const double rC = 1.0 / aux.totalCp;
        jacColumnsTile<NL>(tv, t, jr, colS, j, false, [&](int jc) {
            double acc = 0.0;
            for (int k = 0; k < K; k++) {
                const double val = colS[k * CTB_LDC] + t.vS[k * CTB_LDC];
                colS[k * CTB_LDC] = val;
                acc += t.uS[k * CTB_LDC] * val;
            }
            double cp, h, sr, rmw;
            nasaTile(spTab, jc, sc.T, sc.logT, sc.recipT, cp, h, sr, rmw);
            colS[K * CTB_LDC] = -(acc * RT + Tdot0 * (CTB_GAS_CONSTANT * cp)) * rC; // T row entry of this column
            __syncwarp(halfMask);
            for (int ss = 0; ss < nvalid; ss++) {
                double* dst = io.M(e0 + ss) + (size_t) (1 + jc) * n;
                for (int i = l16; i < n; i += 16) dst[i] = colBuf[(i == 0 ? K : i - 1) * CTB_LDC + ss];
            }
            __syncwarp(halfMask);
        });
        tileSync(b.barId, TT);
    }
}

//! (I - gamma J)^-1 of the flagged slots by in-place Gauss-Jordan elimination with partial (row)
//! pivoting — the role of SUNLinSol_Dense's factorisation (CVodesIntegrator.cpp:382-411), with the
//! inverse instead of the factors so that the controller's solves have no dependent chain.
//! One CTA of RP * SPLIT threads per matrix: thread (h, row) keeps columns [h CPT, (h+1) CPT) of
//! matrix row `row` in REGISTERS for the whole elimination (the k loop is unrolled, so every register
//! index is static). Rows are never exchanged: step k picks the largest |a_ik| among the rows not
//! yet used as a pivot (every warp finds it redundantly from the published column, no cross-warp
//! traffic), the pivot row broadcasts its scaled segment through shared memory, and every other row
//! does CPT fused multiply-adds. The implied permutation is undone for free when the result is
//! stored: thread (h, row) pivoted at step kk holds row kk of the inverse, column r_j at position j.
template <int RP, int SPLIT, int CPT>
__global__ void __launch_bounds__(RP * SPLIT) batchInvertKernel(const SlotMem mem)
{
    static_assert(CPT % 2 == 0, "segments are read with 128-bit loads");
    __shared__ __align__(16) double colbuf[RP];
    __shared__ __align__(16) double rowbuf[SPLIT * CPT];
    __shared__ int piv[SPLIT * CPT];
    const int n = mem.n;
    const int tid = threadIdx.x, h = tid / RP, row = tid % RP, lane = tid & 31;
    const int count = mem.setupCount[mem.parity];
    const long long tk = mem.prof ? clock64() : 0;
    for (int e = blockIdx.x; e < count; e += gridDim.x) {
        const int s = mem.setupList[e];
        double* Mg = mem.M(s);
        const double gamma = mem.ctl(s)->gamma;
        const bool rowOk = row < n;
        double a[CPT];
#pragma unroll
        for (int j = 0; j < CPT; j++) {
            const int col = h * CPT + j;
            double val = 0.0;
            if (rowOk && col < n) val = (row == col ? 1.0 : 0.0) - gamma * Mg[(size_t) col * n + row];
            a[j] = val;
        }
        // rows lane, lane + 32, ... are the ones this lane scans in the pivot search
        unsigned usedMask = 0;
#pragma unroll
        for (int r = 0; r < RP / 32; r++) if (lane + 32 * r >= n) usedMask |= 1u << r;
        int myStep = -1;
        bool sing = false;
#pragma unroll
        for (int k = 0; k < SPLIT * CPT; k++) {
            if (k >= n || sing) break;
            if (h == k / CPT) colbuf[row] = a[k % CPT];
            __syncthreads();
            double best = -1.0;
            int bi = RP;
#pragma unroll
            for (int r = 0; r < RP / 32; r++) {
                const double val = fabs(colbuf[lane + 32 * r]);
                if (!((usedMask >> r) & 1u) && val > best) { best = val; bi = lane + 32 * r; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double ob = __shfl_xor_sync(0xffffffffu, best, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
            }
            if (!(best > 0.0) || !(best <= 1.7e308)) {
                sing = true;      // zero, NaN or infinite column: the controller retries with a smaller step
                __syncthreads();
                continue;
            }
            const int p = bi;
            if ((p & 31) == lane) usedMask |= 1u << (p >> 5);
            const double f = colbuf[row];
            const double pinv = 1.0 / colbuf[p];
            if (row == p) {
                myStep = k;
#pragma unroll
                for (int j = 0; j < CPT; j++) {
                    double val = a[j] * pinv;
                    if (j == k % CPT && h == k / CPT) val = pinv;
                    a[j] = val;
                    rowbuf[h * CPT + j] = val;
                }
                if (h == 0) piv[k] = p;
            }
            __syncthreads();
            if (row != p) {
                const double2* rb = reinterpret_cast<const double2*>(rowbuf + h * CPT);
#pragma unroll
                for (int j = 0; j < CPT; j += 2) {
                    const double2 r2 = rb[j / 2];
                    a[j] = (j == k % CPT && h == k / CPT) ? -f * r2.x : fma(-f, r2.x, a[j]);
                    a[j + 1] = (j + 1 == k % CPT && h == k / CPT) ? -f * r2.y : fma(-f, r2.y, a[j + 1]);
                }
            }
        }
        __syncthreads();
        if (!sing && myStep >= 0) {
#pragma unroll
            for (int j = 0; j < CPT; j++) {
                const int col = h * CPT + j;
                if (col < n) Mg[(size_t) piv[col] * n + myStep] = a[j];
            }
        }
        if (tid == 0) mem.ctl(s)->sing = sing ? 1 : 0;
        __syncthreads();
    }
    if (mem.prof && tid == 0) atomicAdd(reinterpret_cast<unsigned long long*>(mem.prof + 5), (unsigned long long) (clock64() - tk));
}

//! Launch description of the tile right-hand-side kernel; tiles == 0: the mechanism does not fit.
struct RhsTilePlan {
    int tiles = 0, grid = 0;
    size_t smemBytes = 0, tileDoubles = 0, tableBytes = 0;
};
constexpr int RHS_TILE_W = 8, RHS_TILE_MAXT = 768;

RhsTilePlan planRhsTile(const DeviceMech& m, long long S)
{
    RhsTilePlan p;
    const TileView& tv = m.tile;
    constexpr int NL = 2 * RHS_TILE_W;
    p.tableBytes = ((size_t) tv.bytes + 127) & ~(size_t) 127;
    const bool uFits = tv.K * CTB_LDC >= NL * CTB_TS, qFits = tv.R - 3 >= 3 * NL; // carveReactorTile
    const size_t d = 3 * (size_t) tv.K * CTB_LDC + std::max((size_t) (tv.R + 1) * CTB_TS, 3 * (size_t) NL * CTB_TS)
                     + (uFits && qFits ? 0 : (size_t) NL * CTB_TS + 3 * CTB_TS);
    p.tileDoubles = (d + 15) & ~(size_t) 15;
    const size_t fixed = p.tableBytes + 128;
    const size_t budget = (size_t) m.smemOptIn;
    if (fixed + p.tileDoubles * 8 > budget) return p;
    int tiles = (int) ((budget - fixed) / (p.tileDoubles * 8));
    tiles = std::min(tiles, std::min(15, RHS_TILE_MAXT / (RHS_TILE_W * 32)));
    const long long nTiles = (S + CTB_TS - 1) / CTB_TS;
    tiles = (int) std::min<long long>(tiles, std::max<long long>(1, (nTiles + m.numSMs - 1) / m.numSMs));
    p.tiles = tiles;
    p.smemBytes = fixed + (size_t) tiles * p.tileDoubles * 8;
    p.grid = (int) std::max<long long>(1, std::min<long long>((nTiles + tiles - 1) / tiles, m.numSMs));
    cudaFuncSetAttribute(batchRhsTileKernel<RHS_TILE_W, RHS_TILE_MAXT, SlotIO>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int) p.smemBytes);
    cudaFuncSetAttribute(batchRhsTileKernel<RHS_TILE_W, RHS_TILE_MAXT, RowIO>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int) p.smemBytes);
    cudaFuncSetAttribute(batchRhsTileKernel<RHS_TILE_W, RHS_TILE_MAXT, FdIO>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int) p.smemBytes);
    return p;
}

//! The same inversion with a ROTATING register row: after every pivot step the row is shifted left by
//! one position (fused into the update: a[j] <- a[j+1] - f b[j]) and the finished inverse column is
//! appended at the end, so the pivot column is always position 0, no register index depends on the
//! step, and the step loop is NOT unrolled: ~250 instructions instead of the ~9 000 of the unrolled
//! kernel, which spent most of its time waiting for instruction fetches (3-4 CTAs per SM walking
//! through 150 KB of code at different places). One thread per matrix row, CPT >= n columns each.
template <int RP, int CPT, int MINB>
__global__ void __launch_bounds__(RP, MINB) batchInvertRotKernel(const SlotMem mem)
{
    static_assert(CPT % 2 == 0, "rows are broadcast with 128-bit accesses");
    __shared__ __align__(16) double colbuf[RP];
    __shared__ __align__(16) double rowbuf[CPT];
    __shared__ int piv[CPT];
    const int n = mem.n;
    const int row = threadIdx.x, lane = threadIdx.x & 31;
    const int count = mem.setupCount[mem.parity];
    const long long tk = mem.prof ? clock64() : 0;
    for (int e = blockIdx.x; e < count; e += gridDim.x) {
        const int s = mem.setupList[e];
        double* Mg = mem.M(s);
        const double gamma = mem.ctl(s)->gamma;
        const bool rowOk = row < n;
        double a[CPT];
#pragma unroll
        for (int j = 0; j < CPT; j++) {
            double val = 0.0;
            if (rowOk && j < n) val = (row == j ? 1.0 : 0.0) - gamma * Mg[(size_t) j * n + row];
            a[j] = val;
        }
        unsigned usedMask = 0;
#pragma unroll
        for (int r = 0; r < RP / 32; r++) if (lane + 32 * r >= n) usedMask |= 1u << r;
        int myStep = -1;
        bool sing = false;
        colbuf[row] = a[0];
        __syncthreads();
#pragma unroll 1
        for (int k = 0; k < n; k++) {
            // Partial pivoting on the HIGH WORDS of |a[0]| (sign stripped: exponent and 20 mantissa bits order the
            // magnitudes well enough for stability): one integer warp reduction instead of five rounds of
            // double-precision shuffles and compares. Every warp scans the whole column and finds the same row.
            unsigned key = 0u;
            int bi = RP;
#pragma unroll
            for (int r = 0; r < RP / 32; r++) {
                const unsigned kk = ((usedMask >> r) & 1u) ? 0u : ((unsigned) __double2hiint(colbuf[lane + 32 * r]) & 0x7fffffffu);
                if (kk > key) { key = kk; bi = lane + 32 * r; }
            }
            const unsigned kmax = __reduce_max_sync(0xffffffffu, key);
            if (kmax == 0u || kmax >= 0x7ff00000u) {
                sing = true; // zero (or denormal), NaN or infinite column: the controller retries with a smaller step
                break;       // uniform: every warp sees the same column
            }
            const int p = __shfl_sync(0xffffffffu, bi, __ffs(__ballot_sync(0xffffffffu, key == kmax)) - 1);
            if ((p & 31) == lane) usedMask |= 1u << (p >> 5);
            const bool isP = row == p;
            if (isP) {
                // the RAW pivot row (shifted by one position) goes to shared memory; the scaling by 1 / pivot is
                // folded into every row's own multiplier, so nobody waits for this thread to scale it
                myStep = k;
                double2* wb = reinterpret_cast<double2*>(rowbuf);
#pragma unroll
                for (int j = 0; j < CPT - 2; j += 2) wb[j / 2] = make_double2(a[j + 1], a[j + 2]);
                wb[CPT / 2 - 1] = make_double2(a[CPT - 1], 1.0);
                piv[k] = p;
            }
            const double pv = colbuf[p];
            // |pivot| in [2^-1012, 2^1012]: reciprocal by the hardware seed and two Newton steps (the inverse only
            // steers a Newton iteration), else the full division
            const double pinv = (kmax > 0x00b00000u && kmax < 0x7f300000u) ? rcpFast(pv) : 1.0 / pv;
            const double alpha = isP ? pinv : -(a[0] * pinv);
            __syncthreads();
            {
                // a[j] <- a[j+1] - (a[0] / pivot) raw[j]; the pivot row itself: a[j] <- raw[j] / pivot, as the
                // same fused multiply-add on a zero addend (no divergent second loop)
                const double2* rb = reinterpret_cast<const double2*>(rowbuf);
#pragma unroll
                for (int j = 0; j < CPT - 2; j += 2) {
                    const double2 r2 = rb[j / 2];
                    a[j] = fma(alpha, r2.x, isP ? 0.0 : a[j + 1]);
                    a[j + 1] = fma(alpha, r2.y, isP ? 0.0 : a[j + 2]);
                }
                const double2 r2 = rb[CPT / 2 - 1];
                a[CPT - 2] = fma(alpha, r2.x, isP ? 0.0 : a[CPT - 1]);
                a[CPT - 1] = alpha * r2.y;
            }
            colbuf[row] = a[0];
            __syncthreads();
        }
        if (!sing) {
            // positions n .. CPT-1 of the original row are padding: rotate them out
#pragma unroll 1
            for (int k = n; k < CPT; k++) {
#pragma unroll
                for (int j = 0; j < CPT - 1; j++) a[j] = a[j + 1];
                a[CPT - 1] = 0.0;
            }
            if (myStep >= 0) {
#pragma unroll
                for (int j = 0; j < CPT; j++) {
                    if (j < n) Mg[(size_t) piv[j] * n + myStep] = a[j];
                }
            }
        }
        if (row == 0) mem.ctl(s)->sing = sing ? 1 : 0;
        __syncthreads();
    }
    if (mem.prof && threadIdx.x == 0) atomicAdd(reinterpret_cast<unsigned long long*>(mem.prof + 5), (unsigned long long) (clock64() - tk));
}

template <int RP, int CPT, int MINB>
void launchInvertRotT(const DeviceMech& m, const SlotMem& mem, cudaStream_t stream)
{
    auto kern = batchInvertRotKernel<RP, CPT, MINB>;
    int perSm = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kern, RP, 0);
    perSm = std::max(1, perSm);
    kern<<<m.numSMs * perSm, RP, 0, stream>>>(mem);
}

template <int NT>
bool launchJacT(const DeviceMech& m, const SlotMem& mem, cudaStream_t stream)
{
    const MechView& v = m.view;
    const size_t groupDoubles = 16 + reactorWorkDoubles(v.K, v.R);
    const size_t fixed = v.hotBytes + 16;
    const size_t budget = (size_t) m.smemOptIn - 1024;
    if (fixed + groupDoubles * 8 > budget) return false;
    int groups = (int) ((budget - fixed) / (groupDoubles * 8));
    groups = std::min(groups, std::min(15, 512 / NT));
    if (groups < 1) return false;
    const size_t smemBytes = fixed + (size_t) groups * groupDoubles * 8;
    auto kern = batchJacKernel<NT>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smemBytes);
    kern<<<m.numSMs, groups * NT, smemBytes, stream>>>(v, m.dBlob, mem, (int) groupDoubles);
    return true;
}

constexpr int RJAC_TILE_W = 4, RJAC_TILE_MAXT = 512;

//! Launch description of the reactor Jacobian tile kernel (tiles == 0: does not fit).
JacTilePlan planReactorJacTile(const DeviceMech& m)
{
    JacTilePlan p;
    const TileView& tv = m.tile;
    constexpr int NL = 2 * RJAC_TILE_W;
    const int K = tv.K, R = tv.R, n = K + 1;
    p.tableBytes = ((size_t) tv.bytesJac + 127) & ~(size_t) 127;
    const size_t qSpan = std::max((size_t) (R + 1) * CTB_TS, 3 * (size_t) NL * CTB_TS);
    const size_t span = std::max(qSpan + (size_t) K * CTB_LDC, (size_t) NL * n * CTB_LDC);
    p.tileDoubles = (3 * (size_t) K * CTB_LDC + span + (size_t) NL * CTB_TS + CTB_TS + 15) & ~(size_t) 15;
    const size_t fixed = p.tableBytes + 128;
    const size_t budget = (size_t) m.smemOptIn;
    if (fixed + p.tileDoubles * 8 > budget) return p;
    int tiles = (int) ((budget - fixed) / (p.tileDoubles * 8));
    tiles = std::min(tiles, std::min(15, RJAC_TILE_MAXT / (RJAC_TILE_W * 32)));
    p.tiles = tiles;
    p.smemBytes = fixed + (size_t) tiles * p.tileDoubles * 8;
    p.grid = m.numSMs;
    p.scratchBytes = (size_t) p.grid * tiles * jacScratchDoubles(R, tv.nTb) * 8;
    return p;
}

void launchJac(const DeviceMech& m, const SlotMem& mem, cudaStream_t stream, const JacTilePlan& tp, void* scratch)
{
    if (tp.tiles > 0 && scratch) {
        batchJacTileKernel<RJAC_TILE_W, RJAC_TILE_MAXT, SlotJacIO><<<tp.grid, tp.tiles * RJAC_TILE_W * 32, tp.smemBytes, stream>>>(
            m.tile, m.dBlob, SlotJacIO{mem}, static_cast<double*>(scratch), (int) tp.tileDoubles, (int) tp.tableBytes);
        return;
    }
    const MechView& v = m.view;
    const size_t groupBytes = (16 + reactorWorkDoubles(v.K, v.R)) * 8;
    const size_t budget = (size_t) m.smemOptIn - 1024 - v.hotBytes - 16;
    const int fit = (int) std::max<size_t>(1, budget / groupBytes);
    int nt = 32;
    while (nt < 256 && fit * nt * 2 <= 512) nt *= 2;
    bool ok;
    switch (nt) {
    case 32: ok = launchJacT<32>(m, mem, stream); break;
    case 64: ok = launchJacT<64>(m, mem, stream); break;
    case 128: ok = launchJacT<128>(m, mem, stream); break;
    default: ok = launchJacT<256>(m, mem, stream); break;
    }
    if (!ok) throw std::runtime_error("mechanism too large for the shared-memory Newton-matrix kernel");
}

template <int RP, int SPLIT, int CPT>
void launchInvertT(const DeviceMech& m, const SlotMem& mem, cudaStream_t stream)
{
    auto kern = batchInvertKernel<RP, SPLIT, CPT>;
    int perSm = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kern, RP * SPLIT, 0);
    perSm = std::max(1, perSm);
    kern<<<m.numSMs * perSm, RP * SPLIT, 0, stream>>>(mem);
}

void launchInvert(const DeviceMech& m, const SlotMem& mem, cudaStream_t stream)
{
    const int n = mem.n;
    static const bool unrolled = [] { const char* e = std::getenv("CTB_INVERT_KERNEL"); return e && std::string(e) == "unrolled"; }();
    if (!unrolled && n <= 104) {
        if (n <= 16) launchInvertRotT<32, 16, 16>(m, mem, stream);
        else if (n <= 32) launchInvertRotT<32, 32, 12>(m, mem, stream);
        else if (n <= 56) launchInvertRotT<64, 56, 6>(m, mem, stream);
        else if (n <= 80) launchInvertRotT<96, 80, 3>(m, mem, stream);
        else launchInvertRotT<128, 104, 2>(m, mem, stream);
        return;
    }
    if (n <= 16) launchInvertT<32, 1, 16>(m, mem, stream);
    else if (n <= 32) launchInvertT<32, 1, 32>(m, mem, stream);
    else if (n <= 56) launchInvertT<64, 1, 56>(m, mem, stream);
    else if (n <= 64) launchInvertT<64, 2, 32>(m, mem, stream);
    else if (n <= 104) launchInvertT<128, 2, 52>(m, mem, stream);
    else if (n <= 128) launchInvertT<128, 2, 64>(m, mem, stream);
    else throw std::runtime_error("more than 128 equations per reactor");
}

} // namespace

bool launchReactorJacobianTile(const DeviceMech& m, long long N, const double* y, const double* P, double* jac,
                               void* scratch, cudaStream_t stream)
{
    if (N <= 0) return true;
    const JacTilePlan tp = planReactorJacTile(m);
    if (tp.tiles == 0 || tp.scratchBytes > reactorJacScratchBytes(m)) return false;
    auto kern = batchJacTileKernel<RJAC_TILE_W, RJAC_TILE_MAXT, RowJacIO>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) tp.smemBytes);
    RowJacIO io = {N, m.tile.K + 1, y, P, jac};
    kern<<<tp.grid, tp.tiles * RJAC_TILE_W * 32, tp.smemBytes, stream>>>(m.tile, m.dBlob, io, static_cast<double*>(scratch),
                                                                        (int) tp.tileDoubles, (int) tp.tableBytes);
    return true;
}

bool launchReactorEvalTile(const DeviceMech& m, int type, long long N, const double* y, const double* P, double* ydot,
                           cudaStream_t stream)
{
    if (N <= 0) return true;
    const RhsTilePlan tp = planRhsTile(m, N);
    if (tp.tiles == 0) return false;
    RowIO io = {N, reactorNeq(type, m.tile.K), type, y, P, ydot};
    batchRhsTileKernel<RHS_TILE_W, RHS_TILE_MAXT, RowIO><<<tp.grid, tp.tiles * RHS_TILE_W * 32, tp.smemBytes, stream>>>(
        m.tile, m.dBlob, io, (int) tp.tileDoubles, (int) tp.tableBytes);
    return true;
}

size_t reactorJacScratchBytes(const DeviceMech& m)
{
    return (size_t) m.numSMs * 4 * jacScratchDoubles(m.tile.R, m.tile.nTb) * 8;
}

long long batchSlotCap(const DeviceMech& m, int neq)
{
    SlotMem tmp = {};
    const unsigned long long slotBytes = slotLayout(neq, tmp);
    if ((neq + 31) / 32 > 4) return 0;
    if ((unsigned long long) tmp.oM * 4 + 256 > (unsigned long long) m.smemOptIn - 1024) return 0;
    // enough slots for several waves of every kernel, bounded by ~8 GB of slot memory. 384 per SM: the 16-slot
    // tiles of the right-hand-side kernel are then a whole number of waves (3 tiles per SM), and the slots that
    // ask for a new Newton matrix in one round (7-8 % in an ignition batch) fit ONE pass of the Jacobian tile
    // kernel (2 tiles per SM), whose duration is the latency of a tile whatever their number
    // (512: 24.6 k, 448: 25.1 k, 384: 25.9 k, 320: 24.9 k gri30 reactors/s on 524 288 reactors)
    int perSm = 384;
    if (const char* e = std::getenv("CTB_SLOTS_PER_SM")) perSm = std::max(16, std::atoi(e)); // diagnostics: same results, other pace
    long long cap = (long long) m.numSMs * perSm;
    cap = std::min<long long>(cap, (8LL << 30) / (long long) slotBytes);
    return std::max<long long>(cap, 1);
}

size_t batchWorkspaceBytes(int neq, long long S, BatchLayout* lay, size_t jacScratchBytesFor)
{
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t) 255; return o; };
    SlotMem tmp = {};
    const unsigned long long slotBytes = slotLayout(neq, tmp);
    BatchLayout l;
    l.slots = take((size_t) slotBytes * S);
    l.pc = take(sizeof(int) * S);
    l.setupList = take(sizeof(int) * S);
    l.counters = take(256);
    l.jacScratch = take(jacScratchBytesFor ? jacScratchBytesFor : 16);
    if (lay) *lay = l;
    return off;
}

long long launchAdvanceBatch(const DeviceMech& m, const ReactorBatch& b, void* workspace, long long S,
                             cudaStream_t stream)
{
    const MechView& v = m.view;
    const int n = b.neq;
    const bool groupJac = std::getenv("CTB_JAC_KERNEL") && std::string(std::getenv("CTB_JAC_KERNEL")) == "group";
    JacTilePlan jp = (groupJac || b.type != 0) ? JacTilePlan() : planReactorJacTile(m);
    if (jp.tiles > 0 && jp.scratchBytes > reactorJacScratchBytes(m)) jp = JacTilePlan();
    if (jp.tiles > 0) {
        cudaFuncSetAttribute(batchJacTileKernel<RJAC_TILE_W, RJAC_TILE_MAXT, SlotJacIO>,
                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int) jp.smemBytes);
    }
    BatchLayout l;
    batchWorkspaceBytes(n, S, &l, b.type == 0 ? reactorJacScratchBytes(m) : sizeof(double) * (size_t) S * (n + 1) * n);
    unsigned char* w = static_cast<unsigned char*>(workspace);
    SlotMem mem = {};
    slotLayout(n, mem);
    mem.rtype = b.type;
    mem.slots = w + l.slots;
    mem.pc = reinterpret_cast<int*>(w + l.pc);
    mem.setupList = reinterpret_cast<int*>(w + l.setupList);
    mem.setupCount = reinterpret_cast<int*>(w + l.counters);
    mem.nActive = mem.setupCount + 2;
    mem.S = (int) S;
    const bool profile = std::getenv("CTB_PROFILE") != nullptr; // diagnostics only: never changes results
    mem.prof = profile ? reinterpret_cast<long long*>(w + l.counters + 64) : nullptr;
    if (profile) cudaMemsetAsync(mem.prof, 0, 16 * sizeof(long long), stream);

    // ---- rhs kernel: groups of 8 lanes, as many as shared memory holds
    const int rhsGroupDoubles = (2 * n + 4 * v.K + v.R + 1) & ~1;
    const size_t budget = (size_t) m.smemOptIn - 1024;
    const size_t rhsFixed = v.hotBytes + 16;
    if (rhsFixed + (size_t) rhsGroupDoubles * 8 * 4 > budget) throw std::runtime_error("mechanism too large for the rhs kernel");
    int rhsGroups = (int) ((budget - rhsFixed) / ((size_t) rhsGroupDoubles * 8));
    rhsGroups = std::min(64, rhsGroups / 4 * 4);
    const size_t rhsSmem = rhsFixed + (size_t) rhsGroups * rhsGroupDoubles * 8;
    cudaFuncSetAttribute(batchRhsKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) rhsSmem);
    const int rhsGrid = (int) std::max<long long>(1, std::min<long long>((S + rhsGroups - 1) / rhsGroups, m.numSMs));
    // the tile kernel (rates_tile.cuh) wherever the mechanism's tables and one tile fit in shared memory
    const bool groupRhs = std::getenv("CTB_RHS_KERNEL") && std::string(std::getenv("CTB_RHS_KERNEL")) == "group";
    const RhsTilePlan tp = (groupRhs && b.type == 0) ? RhsTilePlan() : planRhsTile(m, S);
    if (b.type != 0 && tp.tiles == 0) throw std::runtime_error("mechanism too large for the tile kernels of this reactor model");

    // ---- control kernel: one shared-memory [ctl | vec] buffer per warp
    int ctlWarps = (int) std::min<size_t>(CTB_CTL_WARPS, (budget - 256) / mem.oM);
    if (ctlWarps < 1) throw std::runtime_error("a reactor slot does not fit in shared memory");
    // what is left of the shared memory becomes the per-warp rings of the Newton matrix (MatRing): NBUF chunks of
    // an even number of columns; no room for two columns per chunk: the product loads through registers
    const size_t ctlFixed = (((size_t) ctlWarps * mem.oM + (size_t) ctlWarps * (1 + MatRing::NBUF) * 8 + 127) & ~(size_t) 127);
    int ringCols = 0;
    if (budget > ctlFixed) {
        const size_t perWarp = (budget - ctlFixed) / ctlWarps;
        ringCols = (int) std::min<size_t>(perWarp / ((size_t) MatRing::NBUF * n * 8), (size_t) n) & ~1;
    }
    if (std::getenv("CTB_CTL_GEMV") && std::string(std::getenv("CTB_CTL_GEMV")) == "registers") ringCols = 0; // diagnostics
    const size_t ctlSmem = ctlFixed + (size_t) ctlWarps * MatRing::NBUF * ringCols * n * 8;
    const int npl = (n + 31) / 32;
    if (npl > 4) throw std::runtime_error("more than 128 equations per reactor");
    const int ctlGrid = (int) std::max<long long>(1, std::min<long long>((S + ctlWarps - 1) / ctlWarps, m.numSMs));
    void (*ctlKern)(const ReactorBatch, const SlotMem, const int) =
        npl == 1 ? batchControlKernel<1> : npl == 2 ? batchControlKernel<2> : npl == 3 ? batchControlKernel<3>
                                                                                       : batchControlKernel<4>;
    cudaFuncSetAttribute(ctlKern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) ctlSmem);

    batchInitKernel<<<(unsigned) ((S + 255) / 256), 256, 0, stream>>>(mem);
    long long rounds = 0;
    // every reactor is bounded by maxSteps, every step by MXNCF + MXNEF attempts of at most
    // 2 * MAXCOR + 2 right-hand sides: the loop ends on nActive alone; the bound is a safety net
    const long long perReactor = (long long) b.maxSteps * (BDF_MXNCF + BDF_MXNEF) * (2 * BDF_MAXCOR + 2) + 64;
    const long long maxRounds = ((b.N + S - 1) / S + 1) * perReactor;
    int active = 1;
    // CTB_PROFILE: CUDA events around every kernel of a round, read at the 16-round sync points
    cudaEvent_t ev[16][5];
    double kernMs[4] = {0, 0, 0, 0};
    if (profile) for (auto& r : ev) for (auto& e : r) cudaEventCreate(&e);
    while (rounds < maxRounds) {
        cudaEvent_t* e5 = profile ? ev[rounds & 15] : nullptr;
        mem.parity = (int) (rounds & 1);
        if (e5) cudaEventRecord(e5[0], stream);
        ctlKern<<<ctlGrid, ctlWarps * 32, ctlSmem, stream>>>(b, mem, ringCols);
        if (e5) cudaEventRecord(e5[1], stream);
        if (tp.tiles > 0) {
            batchRhsTileKernel<RHS_TILE_W, RHS_TILE_MAXT, SlotIO><<<tp.grid, tp.tiles * RHS_TILE_W * 32, tp.smemBytes, stream>>>(
                m.tile, m.dBlob, SlotIO{mem}, (int) tp.tileDoubles, (int) tp.tableBytes);
        } else {
            batchRhsKernel<<<rhsGrid, rhsGroups * 8, rhsSmem, stream>>>(v, m.dBlob, mem, rhsGroupDoubles);
        }
        if (e5) cudaEventRecord(e5[2], stream);
        if (b.type == 0) {
            launchJac(m, mem, stream, jp, w + l.jacScratch);
        } else {
            // forward-difference Jacobian through the right-hand-side tile kernel: n + 1 states per flagged slot
            double* Fbuf = reinterpret_cast<double*>(w + l.jacScratch);
            batchRhsTileKernel<RHS_TILE_W, RHS_TILE_MAXT, FdIO><<<m.numSMs, tp.tiles * RHS_TILE_W * 32, tp.smemBytes, stream>>>(
                m.tile, m.dBlob, FdIO{mem, Fbuf}, (int) tp.tileDoubles, (int) tp.tableBytes);
            fdJacFinishKernel<<<m.numSMs * 4, 256, 0, stream>>>(mem, Fbuf);
        }
        if (e5) cudaEventRecord(e5[3], stream);
        launchInvert(m, mem, stream);
        if (e5) cudaEventRecord(e5[4], stream);
        rounds++;
        if ((rounds & 15) == 0) {
            cudaError_t e = cudaMemcpyAsync(&active, mem.nActive, sizeof(int), cudaMemcpyDeviceToHost, stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
            if (e != cudaSuccess) throw std::runtime_error(std::string("batch integrator: ") + cudaGetErrorString(e));
            if (profile) {
                for (auto& r : ev) {
                    for (int k = 0; k < 4; k++) {
                        float ms = 0.f;
                        cudaEventElapsedTime(&ms, r[k], r[k + 1]);
                        kernMs[k] += ms;
                    }
                }
            }
            if (active <= 0) break;
        }
    }
    if (active > 0) {
        // the control kernel of the last round may have emptied the pool after the last read
        cudaMemcpyAsync(&active, mem.nActive, sizeof(int), cudaMemcpyDeviceToHost, stream);
        cudaStreamSynchronize(stream);
        if (active > 0) {
            const long long items = std::max<long long>(S, b.N);
            batchAbortKernel<<<(unsigned) ((items + 255) / 256), 256, 0, stream>>>(b, mem);
        }
    }
    if (profile) {
        for (auto& r : ev) for (auto& e : r) cudaEventDestroy(e);
        std::fprintf(stderr, "[ctb profile] kernel time over %lld rounds: control %.1f ms, rhs %.1f ms, jacobian %.1f ms, invert %.1f ms\n",
                     rounds, kernMs[0], kernMs[1], kernMs[2], kernMs[3]);
        long long h[16];
        cudaMemcpy(h, mem.prof, sizeof(h), cudaMemcpyDeviceToHost);
        const double nc = h[11] ? (double) h[11] : 1.0;
        std::fprintf(stderr, "[ctb profile] control: %lld slot visits; cycles per visit: load %.0f, run %.0f, store %.0f\n",
                     h[11], h[8] / nc, h[9] / nc, h[10] / nc);
        const double ns = h[7] ? (double) h[7] : 1.0;
        std::fprintf(stderr, "[ctb profile] rounds %lld, setups %lld; cycles per setup: rhs+jacdata %.0f, jacobianCi %.0f, "
                     "dense rows %.0f, T column %.0f, writeback %.0f; invert CTA-cycles total %lld\n", rounds, h[7], h[0] / ns,
                     h[1] / ns, h[2] / ns, h[3] / ns, h[6] / ns, h[5]);
    }
    return rounds;
}

} // namespace ctb
