// reactor_batch.cu — the batch-synchronous integrator of BatchReactorNet (SURVEY.md §8 rows
// a12-a14): rounds of three kernels over a pool of SLOTS,
//
//   batchControlKernel  one warp per slot: the resumable BDF/Newton controller of bdf_batch.cuh
//                       (triangular solves stream the slot's LU factors from HBM, coalesced)
//   batchRhsKernel      one thread per slot: reactor right-hand side at the requested state
//                       (reactor_thread.cuh; species arrays in shared memory, [species][thread])
//   batchSetupKernel    one cooperative group per FLAGGED slot: analytic Newton matrix + dense LU
//                       in shared memory (bdf.cuh), factors written back to the slot's block
//
// with the per-slot integrator memory in HBM. Compared with the persistent one-kernel integrator
// of reactor_kernels.cu (3-5 reactors per SM, everything in shared memory) this keeps thousands of
// reactors in flight per SM-round and runs the chemistry at the throughput of the rates kernels.
#include "bdf_batch.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <stdexcept>

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
    if (s == 0) *mem.nActive = mem.S;
}

// ---- warp-level TMA helpers (1-D bulk copies, one mbarrier per warp)
__device__ __forceinline__ void bulkLoad(void* dstSmem, const void* srcGlobal, uint32_t bytes, unsigned long long* bar)
{
    const uint32_t barAddr = (uint32_t) __cvta_generic_to_shared(bar);
    const uint32_t dstAddr = (uint32_t) __cvta_generic_to_shared(dstSmem);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(barAddr), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dstAddr), "l"(srcGlobal), "r"(bytes), "r"(barAddr)
                 : "memory");
}
__device__ __forceinline__ void barWait(unsigned long long* bar, uint32_t parity)
{
    const uint32_t barAddr = (uint32_t) __cvta_generic_to_shared(bar);
    uint32_t done = 0;
    while (!done) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(barAddr), "r"(parity)
            : "memory");
    }
}
__device__ __forceinline__ void bulkStore(void* dstGlobal, const void* srcSmem, uint32_t bytes)
{
    const uint32_t srcAddr = (uint32_t) __cvta_generic_to_shared(srcSmem);
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dstGlobal), "r"(srcAddr), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

//! Persistent CTAs, one warp per slot at a time. The warp brings the slot's whole block (controller
//! memory, Nordsieck array, LU factors: 28 KB for gri30) into its own shared-memory buffer with ONE
//! TMA bulk copy, runs the controller there (BatchBdf::run: no global-memory latency inside the
//! step logic or the triangular solves), and writes the mutable prefix back with a bulk store.
//! The other warps of the CTA cover the copy latency; 8 buffers per SM keep HBM busy.
template <int NPL>
__global__ void __launch_bounds__(256) batchControlKernel(const ReactorBatch b, const SlotMem mem)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, W = blockDim.x >> 5;
    const uint32_t ctlBytes = mem.oJ; // everything but the saved Jacobian
    unsigned char* buf = smem + (size_t) warp * ctlBytes;
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + (size_t) W * ctlBytes) + warp;
    if (lane == 0) {
        const uint32_t barAddr = (uint32_t) __cvta_generic_to_shared(bar);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(barAddr));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const SubWarp<32> g;
    uint32_t parity = 0;
    const uint32_t prefix = mem.oPerm; // ctl + vec
    for (int s = blockIdx.x * W + warp; s < mem.S; s += gridDim.x * W) {
        if (mem.pc[s] == PC_IDLE) continue;
        unsigned char* gblk = mem.block(s);
        const long long t0 = mem.prof ? clock64() : 0;
        if (lane == 0) bulkLoad(buf, gblk, ctlBytes, bar);
        barWait(bar, parity);
        parity ^= 1;
        const long long t1 = mem.prof ? clock64() : 0;
        {
            BatchBdf<SubWarp<32>, NPL> bdf(g, b, mem, s, buf, nullptr);
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
//! lanes per slot (reactorRhs of bdf.cuh on rates.cuh), work areas in shared memory.
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

//! Newton matrix I - gamma J at the slot's requested state and its LU factors, for the slots the
//! controller flagged in this round. One NamedGroup of NT threads per slot.
template <int NT>
__global__ void __launch_bounds__(512) batchSetupKernel(const MechView v, const unsigned char* __restrict__ gmech,
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
    const int count = *mem.jacCount;
    for (int e = blockIdx.x * groups + gInCta; e < count; e += gridDim.x * groups) {
        const int s = mem.jacList[e];
        const double P = mem.ctl(s)->P;
        const double gamma = mem.ctl(s)->gamma;
        const double* ys = mem.vec(s) + 8 * n;
        for (int i = g.lane(); i < n; i += NT) wk.y[i] = ys[i];
        g.sync();
        buildNewtonMatrix(g, v, sm, gmech, wk, P, gamma, mem.prof, false); // J only: the LU kernel forms I - gamma J
        long long tk = mem.prof ? clock64() : 0;
        // J = d(ydot)/dy goes to the slot's block (column-major n x n), where it stays until the next
        // evaluation; batchLuKernel forms and factors I - gamma J from it
        double* dstM = mem.J(s);
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

//! Dense LU with partial pivoting of the flagged slots' Newton matrices (the role of
//! SUNLinSol_Dense, CVodesIntegrator.cpp:382-411): ONE WARP per matrix. The warp brings the slot's
//! [perm | M] region into its shared-memory buffer with a TMA bulk copy, factors it there with
//! warp-synchronous steps only (lanes own rows; the pivot is a shuffle arg-max), leaves the
//! reciprocal pivots on the diagonal and the composed permutation for the controller's solves, and
//! writes the region back with a bulk store. 8 matrices in flight per SM.
__global__ void __launch_bounds__(256) batchLuKernel(const SlotMem mem)
{
    extern __shared__ __align__(128) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, W = blockDim.x >> 5;
    const uint32_t region = (uint32_t) (mem.oJ - mem.oPerm); // [perm | M]
    unsigned char* buf = smem + (size_t) warp * region;
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + (size_t) W * region) + warp;
    if (lane == 0) {
        const uint32_t barAddr = (uint32_t) __cvta_generic_to_shared(bar);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(barAddr));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const int n = mem.n;
    int* perm = reinterpret_cast<int*>(buf);
    double* M = reinterpret_cast<double*>(buf + (mem.oM - mem.oPerm));
    const int count = *mem.setupCount;
    uint32_t parity = 0;
    long long tk = mem.prof ? clock64() : 0;
    for (int e = blockIdx.x * W + warp; e < count; e += gridDim.x * W) {
        const int s = mem.setupList[e];
        unsigned char* gblk = mem.block(s) + mem.oPerm;
        if (lane == 0) bulkLoad(M, mem.J(s), ((uint32_t) (n * n) * 8u + 15u) & ~15u, bar);
        barWait(bar, parity);
        parity ^= 1;
        {
            const double gamma = mem.ctl(s)->gamma;
            for (int col = 0; col < n; col++) {
                double* cc = M + (size_t) col * n;
                for (int row = lane; row < n; row += 32) cc[row] = (row == col ? 1.0 : 0.0) - gamma * cc[row];
            }
            __syncwarp();
        }
        int fail = 0;
        for (int k = 0; k < n; k++) {
            double* ck = M + (size_t) k * n;
            // pivot: arg-max |a_ik|, i >= k (lowest index on ties, as the sequential search)
            double best = -1.0;
            int bi = n;
            for (int i = k + lane; i < n; i += 32) {
                const double a = fabs(ck[i]);
                if (a > best) { best = a; bi = i; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const double ob = __shfl_xor_sync(0xffffffffu, best, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (ob > best || (ob == best && oi < bi)) { best = ob; bi = oi; }
            }
            const int p = bi;
            if (lane == 0) perm[k] = p; // recorded swap; composed below
            if (!(best > 0.0)) {
                if (!fail) fail = k + 1;
                continue;
            }
            if (p != k) {
                for (int j = lane; j < n; j += 32) {
                    double* cj = M + (size_t) j * n;
                    const double t = cj[k];
                    cj[k] = cj[p];
                    cj[p] = t;
                }
            }
            __syncwarp();
            const double rp = 1.0 / ck[k];
            double l[4];
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int i = k + 1 + lane + 32 * r;
                l[r] = 0.0;
                if (i < n) {
                    l[r] = ck[i] * rp;
                    ck[i] = l[r];
                }
            }
            // rank-1 update: lanes own rows (at most 4 chunks of 32 for n <= 128)
            const int rows = n - k - 1;
#pragma unroll 2
            for (int j = k + 1; j < n; j++) {
                double* cj = M + (size_t) j * n;
                const double akj = cj[k];
                double* cji = cj + k + 1 + lane;
                if (lane < rows) cji[0] -= l[0] * akj;
                if (lane + 32 < rows) cji[32] -= l[1] * akj;
                if (lane + 64 < rows) cji[64] -= l[2] * akj;
                if (lane + 96 < rows) cji[96] -= l[3] * akj;
            }
            __syncwarp();
        }
        // reciprocal pivots; composed permutation (P b)[i] = b[id[i]]
        for (int k = lane; k < n; k += 32) M[(size_t) k * n + k] = 1.0 / M[(size_t) k * n + k];
        {
            // id starts as the identity, distributed like the solve's vectors (element i in lane
            // i % 32, register i / 32); the recorded swaps are applied in order by shuffles
            int idr[4];
#pragma unroll
            for (int r = 0; r < 4; r++) idr[r] = lane + 32 * r;
            for (int k = 0; k < n; k++) {
                const int p = perm[k];
                if (p != k) {
                    int srcK = idr[0], srcP = idr[0];
#pragma unroll
                    for (int r = 1; r < 4; r++) {
                        if ((k >> 5) == r) srcK = idr[r];
                        if ((p >> 5) == r) srcP = idr[r];
                    }
                    const int vK = __shfl_sync(0xffffffffu, srcK, k & 31);
                    const int vP = __shfl_sync(0xffffffffu, srcP, p & 31);
#pragma unroll
                    for (int r = 0; r < 4; r++) {
                        if (lane == (k & 31) && (k >> 5) == r) idr[r] = vP;
                        if (lane == (p & 31) && (p >> 5) == r) idr[r] = vK;
                    }
                }
            }
            __syncwarp();
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int i = lane + 32 * r;
                if (i < n) perm[i] = idr[r];
            }
        }
        __syncwarp();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
            mem.ctl(s)->sing = fail;
            bulkStore(gblk, buf, region);
        }
        __syncwarp();
    }
    if (mem.prof && lane == 0) atomicAdd(reinterpret_cast<unsigned long long*>(mem.prof + 5), (unsigned long long) (clock64() - tk));
}

template <int NT>
bool launchSetupT(const DeviceMech& m, const SlotMem& mem, cudaStream_t stream)
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
    auto kern = batchSetupKernel<NT>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smemBytes);
    kern<<<m.numSMs, groups * NT, smemBytes, stream>>>(v, m.dBlob, mem, (int) groupDoubles);
    return true;
}

void launchSetup(const DeviceMech& m, const SlotMem& mem, cudaStream_t stream, cudaEvent_t between = nullptr)
{
    const MechView& v = m.view;
    const size_t groupBytes = (16 + reactorWorkDoubles(v.K, v.R)) * 8;
    const size_t budget = (size_t) m.smemOptIn - 1024 - v.hotBytes - 16;
    const int fit = (int) std::max<size_t>(1, budget / groupBytes);
    int nt = 32;
    while (nt < 256 && fit * nt * 2 <= 512) nt *= 2;
    if (const char* e = std::getenv("CTB_SETUP_NT")) nt = std::atoi(e);
    bool ok;
    switch (nt) {
    case 32: ok = launchSetupT<32>(m, mem, stream); break;
    case 64: ok = launchSetupT<64>(m, mem, stream); break;
    case 128: ok = launchSetupT<128>(m, mem, stream); break;
    case 256: ok = launchSetupT<256>(m, mem, stream); break;
    default: throw std::runtime_error("CTB_SETUP_NT must be 32, 64, 128 or 256");
    }
    if (!ok) throw std::runtime_error("mechanism too large for the shared-memory Newton-matrix kernel");
    if (between) cudaEventRecord(between, stream);
    // factorisation: one warp per flagged matrix
    const size_t region = (size_t) (mem.oJ - mem.oPerm);
    const int warps = (int) std::min<size_t>(8, ((size_t) m.smemOptIn - 1024 - 64) / (region + 256));
    if (warps < 1) throw std::runtime_error("Newton matrix does not fit in shared memory");
    const size_t luSmem = (size_t) warps * region + 64 + (size_t) warps * 32 * 8;
    cudaFuncSetAttribute(batchLuKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) luSmem);
    batchLuKernel<<<m.numSMs, warps * 32, luSmem, stream>>>(mem);
}

} // namespace

long long batchSlotCap(const DeviceMech& m, int neq)
{
    SlotMem tmp = {};
    const unsigned long long slotBytes = slotLayout(neq, tmp);
    (void) slotBytes;
    if (tmp.oJ + 64ull > (unsigned long long) m.smemOptIn - 1024) return 0; // a slot must fit the control kernel's buffer
    if ((neq + 31) / 32 > 4) return 0;
    // enough slots for several waves of every kernel, bounded by ~8 GB of slot memory
    long long cap = (long long) m.numSMs * 512;
    cap = std::min<long long>(cap, (8LL << 30) / (long long) slotBytes);
    return std::max<long long>(cap, 1);
}

size_t batchWorkspaceBytes(int neq, long long S, BatchLayout* lay)
{
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t) 255; return o; };
    SlotMem tmp = {};
    const unsigned long long slotBytes = slotLayout(neq, tmp);
    BatchLayout l;
    l.slots = take((size_t) slotBytes * S);
    l.pc = take(sizeof(int) * S);
    l.setupList = take(sizeof(int) * 2 * S);
    l.counters = take(256);
    if (lay) *lay = l;
    return off;
}

long long launchAdvanceBatch(const DeviceMech& m, const ReactorBatch& b, void* workspace, long long S,
                             cudaStream_t stream)
{
    const MechView& v = m.view;
    const int n = b.neq;
    BatchLayout l;
    batchWorkspaceBytes(n, S, &l);
    unsigned char* w = static_cast<unsigned char*>(workspace);
    SlotMem mem = {};
    slotLayout(n, mem);
    mem.slots = w + l.slots;
    mem.pc = reinterpret_cast<int*>(w + l.pc);
    mem.setupList = reinterpret_cast<int*>(w + l.setupList);
    mem.jacList = mem.setupList + S;
    mem.setupCount = reinterpret_cast<int*>(w + l.counters);
    mem.nActive = mem.setupCount + 1;
    mem.jacCount = mem.setupCount + 2;
    mem.S = (int) S;
    const bool profile = std::getenv("CTB_PROFILE") != nullptr;
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

    // ---- control kernel: one shared-memory slot buffer per warp
    int ctlWarps = (int) std::min<size_t>(8, (budget - 64) / mem.oJ);
    if (ctlWarps < 1) throw std::runtime_error("a reactor slot does not fit in shared memory");
    const size_t ctlSmem = (size_t) ctlWarps * mem.oJ + 64;
    const int npl = (n + 31) / 32;
    if (npl > 4) throw std::runtime_error("more than 128 equations per reactor: use the persistent integrator");
    const int ctlGrid = (int) std::max<long long>(1, std::min<long long>((S + ctlWarps - 1) / ctlWarps, m.numSMs));
    void (*ctlKern)(const ReactorBatch, const SlotMem) =
        npl == 1 ? batchControlKernel<1> : npl == 2 ? batchControlKernel<2> : npl == 3 ? batchControlKernel<3>
                                                                                       : batchControlKernel<4>;
    cudaFuncSetAttribute(ctlKern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) ctlSmem);

    batchInitKernel<<<(unsigned) ((S + 255) / 256), 256, 0, stream>>>(mem);
    long long rounds = 0;
    const long long maxRounds = 64LL * b.maxSteps + 1000;
    int active = 1;
    // CTB_PROFILE: CUDA events around every kernel of a round, read at the 16-round sync points
    cudaEvent_t ev[16][5];
    double kernMs[4] = {0, 0, 0, 0};
    if (profile) for (auto& r : ev) for (auto& e : r) cudaEventCreate(&e);
    while (rounds < maxRounds) {
        cudaEvent_t* e5 = profile ? ev[rounds & 15] : nullptr;
        cudaMemsetAsync(mem.setupCount, 0, sizeof(int), stream);
        cudaMemsetAsync(mem.jacCount, 0, sizeof(int), stream);
        if (e5) cudaEventRecord(e5[0], stream);
        ctlKern<<<ctlGrid, ctlWarps * 32, ctlSmem, stream>>>(b, mem);
        if (e5) cudaEventRecord(e5[1], stream);
        batchRhsKernel<<<rhsGrid, rhsGroups * 8, rhsSmem, stream>>>(v, m.dBlob, mem, rhsGroupDoubles);
        if (e5) cudaEventRecord(e5[2], stream);
        launchSetup(m, mem, stream, e5 ? e5[3] : nullptr);
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
    if (profile) {
        for (auto& r : ev) for (auto& e : r) cudaEventDestroy(e);
        std::fprintf(stderr, "[ctb profile] kernel time over %lld rounds: control %.1f ms, rhs %.1f ms, setup %.1f ms, lu %.1f ms\n",
                     rounds, kernMs[0], kernMs[1], kernMs[2], kernMs[3]);
    }
    if (active > 0) throw std::runtime_error("batch integrator: round limit reached");
    if (profile) {
        long long h[16];
        cudaMemcpy(h, mem.prof, sizeof(h), cudaMemcpyDeviceToHost);
        const double nc = h[11] ? (double) h[11] : 1.0;
        std::fprintf(stderr, "[ctb profile] control: %lld slot visits; cycles per visit: load %.0f, run %.0f, store %.0f\n",
                     h[11], h[8] / nc, h[9] / nc, h[10] / nc);
        const double ns = h[7] ? (double) h[7] : 1.0;
        std::fprintf(stderr, "[ctb profile] rounds %lld, setups %lld; cycles per setup: rhs+jacdata %.0f, jacobianCi %.0f, "
                     "dense rows %.0f, T column %.0f, I-gJ %.0f, LU %.0f, writeback %.0f\n", rounds, h[7], h[0] / ns,
                     h[1] / ns, h[2] / ns, h[3] / ns, h[4] / ns, h[5] / ns, h[6] / ns);
    }
    return rounds;
}

} // namespace ctb
