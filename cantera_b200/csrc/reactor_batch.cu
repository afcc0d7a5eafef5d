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
#include "reactor_thread.cuh"

#include <algorithm>
#include <cstdlib>
#include <stdexcept>

namespace ctb {

namespace {

__global__ void batchInitKernel(SlotMem mem)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < mem.S) {
        mem.ctl[s].pc = PC_NEW;
        mem.sing[s] = 0;
    }
    if (s == 0) *mem.nActive = mem.S;
}

template <int NPL>
__global__ void __launch_bounds__(128) batchControlKernel(const ReactorBatch b, const SlotMem mem)
{
    const int s = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (s >= mem.S) return;
    const SubWarp<32> g;
    BatchBdf<SubWarp<32>, NPL> bdf(g, b, mem, s, nullptr);
    bdf.run();
}

template <int U>
__global__ void __launch_bounds__(256, 1) batchRhsKernel(const MechView v, const unsigned char* __restrict__ gmech,
                                                         const SlotMem mem, const int neq, const int st)
{
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + v.hotBytes);
    tmaStage(smem, gmech, v.hotBytes, bar);
    const unsigned char* sm = smem;
    const int K = v.K;
    double* cAll = reinterpret_cast<double*>(smem + v.hotBytes + 16);
    TArr t;
    t.c = cAll + threadIdx.x;
    t.g = cAll + (size_t) K * st + threadIdx.x;
    t.w = cAll + (size_t) 2 * K * st + threadIdx.x;
    t.st = st;
    double* wAll = cAll + (size_t) 2 * K * st;
    const size_t ld = (size_t) SLOT_VECS * neq;
    for (int base = blockIdx.x * blockDim.x; base < mem.S; base += gridDim.x * blockDim.x) {
        const int nb = min((int) blockDim.x, mem.S - base);
        const double* yb = mem.vec + (size_t) base * ld + 8 * neq; // y block of the first slot
        // moles n_k of the CTA's slots -> c[k][thread] (coalesced rows, transposed through smem)
        loadTransposed(cAll, st, yb + 1, (long long) ld, nb, K);
        const int s = base + threadIdx.x;
        const bool valid = (int) threadIdx.x < nb && mem.ctl[s].pc != PC_IDLE;
        double T = 1000.0, P = 101325.0;
        if (valid) {
            T = yb[(size_t) threadIdx.x * ld];
            P = mem.ctl[s].P;
        }
        __syncthreads();
        double Tdot = 0.0;
        if (valid) Tdot = reactorRhsT<U>(v, sm, T, P, t);
        __syncthreads();
        double* fb = mem.vec + (size_t) base * ld + 9 * neq;
        storeTransposed(fb + 1, (long long) ld, wAll, st, nb, K);
        if (valid) fb[(size_t) threadIdx.x * ld] = Tdot;
        __syncthreads();
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
    const int count = *mem.setupCount;
    for (int e = blockIdx.x * groups + gInCta; e < count; e += gridDim.x * groups) {
        const int s = mem.setupList[e];
        const double P = mem.ctl[s].P;
        const double gamma = mem.ctl[s].gamma;
        const double* ys = mem.vec + ((size_t) s * SLOT_VECS + 8) * n;
        for (int i = g.lane(); i < n; i += NT) wk.y[i] = ys[i];
        g.sync();
        buildNewtonMatrix(g, v, sm, gmech, wk, P, gamma);
        const int fail = luFactor(g, wk);
        g.sync();
        // factors to the slot's block: column-major n x n, reciprocal pivots on the diagonal
        double* dstM = mem.M + (size_t) s * n * n;
        for (int e2 = g.lane(); e2 < n * n; e2 += NT) {
            const int col = e2 / n, row = e2 - col * n;
            const double a = wk.M[(size_t) col * ldm + row];
            dstM[e2] = (row == col) ? 1.0 / a : a;
        }
        // composed permutation: (P b)[i] = b[id[i]] after the recorded sequence of row swaps
        int* id = reinterpret_cast<int*>(wk.tmp);
        if (g.lane() == 0) {
            for (int i = 0; i < n; i++) id[i] = i;
            for (int k = 0; k < n; k++) {
                const int pk = wk.perm[k];
                if (pk != k) { const int t = id[k]; id[k] = id[pk]; id[pk] = t; }
            }
            mem.sing[s] = fail;
        }
        g.sync();
        for (int i = g.lane(); i < n; i += NT) mem.perm[(size_t) s * n + i] = id[i];
        g.sync();
    }
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

void launchSetup(const DeviceMech& m, const SlotMem& mem, cudaStream_t stream)
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
}

} // namespace

int threadKernelThreads(const DeviceMech& m, int arraysPerThread); // kin_kernels.cu

long long batchSlotCap(const DeviceMech& m)
{
    const int nth = threadKernelThreads(m, 3);
    if (nth <= 0) return 0;
    return (long long) m.numSMs * nth * 4;
}

size_t batchWorkspaceBytes(int neq, long long S, BatchLayout* lay)
{
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += (bytes + 255) & ~(size_t) 255; return o; };
    BatchLayout l;
    l.vec = take(sizeof(double) * SLOT_VECS * neq * S);
    l.M = take(sizeof(double) * (size_t) neq * neq * S);
    l.perm = take(sizeof(int) * neq * S);
    l.sing = take(sizeof(int) * S);
    l.ctl = take(sizeof(BatchCtl) * S);
    l.setupList = take(sizeof(int) * S);
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
    mem.vec = reinterpret_cast<double*>(w + l.vec);
    mem.M = reinterpret_cast<double*>(w + l.M);
    mem.perm = reinterpret_cast<int*>(w + l.perm);
    mem.sing = reinterpret_cast<int*>(w + l.sing);
    mem.ctl = reinterpret_cast<BatchCtl*>(w + l.ctl);
    mem.setupList = reinterpret_cast<int*>(w + l.setupList);
    mem.setupCount = reinterpret_cast<int*>(w + l.counters);
    mem.nActive = mem.setupCount + 1;
    mem.S = (int) S;

    const int nth = threadKernelThreads(m, 3);
    if (nth <= 0) throw std::runtime_error("mechanism too large for the thread-per-reactor kernels");
    const int st = nth + 1;
    const size_t rhsSmem = v.hotBytes + 16 + (size_t) 3 * v.K * st * 8;
    auto rhsKern = batchRhsKernel<4>;
    cudaFuncSetAttribute(rhsKern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) rhsSmem);
    const int rhsGrid = (int) std::max<long long>(1, std::min<long long>((S + nth - 1) / nth, m.numSMs));

    const int ctlThreads = 128;
    const int ctlGrid = (int) ((S + ctlThreads / 32 - 1) / (ctlThreads / 32));
    const int npl = (n + 31) / 32;
    if (npl > 4) throw std::runtime_error("more than 128 equations per reactor: use the persistent integrator");

    batchInitKernel<<<(unsigned) ((S + 255) / 256), 256, 0, stream>>>(mem);
    long long rounds = 0;
    const long long maxRounds = 64LL * b.maxSteps + 1000;
    int active = 1;
    while (rounds < maxRounds) {
        cudaMemsetAsync(mem.setupCount, 0, sizeof(int), stream);
        switch (npl) {
        case 1: batchControlKernel<1><<<ctlGrid, ctlThreads, 0, stream>>>(b, mem); break;
        case 2: batchControlKernel<2><<<ctlGrid, ctlThreads, 0, stream>>>(b, mem); break;
        case 3: batchControlKernel<3><<<ctlGrid, ctlThreads, 0, stream>>>(b, mem); break;
        default: batchControlKernel<4><<<ctlGrid, ctlThreads, 0, stream>>>(b, mem); break;
        }
        rhsKern<<<rhsGrid, nth, rhsSmem, stream>>>(v, m.dBlob, mem, n, st);
        launchSetup(m, mem, stream);
        rounds++;
        if ((rounds & 15) == 0) {
            cudaError_t e = cudaMemcpyAsync(&active, mem.nActive, sizeof(int), cudaMemcpyDeviceToHost, stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
            if (e != cudaSuccess) throw std::runtime_error(std::string("batch integrator: ") + cudaGetErrorString(e));
            if (active <= 0) break;
        }
    }
    if (active > 0) throw std::runtime_error("batch integrator: round limit reached");
    return rounds;
}

} // namespace ctb
