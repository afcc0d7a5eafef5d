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

#include <algorithm>
#include <cstdlib>
#include <cstring>
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

void launchJac(const DeviceMech& m, const JacArgs& a, cudaStream_t stream)
{
    if (a.N <= 0) return;
    const MechView& v = m.view;
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

} // namespace ctb
