// reactor_advance.cu — the persistent single-kernel integrator of BatchReactorNet (SURVEY.md §8 rows
// a12-a14; opt-in with CTB_BDF_KERNEL=persistent, and the fallback when the slot pool of
// reactor_batch.cu does not fit).
//
//   advanceKernel<NT>   persistent CTAs; each cooperative group of NT threads pulls reactors from a
//                       global work queue (reactors differ ~10x in step count) and integrates one
//                       at a time with the BDF/Newton solver of bdf.cuh, entirely out of shared
//                       memory: per reactor, HBM sees y in, y + integrator memory out.
//
// Compiled once per group width (-DCTB_ADV_NT=32|64|128|256, cantera_b200/build.py): the four
// instantiations take minutes each and build in parallel this way.
#include "bdf.cuh"

#include <algorithm>
#include <cstdlib>
#include <stdexcept>

#ifndef CTB_ADV_NT
#error "compile with -DCTB_ADV_NT=32, 64, 128 or 256"
#endif

namespace ctb {

namespace {

template <int NT>
__global__ void __launch_bounds__(512) advanceKernel(const MechView v, const unsigned char* __restrict__ gmech,
                                                     const ReactorBatch a, const int groupDoubles)
{
    extern __shared__ __align__(16) unsigned char smem[];
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + v.hotBytes);
    tmaStage(smem, gmech, v.hotBytes, bar);
    const unsigned char* sm = smem;

    using G = NamedGroup<NT>;
    const int gInCta = threadIdx.x / NT;
    double* base = reinterpret_cast<double*>(smem + v.hotBytes + 16) + (size_t) gInCta * groupDoubles;
    const G g(base, 1 + gInCta);
    const ReactorWork wk = carveWork(base + 16, v.K, v.R);
    const int n = wk.n;
    Bdf<G> bdf(g, v, sm, gmech, wk);
    bdf.rtol = a.rtol;
    bdf.atolT = a.atol;
    bdf.crateFloor = a.crateFloor;

    for (;;) {
        // ---- next reactor from the queue
        long long r;
        if (NT == 32) {
            unsigned long long t = 0;
            if (g.lane() == 0) t = atomicAdd(a.counter, 1ULL);
            r = (long long) __shfl_sync(0xffffffffu, t, 0);
        } else {
            g.sync();
            if (g.lane() == 0) base[8] = (double) atomicAdd(a.counter, 1ULL);
            g.sync();
            r = (long long) base[8];
        }
        if (r >= a.N) break;
        if (a.status[r] != CTB_STATUS_OK && !a.restart) continue; // failed earlier: stays failed

        double* yg = a.y + r * n;
        double* hist = a.hist + r * a.histStride;
        bdf.P = a.P[r];
        bdf.nfe = bdf.nsetups = bdf.nni = bdf.ncfn = bdf.netf = 0;
        int status = CTB_STATUS_OK;
        if (a.restart) {
            double moles = 0.0;
            for (int i = g.lane(); i < n; i += NT) {
                const double yi = yg[i];
                wk.zn[i] = yi;
                if (i > 0) moles += yi;
            }
            moles = g.sum(moles);
            // MoleReactor::updateDefaultTolerances (MoleReactor.cpp:64-76): atol * total moles for
            // the species rows unless the user set an absolute tolerance (ReactorNet.cpp:233-245)
            bdf.atolS = (a.atolUser || !(moles > 0.0)) ? a.atol : a.atol * moles;
            g.sync();
            status = bdf.start(a.t0, a.tout);
        } else {
            bdf.load(hist);
            g.sync();
        }
        const int nst0 = bdf.nst;
        const double Tign = a.ignDT > 0.0 ? a.T0[r] + a.ignDT : 0.0;
        bool ignPending = a.ignDT > 0.0 && !(a.tIgn[r] == a.tIgn[r]);
        double tPrev = bdf.tn, TPrev = wk.zn[0];
        while (status == CTB_STATUS_OK && bdf.tn < a.tout) {
            if (bdf.nst - nst0 >= a.maxSteps) {
                status = CTB_STATUS_TOO_MANY_STEPS;
                break;
            }
            if (bdf.nst > 0) bdf.setEwt(wk.zn);
            status = bdf.step();
            if (status != CTB_STATUS_OK) break;
            const double Tnow = wk.zn[0];
            if (!(Tnow == Tnow)) {
                status = CTB_STATUS_NOT_FINITE;
                break;
            }
            if (ignPending && Tnow >= Tign) {
                // first crossing of T0 + dT, linear in the step (openmp_ignition.cpp:70-73)
                if (g.lane() == 0) a.tIgn[r] = tPrev + (bdf.tn - tPrev) * (Tign - TPrev) / (Tnow - TPrev);
                ignPending = false;
            }
            tPrev = bdf.tn;
            TPrev = Tnow;
        }
        g.sync();
        if (status == CTB_STATUS_OK) {
            bdf.interpolate(a.tout, yg);
        } else {
            for (int i = g.lane(); i < n; i += NT) yg[i] = wk.zn[i];
        }
        bdf.save(hist);
        if (g.lane() == 0) {
            a.status[r] = status;
            long long* st = a.stats + r * CTB_NSTATS;
            st[STAT_STEPS] += bdf.nst - nst0;
            st[STAT_RHS] += bdf.nfe;
            st[STAT_JAC] += bdf.nsetups;
            st[STAT_SETUPS] += bdf.nsetups;
            st[STAT_NLITERS] += bdf.nni;
            st[STAT_NCFAILS] += bdf.ncfn;
            st[STAT_ETFAILS] += bdf.netf;
            st[STAT_ORDER] = bdf.q;
        }
    }
}

template <int NT>
bool launchAdvanceT(const DeviceMech& m, const ReactorBatch& b, int maxThreads, cudaStream_t stream)
{
    const MechView& v = m.view;
    const size_t groupDoubles = 16 + reactorWorkDoubles(v.K, v.R);
    const size_t fixed = v.hotBytes + 16;
    const size_t budget = (size_t) m.smemOptIn - 1024;
    if (fixed + groupDoubles * 8 > budget) return false;
    int groups = (int) ((budget - fixed) / (groupDoubles * 8));
    groups = std::min(groups, std::min(15, maxThreads / NT));
    if (groups < 1) return false;
    const size_t smemBytes = fixed + (size_t) groups * groupDoubles * 8;
    auto kern = advanceKernel<NT>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smemBytes);
    const long long want = (b.N + groups - 1) / groups;
    const int grid = (int) std::max<long long>(1, std::min<long long>(want, m.numSMs));
    kern<<<grid, groups * NT, smemBytes, stream>>>(v, m.dBlob, b, (int) groupDoubles);
    return true;
}

} // namespace

#define CTB_ADV_CAT2(a, b) a##b
#define CTB_ADV_CAT(a, b) CTB_ADV_CAT2(a, b)

bool CTB_ADV_CAT(launchAdvance, CTB_ADV_NT)(const DeviceMech& m, const ReactorBatch& b, int maxThreads,
                                            cudaStream_t stream)
{
    return launchAdvanceT<CTB_ADV_NT>(m, b, maxThreads, stream);
}

} // namespace ctb
