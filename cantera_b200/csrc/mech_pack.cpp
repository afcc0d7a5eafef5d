// mech_pack.cpp — host-side packer: ctb_mech_tables -> contiguous device image (mech_pack.h).
#include "mech_pack.h"

#include <algorithm>
#include <cmath>
#include <numeric>
#include <stdexcept>

namespace ctb {

namespace {

struct Blob {
    std::vector<unsigned char> data;
    template <class T>
    uint32_t put(const std::vector<T>& v) {
        while (data.size() % 16) data.push_back(0);
        uint32_t off = (uint32_t) data.size();
        const unsigned char* p = (const unsigned char*) v.data();
        data.insert(data.end(), p, p + v.size() * sizeof(T));
        return off;
    }
    void finish() {
        while (data.size() % 16) data.push_back(0);
    }
};

// The reference routes a reaction side to C1/C2/C3 when every coefficient is a whole number
// equal to its order and the side has at most three molecules; otherwise to C_AnyN
// (StoichManager.h:681-721).
bool simpleSide(int n, const double* stoich, const double* order, int& molecules)
{
    molecules = 0;
    for (int i = 0; i < n; i++) {
        if (std::fmod(stoich[i], 1.0) != 0.0 || stoich[i] != order[i]) return false;
        molecules += (int) stoich[i];
    }
    return n <= 3 && molecules >= 1 && molecules <= 3;
}

} // namespace

PackedMech packMechanism(const ctb_mech_tables& t)
{
    const int K = t.n_species, R = t.n_reactions;
    if (K < 1 || K > 65534) throw std::runtime_error("packMechanism: species count out of range");
    if (R < 0 || R > 65534) throw std::runtime_error("packMechanism: reaction count out of range");
    // the tables come from outside the library: every index becomes a device offset, so check them
    auto checkCsr = [&](const char* what, const int32_t* ptr, const int32_t* idx, int rows, int bound) {
        if (!ptr) throw std::runtime_error(std::string("packMechanism: missing table ") + what);
        if (ptr[0] != 0) throw std::runtime_error(std::string("packMechanism: ") + what + " does not start at 0");
        for (int i = 0; i < rows; i++) {
            if (ptr[i + 1] < ptr[i]) throw std::runtime_error(std::string("packMechanism: ") + what + " is not monotone");
        }
        if (idx) {
            for (int j = 0; j < ptr[rows]; j++) {
                if (idx[j] < 0 || idx[j] >= bound) {
                    throw std::runtime_error(std::string("packMechanism: ") + what + " has a species index out of range");
                }
            }
        }
    };
    checkCsr("reac_ptr/reac_species", t.reac_ptr, t.reac_species, R, K);
    checkCsr("prod_ptr/prod_species", t.prod_ptr, t.prod_species, R, K);
    checkCsr("tb_ptr/tb_species", t.tb_ptr, t.tb_species, R, K);
    checkCsr("plog_ptr", t.plog_ptr, nullptr, R, 0);
    checkCsr("cheb_ptr", t.cheb_ptr, nullptr, R, 0);
    for (int i = 0; i < R; i++) {
        if (t.rxn_type[i] == CTB_RATE_CHEBYSHEV && (long long) t.cheb_nT[i] * t.cheb_nP[i] != t.cheb_ptr[i + 1] - t.cheb_ptr[i]) {
            throw std::runtime_error("packMechanism: Chebyshev coefficient count does not match nT x nP");
        }
    }
    PackedMech pm;
    MechView& v = pm.v;
    std::memset(&v, 0, sizeof(v));
    v.K = K;
    v.R = R;
    v.p_ref = t.p_ref;

    // ---- execution order
    std::vector<int> cls(R);
    for (int i = 0; i < R; i++) {
        switch (t.rxn_type[i]) {
        case CTB_RATE_ARRHENIUS:
            cls[i] = (t.rate_b[i] == 0.0 && t.rate_Ea_R[i] == 0.0) ? CLS_ARR_CONST : CLS_ARR;
            break;
        case CTB_RATE_LINDEMANN: cls[i] = CLS_LINDEMANN; break;
        case CTB_RATE_TROE: cls[i] = CLS_TROE; break;
        case CTB_RATE_SRI: cls[i] = CLS_SRI; break;
        case CTB_RATE_TSANG: cls[i] = CLS_TSANG; break;
        case CTB_RATE_PLOG: cls[i] = CLS_PLOG; break;
        case CTB_RATE_CHEBYSHEV: cls[i] = CLS_CHEB; break;
        default: throw std::runtime_error("packMechanism: unknown rate type");
        }
    }
    // fast signature of the thread-per-state kernels (see RunRec)
    std::vector<int> fcode(R, -1);
    for (int i = 0; i < R; i++) {
        if ((cls[i] != CLS_ARR && cls[i] != CLS_ARR_CONST) || (t.rxn_flags[i] & CTB_RXN_THIRD_BODY)) continue;
        int pat[2] = {-1, -1};
        std::vector<int> seen;
        bool ok = true;
        for (int side = 0; side < 2 && ok; side++) {
            const int b0 = side ? t.prod_ptr[i] : t.reac_ptr[i], e0 = side ? t.prod_ptr[i + 1] : t.reac_ptr[i + 1];
            const int32_t* sp = side ? t.prod_species : t.reac_species;
            const double* st = side ? t.prod_stoich : t.reac_stoich;
            const double* od = side ? t.prod_stoich : t.reac_order;
            int molecules;
            if (e0 == b0 || !simpleSide(e0 - b0, st + b0, od + b0, molecules)) { ok = false; break; }
            const int n = e0 - b0;
            if (n == 1 && molecules == 1) pat[side] = PAT_A;
            else if (n == 2 && molecules == 2) pat[side] = PAT_AB;
            else if (n == 1 && molecules == 2) pat[side] = PAT_2A;
            else if (n == 3 && molecules == 3) pat[side] = PAT_ABC;
            else ok = false;
            for (int j = b0; j < e0 && ok; j++) {
                if (std::find(seen.begin(), seen.end(), sp[j]) != seen.end()) ok = false;
                seen.push_back(sp[j]);
            }
        }
        if (ok) fcode[i] = fastCode(cls[i] == CLS_ARR_CONST, (t.rxn_flags[i] & CTB_RXN_REVERSIBLE) ? 1 : 0, pat[0], pat[1]);
    }
    pm.perm.resize(R);
    std::iota(pm.perm.begin(), pm.perm.end(), 0);
    // within a class: irreversible first (they skip the equilibrium-constant exp), then
    // third-body flag, then reference order
    std::stable_sort(pm.perm.begin(), pm.perm.end(), [&](int a, int b) {
        if (cls[a] != cls[b]) return cls[a] < cls[b];
        if (fcode[a] != fcode[b]) return fcode[a] < fcode[b];
        int ra = t.rxn_flags[a] & CTB_RXN_REVERSIBLE, rb = t.rxn_flags[b] & CTB_RXN_REVERSIBLE;
        if (ra != rb) return ra < rb;
        int ta = t.rxn_flags[a] & CTB_RXN_THIRD_BODY, tb = t.rxn_flags[b] & CTB_RXN_THIRD_BODY;
        return ta < tb;
    });
    pm.invPerm.resize(R);
    for (int e = 0; e < R; e++) pm.invPerm[pm.perm[e]] = e;
    for (int c = 0; c <= CLS_COUNT; c++) v.clsBegin[c] = R;
    for (int e = R - 1; e >= 0; e--) v.clsBegin[cls[pm.perm[e]]] = e;
    for (int c = CLS_COUNT - 1; c >= 0; c--) v.clsBegin[c] = std::min(v.clsBegin[c], v.clsBegin[c + 1]);

    // ---- species tables
    std::vector<double> molwt(t.mol_weights, t.mol_weights + K), rmolwt(K), tmid(t.nasa_tmid, t.nasa_tmid + K);
    for (int k = 0; k < K; k++) rmolwt[k] = 1.0 / molwt[k]; // Phase::m_rmolwts (Phase.cpp:853)
    std::vector<double> low(t.nasa_low, t.nasa_low + 7 * K), high(t.nasa_high, t.nasa_high + 7 * K);

    // ---- reaction tables
    std::vector<double> A(R), b(R), EaR(R), mult(R), dn(R);
    std::vector<RxnRec> rec(R);
    std::vector<int16_t> tbIdx(R, -1), auxIdx(R, -1);
    std::vector<double> tbDefault, tbEff, fallLow, fallCoef;
    std::vector<int32_t> tbPtr{0}, plogPtr{0}, chebPtr{0}, chebDim, genPtr{0};
    std::vector<uint16_t> tbSp, genSp;
    std::vector<double> plogLogP, plogA, plogB, plogEaR, chebLim, chebCoef, genStoich, genOrder;
    int maxChebNT = 0;

    auto packSide = [&](int n, const int32_t* sp, const double* stoich, const double* order,
                        uint16_t* slots, uint8_t& cnt) {
        int molecules;
        slots[0] = slots[1] = slots[2] = 0xFFFF;
        if (n == 0) { cnt = 0; return; }
        if (simpleSide(n, stoich, order, molecules)) {
            int m = 0;
            for (int i = 0; i < n; i++)
                for (int j = 0; j < (int) stoich[i]; j++) slots[m++] = (uint16_t) sp[i];
            cnt = (uint8_t) molecules;
        } else {
            cnt = SIDE_GENERAL;
            slots[0] = (uint16_t) (genPtr.size() - 1);
            for (int i = 0; i < n; i++) {
                genSp.push_back((uint16_t) sp[i]);
                genStoich.push_back(stoich[i]);
                genOrder.push_back(order[i]);
            }
            genPtr.push_back((int32_t) genSp.size());
            if (genPtr.size() - 1 > 65534) throw std::runtime_error("packMechanism: too many general sides");
        }
    };

    for (int e = 0; e < R; e++) {
        int i = pm.perm[e];
        A[e] = t.rate_A[i]; b[e] = t.rate_b[i]; EaR[e] = t.rate_Ea_R[i];
        mult[e] = t.multiplier[i]; dn[e] = t.delta_n[i];
        RxnRec& r = rec[e];
        r.cls = (uint8_t) cls[i];
        r.flags = (uint8_t) t.rxn_flags[i];
        int rb = t.reac_ptr[i], re = t.reac_ptr[i + 1], pb = t.prod_ptr[i], pe = t.prod_ptr[i + 1];
        packSide(re - rb, t.reac_species + rb, t.reac_stoich + rb, t.reac_order + rb, r.r, r.nr);
        packSide(pe - pb, t.prod_species + pb, t.prod_stoich + pb, t.prod_stoich + pb, r.p, r.np);
        if (t.rxn_flags[i] & CTB_RXN_THIRD_BODY) {
            tbIdx[e] = (int16_t) tbDefault.size();
            tbDefault.push_back(t.tb_default[i]);
            for (int j = t.tb_ptr[i]; j < t.tb_ptr[i + 1]; j++) {
                tbSp.push_back((uint16_t) t.tb_species[j]);
                tbEff.push_back(t.tb_eff[j] - t.tb_default[i]); // ThirdBodyCalc::install (:23-44)
            }
            tbPtr.push_back((int32_t) tbSp.size());
        }
        if (cls[i] >= CLS_LINDEMANN && cls[i] <= CLS_TSANG) {
            auxIdx[e] = (int16_t) (fallLow.size() / 3);
            fallLow.push_back(t.low_A[i]); fallLow.push_back(t.low_b[i]); fallLow.push_back(t.low_Ea_R[i]);
            for (int j = 0; j < 5; j++) fallCoef.push_back(t.falloff_coeffs[5 * i + j]);
        } else if (cls[i] == CLS_PLOG) {
            auxIdx[e] = (int16_t) (plogPtr.size() - 1);
            for (int j = t.plog_ptr[i]; j < t.plog_ptr[i + 1]; j++) {
                plogLogP.push_back(std::log(t.plog_P[j])); // PlogRate::setRates (PlogRate.cpp:92-123)
                plogA.push_back(t.plog_A[j]); plogB.push_back(t.plog_b[j]); plogEaR.push_back(t.plog_Ea_R[j]);
            }
            plogPtr.push_back((int32_t) plogLogP.size());
        } else if (cls[i] == CLS_CHEB) {
            auxIdx[e] = (int16_t) (chebPtr.size() - 1);
            int nT = t.cheb_nT[i], nP = t.cheb_nP[i];
            maxChebNT = std::max(maxChebNT, nT);
            chebDim.push_back(nT); chebDim.push_back(nP);
            const double* lim = t.cheb_limits + 4 * i;
            // ChebyshevRate::setLimits (ChebyshevRate.cpp:96-112)
            double logPmin = std::log10(lim[2]), logPmax = std::log10(lim[3]);
            double TminInv = 1.0 / lim[0], TmaxInv = 1.0 / lim[1];
            chebLim.push_back(-TminInv - TmaxInv);
            chebLim.push_back(1.0 / (TmaxInv - TminInv));
            chebLim.push_back(-logPmin - logPmax);
            chebLim.push_back(1.0 / (logPmax - logPmin));
            for (int j = t.cheb_ptr[i]; j < t.cheb_ptr[i + 1]; j++) chebCoef.push_back(t.cheb_coeffs[j]);
            chebPtr.push_back((int32_t) chebCoef.size());
        }
    }
    if (maxChebNT > 16) throw std::runtime_error("packMechanism: Chebyshev nT > 16 not supported");

    // ---- gather lists (species-major net stoichiometry)
    std::vector<std::vector<std::pair<int, double>>> lists(K);
    bool fractional = false;
    for (int e = 0; e < R; e++) {
        int i = pm.perm[e];
        for (int j = t.prod_ptr[i]; j < t.prod_ptr[i + 1]; j++) {
            lists[t.prod_species[j]].push_back({e, t.prod_stoich[j]});
        }
        for (int j = t.reac_ptr[i]; j < t.reac_ptr[i + 1]; j++) {
            // order-only species (stoich 0, Kinetics.cpp:750-777) stay in the list: they add
            // nothing to the production rates but own a Jacobian column entry
            lists[t.reac_species[j]].push_back({e, -t.reac_stoich[j]});
        }
    }
    std::vector<int32_t> gPtr{0};
    std::vector<GatherEnt> gEnt;
    std::vector<double> gNu;
    for (int k = 0; k < K; k++) {
        for (auto& [e, nu] : lists[k]) {
            GatherEnt g;
            g.rxn = (uint16_t) e;
            if (std::fmod(nu, 1.0) != 0.0 || std::abs(nu) > 30000) {
                fractional = true;
                g.nu = nu > 0 ? 1 : (nu < 0 ? -1 : 0);
            } else {
                g.nu = (int16_t) nu;
            }
            gEnt.push_back(g);
            gNu.push_back(nu);
        }
        gPtr.push_back((int32_t) gEnt.size());
    }

    // ---- balanced deal of the gather lists to 8 lanes (a few radicals own most entries: H has 145
    // of gri30's 1245, so "species k to lane k mod 8" leaves seven lanes idle most of the time)
    std::vector<int32_t> gl8Ptr(9, 0);
    std::vector<uint16_t> gl8Sp;
    {
        std::vector<int> order(K);
        std::iota(order.begin(), order.end(), 0);
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return lists[a].size() > lists[b].size(); });
        std::vector<std::vector<int>> lane(8);
        long long load[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int k : order) {
            int best = 0;
            for (int l = 1; l < 8; l++) if (load[l] < load[best]) best = l;
            lane[best].push_back(k);
            load[best] += (long long) lists[k].size() + 2; // + per-species overhead
        }
        for (int l = 0; l < 8; l++) {
            for (int k : lane[l]) gl8Sp.push_back((uint16_t) k);
            gl8Ptr[l + 1] = (int32_t) gl8Sp.size();
        }
    }

    // ---- transposed third-body efficiencies
    std::vector<std::vector<std::pair<int, double>>> tbT(K);
    for (int e = 0; e < R; e++) {
        if (tbIdx[e] < 0) continue;
        for (int j = tbPtr[tbIdx[e]]; j < tbPtr[tbIdx[e] + 1]; j++) tbT[tbSp[j]].push_back({e, tbEff[j]});
    }
    std::vector<int32_t> tbTPtr{0};
    std::vector<uint16_t> tbTRxn;
    std::vector<double> tbTEff;
    for (int k = 0; k < K; k++) {
        for (auto& [e, eff] : tbT[k]) { tbTRxn.push_back((uint16_t) e); tbTEff.push_back(eff); }
        tbTPtr.push_back((int32_t) tbTRxn.size());
    }

    // ---- runs of equal fast signature (general-path reactions form runs of code -1)
    std::vector<RunRec> runs;
    for (int e = 0; e < R; e++) {
        const int code = fcode[pm.perm[e]];
        if (runs.empty() || runs.back().code != code) runs.push_back(RunRec{e, e + 1, code, 0});
        else runs.back().end = e + 1;
    }

    // ---- lay out: hot tables first (staged to shared memory), cold tables after
    Blob bl;
    v.o_molwt = bl.put(molwt); v.o_rmolwt = bl.put(rmolwt); v.o_tmid = bl.put(tmid);
    v.o_nasa_low = bl.put(low); v.o_nasa_high = bl.put(high);
    v.o_A = bl.put(A); v.o_b = bl.put(b); v.o_EaR = bl.put(EaR); v.o_mult = bl.put(mult); v.o_dn = bl.put(dn);
    v.o_rec = bl.put(rec);
    v.o_tbIdx = bl.put(tbIdx); v.o_auxIdx = bl.put(auxIdx);
    v.o_tbDefault = bl.put(tbDefault); v.o_tbPtr = bl.put(tbPtr); v.o_tbSp = bl.put(tbSp); v.o_tbEff = bl.put(tbEff);
    v.o_tbTPtr = bl.put(tbTPtr); v.o_tbTRxn = bl.put(tbTRxn); v.o_tbTEff = bl.put(tbTEff);
    v.o_fallLow = bl.put(fallLow); v.o_fallCoef = bl.put(fallCoef);
    v.o_plogPtr = bl.put(plogPtr); v.o_plogLogP = bl.put(plogLogP); v.o_plogA = bl.put(plogA);
    v.o_plogB = bl.put(plogB); v.o_plogEaR = bl.put(plogEaR);
    v.o_chebPtr = bl.put(chebPtr); v.o_chebDim = bl.put(chebDim); v.o_chebLim = bl.put(chebLim);
    v.o_chebCoef = bl.put(chebCoef);
    v.o_genPtr = bl.put(genPtr); v.o_genSp = bl.put(genSp); v.o_genStoich = bl.put(genStoich);
    v.o_genOrder = bl.put(genOrder);
    v.o_gPtr = bl.put(gPtr); v.o_gEnt = bl.put(gEnt);
    {
        std::vector<uint16_t> gCol, tbTCol;
        for (int k = 0; k < K; k++) {
            for (int i = gPtr[k]; i < gPtr[k + 1]; i++) gCol.push_back((uint16_t) k);
            for (int i = tbTPtr[k]; i < tbTPtr[k + 1]; i++) tbTCol.push_back((uint16_t) k);
        }
        v.o_gCol = bl.put(gCol);
        v.o_tbTCol = bl.put(tbTCol);
    }
    v.o_gl8Ptr = bl.put(gl8Ptr); v.o_gl8Sp = bl.put(gl8Sp);
    v.o_runs = bl.put(runs);
    v.nRuns = (int32_t) runs.size();
    v.o_view = bl.put(std::vector<unsigned char>(sizeof(MechView), 0)); // patched below
    bl.finish();
    v.hotBytes = (uint32_t) bl.data.size();
    v.o_gNu = bl.put(gNu);
    v.o_perm = bl.put(pm.perm);
    bl.finish();
    v.bytes = (uint32_t) bl.data.size();
    v.nTb = (int32_t) tbDefault.size();
    v.nFall = (int32_t) (fallLow.size() / 3);
    v.nPlogRx = (int32_t) plogPtr.size() - 1;
    v.nChebRx = (int32_t) chebPtr.size() - 1;
    v.nGen = (int32_t) genPtr.size() - 1;
    v.nGather = (int32_t) gEnt.size();
    v.hasFractional = fractional ? 1 : 0;
    v.maxChebNT = maxChebNT;
    std::memcpy(bl.data.data() + v.o_view, &v, sizeof(MechView));

    // ---- tile region (rates_tile.cuh): a second self-contained segment, offsets relative to it
    {
        while (bl.data.size() % 128) bl.data.push_back(0);
        TileView& tv = pm.tv;
        std::memset(&tv, 0, sizeof(tv));
        tv.K = K; tv.R = R; tv.nRuns = (int32_t) runs.size(); tv.nTb = (int32_t) tbDefault.size();
        tv.begin = (uint32_t) bl.data.size();
        tv.magicK = (uint32_t) ((0x100000000ULL + (unsigned long long) K - 1) / (unsigned long long) K);
        tv.lnRp0 = std::log(6.02214076e26 * 1.380649e-23 / t.p_ref);
        for (int e = 0; e < R; e++) {
            tv.maxAbsB = std::max(tv.maxAbsB, std::abs(b[e]));
            tv.maxAbsEaR = std::max(tv.maxAbsEaR, std::abs(EaR[e]));
        }
        for (size_t a = 0; a + 2 < fallLow.size(); a += 3) { // low-pressure limits of the falloff reactions
            tv.maxAbsB = std::max(tv.maxAbsB, std::abs(fallLow[a + 1]));
            tv.maxAbsEaR = std::max(tv.maxAbsEaR, std::abs(fallLow[a + 2]));
        }
        Blob tb;
        std::vector<Sp2> sp(K);
        for (int k = 0; k < K; k++) {
            for (int j = 0; j < 7; j++) { sp[k].low[j] = low[7 * k + j]; sp[k].high[j] = high[7 * k + j]; }
            sp[k].tmid = tmid[k];
            sp[k].rmolwt = rmolwt[k];
        }
        std::vector<Rx2> rx(R);
        std::vector<double> auxMult;
        for (int e = 0; e < R; e++) {
            Rx2& x = rx[e];
            std::memset(&x, 0, sizeof(x));
            const bool arr = rec[e].cls == CLS_ARR || rec[e].cls == CLS_ARR_CONST;
            x.A = arr ? A[e] * mult[e] : A[e];
            x.b = b[e]; x.EaR = EaR[e];
            // unit sides: heaviest molecule first, so that pairedProduct (rates_tile.cuh) multiplies
            // factors of similar magnitude
            uint16_t sr[3] = {rec[e].r[0], rec[e].r[1], rec[e].r[2]}, spd[3] = {rec[e].p[0], rec[e].p[1], rec[e].p[2]};
            auto heavyFirst = [&](uint8_t cnt, uint16_t* sl) {
                if (cnt == SIDE_GENERAL) return;
                std::stable_sort(sl, sl + cnt, [&](uint16_t a, uint16_t b2) { return molwt[a] > molwt[b2]; });
            };
            heavyFirst(rec[e].nr, sr);
            heavyFirst(rec[e].np, spd);
            auto slot = [&](uint8_t cnt, const uint16_t* sl, int i) -> uint32_t {
                if (cnt == SIDE_GENERAL) return sl[i];
                return sl[i] == 0xFFFF ? 0u : (uint32_t) sl[i] * CTB_LDC * 8u;
            };
            x.r0 = slot(rec[e].nr, sr, 0); x.r1 = slot(rec[e].nr, sr, 1); x.r2 = slot(rec[e].nr, sr, 2);
            x.p0 = slot(rec[e].np, spd, 0); x.p1 = slot(rec[e].np, spd, 1); x.p2 = slot(rec[e].np, spd, 2);
            x.nr = rec[e].nr; x.np = rec[e].np; x.cls = rec[e].cls; x.flags = rec[e].flags;
            x.tbIdx = tbIdx[e]; x.auxIdx = auxIdx[e];
            if (rec[e].nr != SIDE_GENERAL && rec[e].np != SIDE_GENERAL) {
                std::vector<uint16_t> all(sr, sr + rec[e].nr);
                all.insert(all.end(), spd, spd + rec[e].np);
                std::sort(all.begin(), all.end());
                if (std::adjacent_find(all.begin(), all.end()) == all.end()) x.hints |= RX2_HINT_DISTINCT;
            }
        }
        {
            // aux ordinals are per class family (falloff / plog / chebyshev count separately): one
            // multiplier table per family, concatenated: [falloff | plog | cheb]
            const int nF = (int) (fallLow.size() / 3), nP = (int) plogPtr.size() - 1, nC = (int) chebPtr.size() - 1;
            auxMult.assign((size_t) nF + nP + nC + 1, 1.0);
            for (int e = 0; e < R; e++) {
                if (auxIdx[e] < 0) continue;
                const int c = rec[e].cls;
                const int base = (c >= CLS_LINDEMANN && c <= CLS_TSANG) ? 0 : (c == CLS_PLOG ? nF : nF + nP);
                auxMult[(size_t) base + auxIdx[e]] = mult[e];
                rx[e].auxMultIdx = (uint32_t) (base + auxIdx[e]);
            }
        }
        // gather entries: per species [+1 pairs][-1 pairs], gen entries aside
        std::vector<uint32_t> gEnt2;
        std::vector<GGen2> gGen2;
        std::vector<GSp2> gsp(K);
        const uint32_t zeroRow = (uint32_t) R * CTB_TS * 8;
        for (int k = 0; k < K; k++) {
            GSp2& g = gsp[k];
            g.k = (uint16_t) k;
            g.start = (uint32_t) (gEnt2.size() * 4);
            g.genStart = (uint32_t) gGen2.size();
            std::vector<uint32_t> plus, minus;
            for (auto& [e, nu] : lists[k]) {
                if (nu == 0.0) continue;
                const uint32_t off = (uint32_t) e * CTB_TS * 8;
                const double an = std::abs(nu);
                if (std::fmod(an, 1.0) == 0.0 && an <= 3.0) {
                    for (int r = 0; r < (int) an; r++) (nu > 0 ? plus : minus).push_back(off);
                } else {
                    gGen2.push_back(GGen2{off, 0, nu});
                }
            }
            while (plus.size() % 4) plus.push_back(zeroRow);
            while (minus.size() % 4) minus.push_back(zeroRow);
            if (plus.size() / 4 > 65535 || minus.size() / 4 > 65535) throw std::runtime_error("packMechanism: gather list too long");
            g.nPlus4 = (uint16_t) (plus.size() / 4);
            g.nMinus4 = (uint16_t) (minus.size() / 4);
            g.nGen = (uint16_t) (gGen2.size() - g.genStart);
            gEnt2.insert(gEnt2.end(), plus.begin(), plus.end());
            gEnt2.insert(gEnt2.end(), minus.begin(), minus.end());
        }
        tv.hasGen = gGen2.empty() ? 0 : 1;
        tv.hasPdep = (plogLogP.empty() && chebCoef.empty()) ? 0 : 1;
        tv.magicN = (uint32_t) ((0x100000000ULL + (unsigned long long) K) / (unsigned long long) (K + 1));
        tv.o_sp = tb.put(sp); tv.o_molwt = tb.put(molwt); tv.o_rx = tb.put(rx); tv.o_runs = tb.put(runs);
        tv.o_gEnt = tb.put(gEnt2); tv.o_gGen = tb.put(gGen2);
        // the deal for 16 reaction lanes (wdot and right-hand-side kernels) sits in the common part, the
        // deal for 8 lanes (Jacobian kernels only) in the Jacobian-only tail below
        auto putDeal = [&](int d) {
            const int NL = 8 << d;
            std::vector<int> order(K);
            std::iota(order.begin(), order.end(), 0);
            auto cost = [&](int k) { return 2 * ((long long) gsp[k].nPlus4 + gsp[k].nMinus4) + gsp[k].nGen + 3; };
            std::stable_sort(order.begin(), order.end(), [&](int a, int b2) { return cost(a) > cost(b2); });
            std::vector<std::vector<int>> lane(NL);
            std::vector<long long> load(NL, 0);
            for (int k : order) {
                int best = 0;
                for (int l = 1; l < NL; l++) if (load[l] < load[best]) best = l;
                lane[best].push_back(k);
                load[best] += cost(k);
            }
            std::vector<GSp2> dealt;
            std::vector<int32_t> lanePtr(NL + 1, 0);
            for (int l = 0; l < NL; l++) {
                for (int k : lane[l]) dealt.push_back(gsp[k]);
                lanePtr[l + 1] = (int32_t) dealt.size();
            }
            tv.o_gSp[d] = tb.put(dealt);
            tv.o_gLane[d] = tb.put(lanePtr);
        };
        putDeal(1);
        tv.o_tbDefault = tb.put(tbDefault); tv.o_tbPtr = tb.put(tbPtr); tv.o_tbSp = tb.put(tbSp); tv.o_tbEff = tb.put(tbEff);
        tv.o_fallLow = tb.put(fallLow); tv.o_fallCoef = tb.put(fallCoef); tv.o_auxMult = tb.put(auxMult);
        tv.o_plogPtr = tb.put(plogPtr); tv.o_plogLogP = tb.put(plogLogP); tv.o_plogA = tb.put(plogA);
        tv.o_plogB = tb.put(plogB); tv.o_plogEaR = tb.put(plogEaR);
        tv.o_chebPtr = tb.put(chebPtr); tv.o_chebDim = tb.put(chebDim); tv.o_chebLim = tb.put(chebLim);
        tv.o_chebCoef = tb.put(chebCoef);
        tv.o_genPtr = tb.put(genPtr); tv.o_genSp = tb.put(genSp); tv.o_genStoich = tb.put(genStoich);
        tv.o_genOrder = tb.put(genOrder);
        tb.finish();
        tv.bytes = (uint32_t) tb.data.size();
        // Jacobian-only tables
        putDeal(0);
        // column items of the Jacobian: the gather entries without the product side of irreversible
        // reactions (their reverse rate of progress is identically zero)
        std::vector<int32_t> cPtr{0};
        std::vector<GatherEnt> cEnt;
        for (int k = 0; k < K; k++) {
            for (int i = gPtr[k]; i < gPtr[k + 1]; i++) {
                if (gEnt[i].nu > 0 && !(rec[gEnt[i].rxn].flags & CTB_RXN_REVERSIBLE)) continue;
                cEnt.push_back(gEnt[i]);
            }
            cPtr.push_back((int32_t) cEnt.size());
        }
        tv.o_colPtr = tb.put(cPtr); tv.o_colEnt = tb.put(cEnt);
        tv.o_tbTPtr = tb.put(tbTPtr); tv.o_tbTRxn = tb.put(tbTRxn); tv.o_tbTEff = tb.put(tbTEff);
        {
            std::vector<uint16_t> tbTW;
            for (uint16_t e : tbTRxn) tbTW.push_back((uint16_t) tbIdx[e]);
            tv.o_tbTW = tb.put(tbTW);
        }
        for (int d = 0; d < 2; d++) {
            const int NL = 4 << d;
            std::vector<int> order(K);
            std::iota(order.begin(), order.end(), 0);
            auto cost = [&](int k) { return (long long) (cPtr[k + 1] - cPtr[k]) + (tbTPtr[k + 1] - tbTPtr[k]) + 4; };
            std::stable_sort(order.begin(), order.end(), [&](int a, int b2) { return cost(a) > cost(b2); });
            std::vector<std::vector<int>> lane(NL);
            std::vector<long long> load(NL, 0);
            for (int k : order) {
                int best = 0;
                for (int l = 1; l < NL; l++) if (load[l] < load[best]) best = l;
                lane[best].push_back(k);
                load[best] += cost(k);
            }
            std::vector<uint16_t> colOrder;
            std::vector<int32_t> lanePtr(NL + 1, 0);
            for (int l = 0; l < NL; l++) {
                for (int k : lane[l]) colOrder.push_back((uint16_t) k);
                lanePtr[l + 1] = (int32_t) colOrder.size();
            }
            tv.o_colLane[d] = tb.put(lanePtr);
            tv.o_colOrder[d] = tb.put(colOrder);
        }
        tb.finish();
        tv.bytesJac = (uint32_t) tb.data.size();
        bl.data.insert(bl.data.end(), tb.data.begin(), tb.data.end());
        bl.finish();
        v.bytes = (uint32_t) bl.data.size();
        std::memcpy(bl.data.data() + v.o_view, &v, sizeof(MechView));
    }
    pm.blob = std::move(bl.data);
    return pm;
}

} // namespace ctb
