// MechTables.h — owning container for the flat mechanism tables of ctb_clib.h and the .ctbm
// file format (a little-endian list of named arrays). No dependency on the reference: this is
// the data contract between the host adaptor (MechanismExtract.h), the C ABI and the kernels.
#pragma once

#include "../ctb_clib.h"

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

namespace cantera_b200 {

struct MechTables {
    int32_t n_species = 0;
    int32_t n_reactions = 0;
    double p_ref = 101325.0;
    std::vector<double> mol_weights, nasa_tmid, nasa_tmin, nasa_tmax, nasa_low, nasa_high;
    std::vector<int32_t> rxn_type, rxn_flags;
    std::vector<double> rate_A, rate_b, rate_Ea_R, low_A, low_b, low_Ea_R, falloff_coeffs;
    std::vector<double> multiplier, delta_n;
    std::vector<int32_t> reac_ptr, reac_species;
    std::vector<double> reac_stoich, reac_order;
    std::vector<int32_t> prod_ptr, prod_species;
    std::vector<double> prod_stoich;
    std::vector<double> tb_default;
    std::vector<int32_t> tb_ptr, tb_species;
    std::vector<double> tb_eff;
    std::vector<int32_t> plog_ptr;
    std::vector<double> plog_P, plog_A, plog_b, plog_Ea_R;
    std::vector<int32_t> cheb_ptr, cheb_nT, cheb_nP;
    std::vector<double> cheb_limits, cheb_coeffs;
    // informational only (error messages, Python front end)
    std::vector<std::string> species_names, equations;

    //! Non-owning C view; valid while *this is alive and unmodified.
    ctb_mech_tables view() const {
        ctb_mech_tables t;
        std::memset(&t, 0, sizeof(t));
        t.n_species = n_species;
        t.n_reactions = n_reactions;
        t.p_ref = p_ref;
        t.mol_weights = mol_weights.data();
        t.nasa_tmid = nasa_tmid.data();
        t.nasa_tmin = nasa_tmin.data();
        t.nasa_tmax = nasa_tmax.data();
        t.nasa_low = nasa_low.data();
        t.nasa_high = nasa_high.data();
        t.rxn_type = rxn_type.data();
        t.rxn_flags = rxn_flags.data();
        t.rate_A = rate_A.data();
        t.rate_b = rate_b.data();
        t.rate_Ea_R = rate_Ea_R.data();
        t.low_A = low_A.data();
        t.low_b = low_b.data();
        t.low_Ea_R = low_Ea_R.data();
        t.falloff_coeffs = falloff_coeffs.data();
        t.multiplier = multiplier.data();
        t.delta_n = delta_n.data();
        t.reac_ptr = reac_ptr.data();
        t.reac_species = reac_species.data();
        t.reac_stoich = reac_stoich.data();
        t.reac_order = reac_order.data();
        t.prod_ptr = prod_ptr.data();
        t.prod_species = prod_species.data();
        t.prod_stoich = prod_stoich.data();
        t.tb_default = tb_default.data();
        t.tb_ptr = tb_ptr.data();
        t.tb_species = tb_species.data();
        t.tb_eff = tb_eff.data();
        t.plog_ptr = plog_ptr.data();
        t.plog_P = plog_P.data();
        t.plog_A = plog_A.data();
        t.plog_b = plog_b.data();
        t.plog_Ea_R = plog_Ea_R.data();
        t.cheb_ptr = cheb_ptr.data();
        t.cheb_nT = cheb_nT.data();
        t.cheb_nP = cheb_nP.data();
        t.cheb_limits = cheb_limits.data();
        t.cheb_coeffs = cheb_coeffs.data();
        return t;
    }

    //! Deep copy from a C view (sizes follow from the CSR pointers).
    static MechTables fromView(const ctb_mech_tables& t) {
        MechTables m;
        const size_t K = t.n_species, R = t.n_reactions;
        m.n_species = t.n_species;
        m.n_reactions = t.n_reactions;
        m.p_ref = t.p_ref;
        auto cpd = [](std::vector<double>& dst, const double* src, size_t n) {
            dst.assign(n, 0.0);
            if (src && n) std::memcpy(dst.data(), src, n * sizeof(double));
        };
        auto cpi = [](std::vector<int32_t>& dst, const int32_t* src, size_t n) {
            dst.assign(n, 0);
            if (src && n) std::memcpy(dst.data(), src, n * sizeof(int32_t));
        };
        cpd(m.mol_weights, t.mol_weights, K);
        cpd(m.nasa_tmid, t.nasa_tmid, K);
        cpd(m.nasa_tmin, t.nasa_tmin, K);
        cpd(m.nasa_tmax, t.nasa_tmax, K);
        cpd(m.nasa_low, t.nasa_low, 7 * K);
        cpd(m.nasa_high, t.nasa_high, 7 * K);
        cpi(m.rxn_type, t.rxn_type, R);
        cpi(m.rxn_flags, t.rxn_flags, R);
        cpd(m.rate_A, t.rate_A, R);
        cpd(m.rate_b, t.rate_b, R);
        cpd(m.rate_Ea_R, t.rate_Ea_R, R);
        cpd(m.low_A, t.low_A, R);
        cpd(m.low_b, t.low_b, R);
        cpd(m.low_Ea_R, t.low_Ea_R, R);
        cpd(m.falloff_coeffs, t.falloff_coeffs, 5 * R);
        cpd(m.multiplier, t.multiplier, R);
        cpd(m.delta_n, t.delta_n, R);
        auto csr = [&](std::vector<int32_t>& ptr, const int32_t* src) -> size_t {
            cpi(ptr, src, R + 1);
            return src ? (size_t) src[R] : 0;
        };
        size_t n = csr(m.reac_ptr, t.reac_ptr);
        cpi(m.reac_species, t.reac_species, n);
        cpd(m.reac_stoich, t.reac_stoich, n);
        cpd(m.reac_order, t.reac_order, n);
        n = csr(m.prod_ptr, t.prod_ptr);
        cpi(m.prod_species, t.prod_species, n);
        cpd(m.prod_stoich, t.prod_stoich, n);
        cpd(m.tb_default, t.tb_default, R);
        n = csr(m.tb_ptr, t.tb_ptr);
        cpi(m.tb_species, t.tb_species, n);
        cpd(m.tb_eff, t.tb_eff, n);
        n = csr(m.plog_ptr, t.plog_ptr);
        cpd(m.plog_P, t.plog_P, n);
        cpd(m.plog_A, t.plog_A, n);
        cpd(m.plog_b, t.plog_b, n);
        cpd(m.plog_Ea_R, t.plog_Ea_R, n);
        n = csr(m.cheb_ptr, t.cheb_ptr);
        cpi(m.cheb_nT, t.cheb_nT, R);
        cpi(m.cheb_nP, t.cheb_nP, R);
        cpd(m.cheb_limits, t.cheb_limits, 4 * R);
        cpd(m.cheb_coeffs, t.cheb_coeffs, n);
        return m;
    }

    // ---- .ctbm: "CTBM0001", u32 nArrays, then per array: u32 nameLen, name, u32 dtype
    //      (0 = f64, 1 = i32, 2 = bytes), u64 count, payload.
    void save(const std::string& path) const {
        FILE* f = std::fopen(path.c_str(), "wb");
        if (!f) throw std::runtime_error("MechTables::save: cannot open " + path);
        std::vector<std::pair<std::string, std::pair<uint32_t, std::string>>> arrs;
        auto addd = [&](const char* name, const std::vector<double>& v) {
            arrs.push_back({name, {0u, std::string((const char*) v.data(), v.size() * 8)}});
        };
        auto addi = [&](const char* name, const std::vector<int32_t>& v) {
            arrs.push_back({name, {1u, std::string((const char*) v.data(), v.size() * 4)}});
        };
        auto adds = [&](const char* name, const std::vector<std::string>& v) {
            std::string blob;
            for (size_t i = 0; i < v.size(); i++) {
                if (i) blob += '\n';
                blob += v[i];
            }
            arrs.push_back({name, {2u, blob}});
        };
        addi("dims", {n_species, n_reactions});
        addd("p_ref", {p_ref});
#define CTB_D(x) addd(#x, x)
#define CTB_I(x) addi(#x, x)
        CTB_D(mol_weights); CTB_D(nasa_tmid); CTB_D(nasa_tmin); CTB_D(nasa_tmax);
        CTB_D(nasa_low); CTB_D(nasa_high);
        CTB_I(rxn_type); CTB_I(rxn_flags);
        CTB_D(rate_A); CTB_D(rate_b); CTB_D(rate_Ea_R);
        CTB_D(low_A); CTB_D(low_b); CTB_D(low_Ea_R); CTB_D(falloff_coeffs);
        CTB_D(multiplier); CTB_D(delta_n);
        CTB_I(reac_ptr); CTB_I(reac_species); CTB_D(reac_stoich); CTB_D(reac_order);
        CTB_I(prod_ptr); CTB_I(prod_species); CTB_D(prod_stoich);
        CTB_D(tb_default); CTB_I(tb_ptr); CTB_I(tb_species); CTB_D(tb_eff);
        CTB_I(plog_ptr); CTB_D(plog_P); CTB_D(plog_A); CTB_D(plog_b); CTB_D(plog_Ea_R);
        CTB_I(cheb_ptr); CTB_I(cheb_nT); CTB_I(cheb_nP); CTB_D(cheb_limits); CTB_D(cheb_coeffs);
#undef CTB_D
#undef CTB_I
        adds("species_names", species_names);
        adds("equations", equations);
        std::fwrite("CTBM0001", 1, 8, f);
        uint32_t n = (uint32_t) arrs.size();
        std::fwrite(&n, 4, 1, f);
        for (auto& [name, td] : arrs) {
            uint32_t nl = (uint32_t) name.size();
            uint64_t cnt = td.first == 0 ? td.second.size() / 8
                         : td.first == 1 ? td.second.size() / 4 : td.second.size();
            std::fwrite(&nl, 4, 1, f);
            std::fwrite(name.data(), 1, nl, f);
            std::fwrite(&td.first, 4, 1, f);
            std::fwrite(&cnt, 8, 1, f);
            std::fwrite(td.second.data(), 1, td.second.size(), f);
        }
        std::fclose(f);
    }

    static MechTables load(const std::string& path) {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) throw std::runtime_error("MechTables::load: cannot open '" + path + "'");
        auto rd = [&](void* p, size_t n) {
            if (n && std::fread(p, 1, n, f) != n) {
                std::fclose(f);
                throw std::runtime_error("MechTables::load: truncated file '" + path + "'");
            }
        };
        char magic[8];
        rd(magic, 8);
        if (std::memcmp(magic, "CTBM0001", 8) != 0) {
            std::fclose(f);
            throw std::runtime_error("MechTables::load: '" + path + "' is not a CTBM0001 file");
        }
        uint32_t n;
        rd(&n, 4);
        MechTables m;
        for (uint32_t a = 0; a < n; a++) {
            uint32_t nl, dtype;
            uint64_t cnt;
            rd(&nl, 4);
            std::string name(nl, '\0');
            rd(name.data(), nl);
            rd(&dtype, 4);
            rd(&cnt, 8);
            std::vector<double> vd;
            std::vector<int32_t> vi;
            std::string vs;
            if (dtype == 0) { vd.resize(cnt); rd(vd.data(), cnt * 8); }
            else if (dtype == 1) { vi.resize(cnt); rd(vi.data(), cnt * 4); }
            else { vs.resize(cnt); rd(vs.data(), cnt); }
            auto splitLines = [](const std::string& s) {
                std::vector<std::string> out;
                size_t b = 0;
                while (b <= s.size() && !s.empty()) {
                    size_t e = s.find('\n', b);
                    if (e == std::string::npos) e = s.size();
                    out.push_back(s.substr(b, e - b));
                    b = e + 1;
                }
                return out;
            };
            if (name == "dims") { m.n_species = vi.at(0); m.n_reactions = vi.at(1); }
            else if (name == "p_ref") m.p_ref = vd.at(0);
            else if (name == "species_names") m.species_names = splitLines(vs);
            else if (name == "equations") m.equations = splitLines(vs);
#define CTB_D(x) else if (name == #x) m.x = std::move(vd)
#define CTB_I(x) else if (name == #x) m.x = std::move(vi)
            CTB_D(mol_weights); CTB_D(nasa_tmid); CTB_D(nasa_tmin); CTB_D(nasa_tmax);
            CTB_D(nasa_low); CTB_D(nasa_high);
            CTB_I(rxn_type); CTB_I(rxn_flags);
            CTB_D(rate_A); CTB_D(rate_b); CTB_D(rate_Ea_R);
            CTB_D(low_A); CTB_D(low_b); CTB_D(low_Ea_R); CTB_D(falloff_coeffs);
            CTB_D(multiplier); CTB_D(delta_n);
            CTB_I(reac_ptr); CTB_I(reac_species); CTB_D(reac_stoich); CTB_D(reac_order);
            CTB_I(prod_ptr); CTB_I(prod_species); CTB_D(prod_stoich);
            CTB_D(tb_default); CTB_I(tb_ptr); CTB_I(tb_species); CTB_D(tb_eff);
            CTB_I(plog_ptr); CTB_D(plog_P); CTB_D(plog_A); CTB_D(plog_b); CTB_D(plog_Ea_R);
            CTB_I(cheb_ptr); CTB_I(cheb_nT); CTB_I(cheb_nP); CTB_D(cheb_limits); CTB_D(cheb_coeffs);
#undef CTB_D
#undef CTB_I
            // unknown arrays are skipped (forward compatibility)
        }
        std::fclose(f);
        return m;
    }
};

} // namespace cantera_b200
