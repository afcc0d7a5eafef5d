// MechanismExtract.h — host adaptor on the REFERENCE side of the boundary.
//
// Walks a Solution loaded by the reference's own newSolution() (src/base/Solution.cpp:267) through
// public API only and flattens it into MechTables. Unit conversion of A, the implicit third-body
// detection (Reaction.cpp:389-555) and duplicate handling are therefore inherited from the
// reference's loader, bit for bit. Header-only; compiled only where Cantera headers exist (by an
// integrator, or by oracle/ref_build for the table extractor). The CUDA library never includes it.
#pragma once

#include "MechTables.h"

#include "cantera/base/Solution.h"
#include "cantera/kinetics/Arrhenius.h"
#include "cantera/kinetics/ChebyshevRate.h"
#include "cantera/kinetics/Falloff.h"
#include "cantera/kinetics/Kinetics.h"
#include "cantera/kinetics/PlogRate.h"
#include "cantera/kinetics/Reaction.h"
#include "cantera/thermo/MultiSpeciesThermo.h"
#include "cantera/thermo/ThermoPhase.h"
#include "cantera/thermo/speciesThermoTypes.h"

#include <algorithm>
#include <cmath>
#include <map>

namespace cantera_b200 {

namespace detail {
//! ArrheniusBase keeps Ea/R (m_Ea_R) but only exposes m_Ea_R*GasConstant (Arrhenius.h:116-118).
//! Pick the double x nearest Ea/GasConstant whose product reproduces the getter exactly.
inline double recoverEaOverR(double Ea) {
    double x = Ea / Cantera::GasConstant;
    if (x * Cantera::GasConstant == Ea) return x;
    double lo = std::nextafter(x, -INFINITY), hi = std::nextafter(x, INFINITY);
    if (lo * Cantera::GasConstant == Ea) return lo;
    if (hi * Cantera::GasConstant == Ea) return hi;
    return x;
}
inline void putArrhenius(const Cantera::ArrheniusRate& r, double& A, double& b, double& EaR) {
    A = r.preExponentialFactor();
    b = r.temperatureExponent();
    EaR = recoverEaOverR(r.activationEnergy());
}
//! TroeRate stores 1/T3, 1/T1 (inf for T=0) and reports their reciprocals (Falloff.cpp:243-281).
inline double recoverRecip(double c) {
    if (std::abs(c) < Cantera::SmallNumber) return INFINITY;
    double x = 1.0 / c;
    if (1.0 / x == c) return x;
    double lo = std::nextafter(x, -INFINITY), hi = std::nextafter(x, INFINITY);
    if (1.0 / lo == c) return lo;
    if (1.0 / hi == c) return hi;
    return x;
}
} // namespace detail

//! Flatten thermo + kinetics of a loaded ideal-gas Solution.
inline MechTables tablesFromSolution(Cantera::Solution& sol) {
    using namespace Cantera;
    auto thermo = sol.thermo();
    auto kin = sol.kinetics();
    if (!thermo || !kin) {
        throw CanteraError("cantera_b200::tablesFromSolution", "Solution has no thermo/kinetics");
    }
    if (thermo->type() != "ideal-gas") {
        throw CanteraError("cantera_b200::tablesFromSolution",
            "only the ideal-gas phase model is on the batched path; got '{}'", thermo->type());
    }
    if (kin->kineticsType() != "bulk") {
        throw CanteraError("cantera_b200::tablesFromSolution",
            "only bulk (gas) kinetics is on the batched path; got '{}'", kin->kineticsType());
    }
    MechTables m;
    const size_t K = thermo->nSpecies();
    const size_t R = kin->nReactions();
    m.n_species = (int32_t) K;
    m.n_reactions = (int32_t) R;
    m.p_ref = thermo->refPressure();

    auto mw = thermo->molecularWeights();
    m.mol_weights.assign(mw.begin(), mw.end());
    m.nasa_tmid.resize(K); m.nasa_tmin.resize(K); m.nasa_tmax.resize(K);
    m.nasa_low.resize(7 * K); m.nasa_high.resize(7 * K);
    auto& sp = thermo->speciesThermo();
    for (size_t k = 0; k < K; k++) {
        int type;
        double c[15], tmin, tmax, pref;
        m.species_names.push_back(thermo->speciesName(k));
        type = sp.reportType(k);
        if (type != NASA2) {
            throw CanteraError("cantera_b200::tablesFromSolution",
                "species '{}' uses thermo parameterisation {}; only two-range NASA7 is on the "
                "batched path", thermo->speciesName(k), type);
        }
        sp.reportParams(k, type, c, tmin, tmax, pref);
        if (pref != m.p_ref) {
            throw CanteraError("cantera_b200::tablesFromSolution",
                "species '{}' has a different reference pressure", thermo->speciesName(k));
        }
        m.nasa_tmid[k] = c[0];
        m.nasa_tmin[k] = tmin;
        m.nasa_tmax[k] = tmax;
        for (int j = 0; j < 7; j++) {
            m.nasa_high[7 * k + j] = c[1 + j];
            m.nasa_low[7 * k + j] = c[8 + j];
        }
    }

    m.rxn_type.assign(R, 0); m.rxn_flags.assign(R, 0);
    m.rate_A.assign(R, 0); m.rate_b.assign(R, 0); m.rate_Ea_R.assign(R, 0);
    m.low_A.assign(R, 0); m.low_b.assign(R, 0); m.low_Ea_R.assign(R, 0);
    m.falloff_coeffs.assign(5 * R, 0);
    m.multiplier.assign(R, 1); m.delta_n.assign(R, 0);
    m.tb_default.assign(R, 0);
    m.cheb_nT.assign(R, 0); m.cheb_nP.assign(R, 0); m.cheb_limits.assign(4 * R, 0);
    m.reac_ptr = {0}; m.prod_ptr = {0}; m.tb_ptr = {0}; m.plog_ptr = {0}; m.cheb_ptr = {0};

    for (size_t i = 0; i < R; i++) {
        auto r = kin->reaction(i);
        auto rate = r->rate();
        m.equations.push_back(r->equation());
        m.multiplier[i] = kin->multiplier(i);
        int32_t flags = r->reversible ? CTB_RXN_REVERSIBLE : 0;

        // -- stoichiometry, exactly as Kinetics::addReaction assembles it (Kinetics.cpp:734-777)
        std::vector<size_t> rk, pk;
        std::vector<double> rstoich, pstoich;
        for (const auto& [name, stoich] : r->reactants) {
            rk.push_back(kin->kineticsSpeciesIndex(name));
            rstoich.push_back(stoich);
        }
        for (const auto& [name, stoich] : r->products) {
            pk.push_back(kin->kineticsSpeciesIndex(name));
            pstoich.push_back(stoich);
        }
        std::vector<double> rorder = rstoich;
        for (const auto& [name, order] : r->orders) {
            size_t k = kin->kineticsSpeciesIndex(name);
            auto loc = std::find(rk.begin(), rk.end(), k);
            if (loc != rk.end()) {
                rorder[loc - rk.begin()] = order;
            } else {
                rk.push_back(k);
                rstoich.push_back(0.0);
                rorder.push_back(order);
            }
        }
        for (size_t n = 0; n < rk.size(); n++) {
            m.reac_species.push_back((int32_t) rk[n]);
            m.reac_stoich.push_back(rstoich[n]);
            m.reac_order.push_back(rorder[n]);
        }
        for (size_t n = 0; n < pk.size(); n++) {
            m.prod_species.push_back((int32_t) pk[n]);
            m.prod_stoich.push_back(pstoich[n]);
        }
        m.reac_ptr.push_back((int32_t) m.reac_species.size());
        m.prod_ptr.push_back((int32_t) m.prod_species.size());
        // m_dn (BulkKinetics.cpp:28-36): products first, then reactants
        double dn = 0.0;
        for (const auto& [name, stoich] : r->products) dn += stoich;
        for (const auto& [name, stoich] : r->reactants) dn -= stoich;
        m.delta_n[i] = dn;

        // -- third body (BulkKinetics.cpp:66-84)
        if (auto tb = r->thirdBody()) {
            flags |= CTB_RXN_THIRD_BODY;
            if (tb->mass_action) flags |= CTB_RXN_MASS_ACTION;
            m.tb_default[i] = tb->default_efficiency;
            std::map<size_t, double> eff;
            for (const auto& [name, e] : tb->efficiencies) {
                size_t k = kin->kineticsSpeciesIndex(name, false);
                if (k != npos) eff[k] = e;
            }
            for (const auto& [k, e] : eff) {
                m.tb_species.push_back((int32_t) k);
                m.tb_eff.push_back(e);
            }
        }
        m.tb_ptr.push_back((int32_t) m.tb_species.size());

        // -- rate parameterisation
        string type = rate->type();
        if (auto arr = std::dynamic_pointer_cast<ArrheniusRate>(rate); arr && type == "Arrhenius") {
            m.rxn_type[i] = CTB_RATE_ARRHENIUS;
            detail::putArrhenius(*arr, m.rate_A[i], m.rate_b[i], m.rate_Ea_R[i]);
        } else if (auto fo = std::dynamic_pointer_cast<FalloffRate>(rate)) {
            string sub = fo->subType();
            detail::putArrhenius(fo->highRate(), m.rate_A[i], m.rate_b[i], m.rate_Ea_R[i]);
            detail::putArrhenius(fo->lowRate(), m.low_A[i], m.low_b[i], m.low_Ea_R[i]);
            if (fo->chemicallyActivated()) flags |= CTB_RXN_CHEM_ACTIVATED;
            double c[5] = {0, 0, 0, 0, 0};
            size_t np = fo->nParameters();
            fo->getFalloffCoeffs(span<double>(c, np));
            double* out = &m.falloff_coeffs[5 * i];
            if (sub == "Lindemann") {
                m.rxn_type[i] = CTB_RATE_LINDEMANN;
            } else if (sub == "Troe") {
                m.rxn_type[i] = CTB_RATE_TROE;
                out[0] = c[0];
                out[1] = detail::recoverRecip(c[1]);
                out[2] = detail::recoverRecip(c[2]);
                out[3] = (np >= 4) ? c[3] : 0.0;
            } else if (sub == "SRI") {
                m.rxn_type[i] = CTB_RATE_SRI;
                out[0] = c[0]; out[1] = c[1]; out[2] = c[2];
                out[3] = (np >= 5) ? c[3] : 1.0;
                out[4] = (np >= 5) ? c[4] : 0.0;
            } else if (sub == "Tsang") {
                m.rxn_type[i] = CTB_RATE_TSANG;
                out[0] = c[0];
                out[1] = (np >= 2) ? c[1] : 0.0;
            } else {
                throw CanteraError("cantera_b200::tablesFromSolution",
                    "falloff function '{}' of reaction {} is not on the batched path", sub, i);
            }
        } else if (auto pl = std::dynamic_pointer_cast<PlogRate>(rate)) {
            m.rxn_type[i] = CTB_RATE_PLOG;
            for (const auto& [P, a] : pl->getRates()) {
                double A, b, EaR;
                detail::putArrhenius(a, A, b, EaR);
                m.plog_P.push_back(P);
                m.plog_A.push_back(A);
                m.plog_b.push_back(b);
                m.plog_Ea_R.push_back(EaR);
            }
        } else if (auto ch = std::dynamic_pointer_cast<ChebyshevRate>(rate)) {
            m.rxn_type[i] = CTB_RATE_CHEBYSHEV;
            const Array2D& d = ch->data();
            m.cheb_nT[i] = (int32_t) d.nRows();
            m.cheb_nP[i] = (int32_t) d.nColumns();
            m.cheb_limits[4 * i + 0] = ch->Tmin();
            m.cheb_limits[4 * i + 1] = ch->Tmax();
            m.cheb_limits[4 * i + 2] = ch->Pmin();
            m.cheb_limits[4 * i + 3] = ch->Pmax();
            for (size_t t = 0; t < d.nRows(); t++) {
                for (size_t p = 0; p < d.nColumns(); p++) {
                    m.cheb_coeffs.push_back(d(t, p));
                }
            }
        } else {
            throw CanteraError("cantera_b200::tablesFromSolution",
                "rate type '{}' of reaction {} ('{}') is not on the batched path",
                type, i, r->equation());
        }
        m.plog_ptr.push_back((int32_t) m.plog_P.size());
        m.cheb_ptr.push_back((int32_t) m.cheb_coeffs.size());
        m.rxn_flags[i] = flags;
    }
    return m;
}

} // namespace cantera_b200
