// ctb_cantera.cpp — libctb_cantera.so, the OPTIONAL bridge between the reference and the batched
// engine: the mechanism is loaded by the reference's own newSolution() in this process (same YAML,
// same unit conversion, same implicit third-body detection), flattened through the reference's
// public API (include/cantera_b200/MechanismExtract.h) and handed to ctb_kin_newFromTables of
// libctb.so. This file is the only product source that includes reference headers; libctb.so itself
// has no dependency on the reference.
#include "cantera_b200/MechanismExtract.h"
#include "ctb_clib.h"
#include "ctb_cantera.h"

#include "cantera/base/Solution.h"
#include "cantera/base/global.h"
#include "cantera/thermo/ThermoPhase.h"
#include "cantera/transport/Transport.h"

#include <string>
#include <vector>

namespace {

template <class F>
int32_t bridged(F&& f)
{
    try {
        return f();
    } catch (Cantera::CanteraError& e) {
        ctb_setError(e.what());
        return -1;
    } catch (std::exception& e) {
        ctb_setError(e.what());
        return CTB_ERR;
    } catch (...) {
        ctb_setError("unknown exception");
        return CTB_ERR;
    }
}

} // namespace

extern "C" {

int32_t ctb_addDataDirectory(const char* dir)
{
    return bridged([&] {
        Cantera::addDataDirectory(dir ? dir : "");
        return 0;
    });
}

int32_t ctb_kin_newFromSolution(void* solution, int32_t device)
{
    return bridged([&] {
        if (!solution) throw Cantera::CanteraError("ctb_kin_newFromSolution", "null Solution");
        cantera_b200::MechTables m = cantera_b200::tablesFromSolution(*static_cast<Cantera::Solution*>(solution));
        const ctb_mech_tables v = m.view();
        return ctb_kin_newFromTables(&v, device);
    });
}

int32_t ctb_kin_new(int32_t solHandle, const char* yaml, const char* phase, int32_t device)
{
    return bridged([&] {
        if (solHandle != -1) {
            throw Cantera::CanteraError("ctb_kin_new", "Solution handles of the generated CLib are not available in this "
                                        "build (pass -1 and a YAML file, or use ctb_kin_newFromSolution)");
        }
        if (!yaml) throw Cantera::CanteraError("ctb_kin_new", "null file name");
        auto sol = Cantera::newSolution(yaml, phase ? phase : "", "none");
        return ctb_kin_newFromSolution(sol.get(), device);
    });
}

int32_t ctb_trans_newFromSolution(void* solution, int32_t device)
{
    return bridged([&] {
        if (!solution) throw Cantera::CanteraError("ctb_trans_newFromSolution", "null Solution");
        auto& sol = *static_cast<Cantera::Solution*>(solution);
        auto thermo = sol.thermo();
        auto tr = sol.transport();
        if (!thermo || !tr) throw Cantera::CanteraError("ctb_trans_newFromSolution", "Solution has no thermo / transport");
        if (tr->transportModel() != "mixture-averaged") {
            throw Cantera::CanteraError("ctb_trans_newFromSolution", "only the mixture-averaged model (non-CK fits) is on the "
                                        "batched path; got '{}'", tr->transportModel());
        }
        // the fitted temperature polynomials, through the public getters (GasTransport.h:107-115)
        const size_t K = thermo->nSpecies();
        std::vector<double> visc(5 * K), cond(5 * K), diff;
        for (size_t k = 0; k < K; k++) {
            tr->getViscosityPolynomial(k, Cantera::span<double>(&visc[5 * k], 5));
            tr->getConductivityPolynomial(k, Cantera::span<double>(&cond[5 * k], 5));
        }
        for (size_t i = 0; i < K; i++) {
            for (size_t j = i; j < K; j++) {
                double c[5];
                tr->getBinDiffusivityPolynomial(i, j, Cantera::span<double>(c, 5));
                diff.insert(diff.end(), c, c + 5);
            }
        }
        auto mw = thermo->molecularWeights();
        std::vector<double> w(mw.begin(), mw.end());
        return ctb_trans_newFromArrays((int32_t) K, w.data(), visc.data(), cond.data(), diff.data(), device);
    });
}

int32_t ctb_trans_new(const char* yaml, const char* phase, int32_t device)
{
    return bridged([&] {
        if (!yaml) throw Cantera::CanteraError("ctb_trans_new", "null file name");
        auto sol = Cantera::newSolution(yaml, phase ? phase : "", "mixture-averaged");
        return ctb_trans_newFromSolution(sol.get(), device);
    });
}

} // extern "C"
