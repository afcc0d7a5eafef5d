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

#include <string>

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

} // extern "C"
