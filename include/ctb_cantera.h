/*
 * ctb_cantera.h — C ABI of the OPTIONAL reference bridge libctb_cantera.so
 * (cantera_b200/csrc/ctb_cantera.cpp). It links the reference (libcantera) and libctb.so; handles
 * it returns are handles of libctb.so (include/ctb_clib.h), errors are read with ctb_getError.
 */
#ifndef CTB_CANTERA_H
#define CTB_CANTERA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* newSolution(yaml, phase) (src/base/Solution.cpp:267; CLib sol_newSolution, ctsol.yaml:11-12) IN PROCESS,
 * flattened through the reference's public API (include/cantera_b200/MechanismExtract.h) and handed to
 * ctb_kin_newFromTables. solHandle: reserved for a handle of the reference's generated CLib Solution
 * cabinet (that layer is generated with doxygen + sourcegen and is not part of this build): must be -1.
 * Returns a kinetics handle of THIS library, or -1 with the reference's error message set. */
int32_t ctb_kin_new(int32_t solHandle, const char* yaml, const char* phase, int32_t device);
/* The same from a live reference object: solution = Cantera::Solution* (C++ callers of the bridge). */
int32_t ctb_kin_newFromSolution(void* solution, int32_t device);
/* newSolution(yaml, phase, "mixture-averaged") in process; the fitted temperature polynomials of its transport
 * object (GasTransport::getViscosityPolynomial / getConductivityPolynomial / getBinDiffusivityPolynomial,
 * GasTransport.h:107-115) and the molecular weights go to ctb_trans_newFromArrays. Returns a transport handle of
 * libctb.so (ctb_trans_*), or -1 with the message set. */
int32_t ctb_trans_new(const char* yaml, const char* phase, int32_t device);
/* The same from a live Cantera::Solution* whose transport model is "mixture-averaged". */
int32_t ctb_trans_newFromSolution(void* solution, int32_t device);
/* Directory searched for mechanism files (Cantera::addDataDirectory). */
int32_t ctb_addDataDirectory(const char* dir);

#ifdef __cplusplus
}
#endif
#endif
