// mech_pack.h — packed device image of a mechanism.
//
// The flat tables of include/ctb_clib.h (reference reaction order, CSR stoichiometry) are
// re-laid out for the kernels: reactions are permuted into an EXECUTION ORDER grouped by rate
// parameterisation so that the lanes of a cooperative group run the same code in one round,
// sides with unit stoichiometry (the reference's C1/C2/C3 classes, StoichManager.h:131-326) are
// packed into one 16-byte record, everything else goes through a general (C_AnyN) CSR side table,
// and the net stoichiometric matrix is stored species-major ("gather lists") so that production
// rates are a conflict-free gather instead of the reference's scatter-add
// (StoichManager.h:732-744).
//
// The image is ONE contiguous 16-byte aligned blob addressed by byte offsets (MechView), so the
// same view works on the copy in global memory and on the copy a CTA stages into shared memory
// with a single TMA bulk copy (cp.async.bulk).
#pragma once

#include "../../include/ctb_clib.h"

#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

namespace ctb {

// Execution classes (sort key). Reactions of one class are contiguous in execution order.
enum RxnClass : int {
    CLS_ARR_CONST = 0, // Arrhenius with b == 0 and Ea == 0: kf = A, no exp
    CLS_ARR = 1,       // Arrhenius (elementary or three-body)
    CLS_LINDEMANN = 2,
    CLS_TROE = 3,
    CLS_SRI = 4,
    CLS_TSANG = 5,
    CLS_PLOG = 6,
    CLS_CHEB = 7,
    CLS_COUNT = 8
};

// 16-byte per-reaction record (one 128-bit load).
struct RxnRec {
    uint16_t r[3];  // reactant molecule slots (species index per unit molecule), 0xFFFF unused;
                    // if nr == SIDE_GENERAL, r[0] = index into the general-side table
    uint16_t p[3];  // product slots, same convention
    uint8_t nr;     // 0..3 or SIDE_GENERAL
    uint8_t np;
    uint8_t cls;    // RxnClass
    uint8_t flags;  // CTB_RXN_* bits
};
static_assert(sizeof(RxnRec) == 16, "RxnRec must be 16 bytes");
static const uint8_t SIDE_GENERAL = 0xFF;

// Thread-per-state kernels (rates_thread.cuh) walk the execution order as a list of RUNS: maximal
// segments of reactions that share one "fast signature" (plain Arrhenius, no third body, 1-3
// distinct unit-stoichiometry molecules per side or one doubled species, no species on both
// sides) are evaluated by an unrolled, branch-free body; everything else takes the general path.
struct RunRec {
    int32_t begin, end; // execution-order range
    int32_t code;       // fastCode(constK, rev, sideR, sideP) or -1 (general path)
    int32_t pad;
};
static_assert(sizeof(RunRec) == 16, "RunRec must be 16 bytes");
// side patterns of the fast signature
enum SidePat : int { PAT_A = 0, PAT_AB = 1, PAT_2A = 2, PAT_ABC = 3, PAT_COUNT = 4 };
// constK: kf = A (b == 0 and Ea == 0, no exp). Reactant pattern A+B+C is not instantiated.
inline int fastCode(int constK, int rev, int patR, int patP)
{
    if (patR == PAT_ABC) return -1;
    return (((constK * 2 + rev) * PAT_COUNT) + patR) * PAT_COUNT + patP;
}

// Species-major gather entry: wdot_k = sum_e nu_e * q[rxn_e]. nu is signed: > 0 for a
// product-side entry, < 0 for a reactant-side entry (the sign doubles as the role flag for
// creation / destruction rates). Mechanisms with fractional coefficients read the signed
// double array o_gNu instead (nu is then only the sign).
struct GatherEnt {
    uint16_t rxn; // execution index of the reaction
    int16_t nu;
};
static_assert(sizeof(GatherEnt) == 4, "GatherEnt must be 4 bytes");

// Byte offsets into the blob. Plain POD: passed to kernels by value.
struct MechView {
    int32_t K, R;
    int32_t nTb;      // reactions with a third body
    int32_t nFall;    // falloff reactions
    int32_t nPlogRx, nChebRx, nGen; // counts of plog / chebyshev reactions and general sides
    int32_t nGather;  // total gather entries
    int32_t clsBegin[CLS_COUNT + 1]; // execution-order ranges per class
    int32_t hasFractional;           // any non-integer nu in the gather lists
    int32_t maxChebNT;
    double p_ref;
    // species tables
    uint32_t o_molwt, o_rmolwt, o_tmid, o_nasa_low, o_nasa_high; // double[K], [K], [K], [K*7], [K*7]
    // reaction tables, execution order
    uint32_t o_A, o_b, o_EaR, o_mult, o_dn; // double[R]
    uint32_t o_rec;                         // RxnRec[R]
    uint32_t o_perm;                        // int32[R]: execution index -> reference index
    uint32_t o_tbIdx;                       // int16[R]: third-body ordinal or -1
    uint32_t o_auxIdx;                      // int16[R]: falloff / plog / cheb ordinal or -1
    // third-body table
    uint32_t o_tbDefault;                   // double[nTb]
    uint32_t o_tbPtr;                       // int32[nTb+1]
    uint32_t o_tbSp;                        // uint16[nTbEnt]
    uint32_t o_tbEff;                       // double[nTbEnt]  (eff - default)
    // the same efficiencies transposed (species-major), for Jacobian columns
    uint32_t o_tbTPtr;                      // int32[K+1]
    uint32_t o_tbTRxn;                      // uint16[nTbEnt] execution index of the reaction
    uint32_t o_tbTEff;                      // double[nTbEnt] (eff - default)
    // falloff table
    uint32_t o_fallLow;                     // double[nFall*3]: A, b, Ea/R of the low-pressure limit
    uint32_t o_fallCoef;                    // double[nFall*5]
    // plog: per reaction group table
    uint32_t o_plogPtr;                     // int32[nPlogRx+1] into the entry arrays
    uint32_t o_plogLogP, o_plogA, o_plogB, o_plogEaR; // double[nPlogEnt]
    // chebyshev
    uint32_t o_chebPtr;                     // int32[nChebRx+1] into coefficient data
    uint32_t o_chebDim;                     // int32[nChebRx*2]: nT, nP
    uint32_t o_chebLim;                     // double[nChebRx*4]: TrNum, TrDen, PrNum, PrDen
    uint32_t o_chebCoef;                    // double[...]
    // general (C_AnyN) sides
    uint32_t o_genPtr;                      // int32[nGen+1]
    uint32_t o_genSp;                       // uint16[nGenEnt]
    uint32_t o_genStoich, o_genOrder;       // double[nGenEnt]
    // gather lists
    uint32_t o_gPtr;                        // int32[K+1]
    uint32_t o_gEnt;                        // GatherEnt[nGather]
    uint32_t o_gNu;                         // double[nGather] (only read if hasFractional)
    uint32_t o_gCol;                        // uint16[nGather]: species (list owner) of gather entry i
    uint32_t o_tbTCol;                      // uint16[nTbEnt]: species of transposed third-body entry i
    // species of the gather, dealt to 8 lanes with balanced entry counts (longest list first):
    uint32_t o_gl8Ptr;                      // int32[9]
    uint32_t o_gl8Sp;                       // uint16[K]
    uint32_t o_runs;                        // RunRec[nRuns]
    int32_t nRuns;
    uint32_t o_view;                        // a copy of this MechView inside the hot region
    uint32_t bytes;                         // blob size, multiple of 16
    uint32_t hotBytes;                      // prefix staged into shared memory (multiple of 16)
};

// ------------------------------------------------------------------------------------------------
// Tile region: the tables of the state-tile kernels (rates_tile.cuh). A TILE is CTB_TS states whose
// per-species / per-reaction values sit in shared memory as rows [index][CTB_TS]; the lanes of a
// half-warp are the 16 states of one tile working on ONE reaction (or species), so every table
// read below is a broadcast and every row access is conflict-free. The region is a second,
// self-contained, 128-byte aligned segment of the blob (offsets relative to its start) that the
// tile kernels stage into shared memory with one TMA bulk copy.
// ------------------------------------------------------------------------------------------------
#define CTB_TS 16               // states per tile
#define CTB_LDC (CTB_TS + 1)    // padded leading dimension of the c / wdot rows (transposed global I/O)

// species record, 128 bytes: NASA7 low range, Tmid, high range, 1/W (NasaPoly2.h:105-112)
struct Sp2 {
    double low[7];
    double tmid;
    double high[7];
    double rmolwt;
};
static_assert(sizeof(Sp2) == 128, "Sp2 must be 128 bytes");

// reaction record, 64 bytes (four 128-bit words; a fast run reads the first two or three). A
// already carries the reaction's multiplier for the Arrhenius classes. The species slots are BYTE
// OFFSETS of the species' row in the c / u / v arrays (k * CTB_LDC * 8) so that an address is one
// add; a SIDE_GENERAL side keeps the index of its general-side entry in slot 0.
struct Rx2 {
    double A, b;              // word 0
    double EaR;               // word 1
    uint32_t r0, r1;
    uint32_t r2, p0, p1, p2;  // word 2
    uint8_t nr, np, cls, flags; // word 3
    int16_t tbIdx, auxIdx;
    uint32_t auxMultIdx;      // ordinal into the auxMult table (non-Arrhenius classes)
    uint32_t hints;           // RX2_HINT_*
};
//! every side has at most three unit slots and all slots of the reaction name different species: the
//! read-modify-writes of its stoichiometric column are independent (jac_tile.cuh scatterTile)
#define RX2_HINT_DISTINCT 1u
static_assert(sizeof(Rx2) == 64, "Rx2 must be 64 bytes");

// gather record of one species, 16 bytes: its entries are QUADS of 32-bit byte offsets of q rows
// (execution index * CTB_TS * 8), nu = +1 entries first, then nu = -1 entries; |nu| = 2, 3 repeat the
// entry, counts are padded to a multiple of four with the offset of the all-zero row q[R].
// Everything else (fractional or large coefficients) goes through the gen list with an explicit double.
struct GSp2 {
    uint32_t start;          // byte offset of the first quad inside the entry table (multiple of 16)
    uint16_t nPlus4, nMinus4; // number of QUADS
    uint16_t k;              // species
    uint16_t nGen;
    uint32_t genStart;       // index of the first gen entry
};
static_assert(sizeof(GSp2) == 16, "GSp2 must be 16 bytes");
struct GGen2 {
    uint32_t off;
    uint32_t pad;
    double nu;
};
static_assert(sizeof(GGen2) == 16, "GGen2 must be 16 bytes");

#define CTB_TILE_DEALS 2 // gather deals for 8 and 16 reaction lanes per state
struct TileView {
    int32_t K, R, nRuns, nTb;
    uint32_t begin;   // offset of the region inside the blob (multiple of 128)
    uint32_t bytes;   // size of the region (multiple of 16)
    uint32_t magicK;  // ceil(2^32 / K): i / K == __umulhi(i, magicK) for i < 65536
    int32_t hasGen;   // any GGen2 entries
    int32_t hasPdep;  // any Plog / Chebyshev reaction (the kernels then form log P)
    int32_t pad0;
    double lnRp0;     // log(GasConstant / p_ref): g' = (h/RT - s/R) + log T + lnRp0
    double maxAbsB, maxAbsEaR; // over all Arrhenius parameter sets: bound of |b log T - (Ea/R)/T| per state
    uint32_t o_sp;    // Sp2[K]
    uint32_t o_molwt; // double[K] molecular weights (reactor kernels)
    uint32_t magicN;  // ceil(2^32 / (K + 1))
    uint32_t o_rx;    // Rx2[R]
    uint32_t o_runs;  // RunRec[nRuns]
    uint32_t o_gEnt;  // uint32 pairs
    uint32_t o_gGen;  // GGen2[]
    uint32_t o_gSp[CTB_TILE_DEALS];   // GSp2[K], species in lane order of the deal
    uint32_t o_gLane[CTB_TILE_DEALS]; // int32[NL + 1]
    uint32_t o_tbDefault, o_tbPtr, o_tbSp, o_tbEff; // third bodies (tbSp: uint16 species)
    uint32_t o_fallLow, o_fallCoef, o_auxMult;      // falloff tables; multiplier of non-Arrhenius classes by aux ordinal
    uint32_t o_plogPtr, o_plogLogP, o_plogA, o_plogB, o_plogEaR;
    uint32_t o_chebPtr, o_chebDim, o_chebLim, o_chebCoef;
    uint32_t o_genPtr, o_genSp, o_genStoich, o_genOrder;
    // ---- tables of the Jacobian tile kernel only (jac_tile.cuh), laid out directly after the tables
    // above: it stages [begin, begin + bytesJac), the other tile kernels only [begin, begin + bytes)
    uint32_t bytesJac;
    uint32_t o_colPtr;   // int32[K+1]: species-major (reaction, nu) entries = the items of Jacobian column j
    uint32_t o_colEnt;   // GatherEnt[]
    uint32_t o_tbTPtr;   // int32[K+1]: explicit third-body efficiencies transposed (species-major)
    uint32_t o_tbTRxn;   // uint16[]
    uint32_t o_tbTW;     // uint16[] third-body ordinal (row of the we scratch) of the same entries
    uint32_t o_tbTEff;   // double[] (eff - default)
    uint32_t o_colLane[2];  // int32[NL+1] for NL = 4, 8: columns dealt to the lanes, balanced by item count
    uint32_t o_colOrder[2]; // uint16[K]
    uint32_t pad1;
};

struct PackedMech {
    MechView v;
    TileView tv;
    std::vector<unsigned char> blob;
    std::vector<int32_t> perm;    // execution -> reference
    std::vector<int32_t> invPerm; // reference -> execution
};

//! Build the packed image. Throws std::runtime_error on unsupported input (K > 65534 ...).
PackedMech packMechanism(const ctb_mech_tables& t);

} // namespace ctb
