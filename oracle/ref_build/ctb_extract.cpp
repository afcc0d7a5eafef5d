// TEST INFRASTRUCTURE / reference-side tool. Loads a mechanism with the UNMODIFIED reference
// (newSolution, src/base/Solution.cpp:267) and writes the flat tables the CUDA library consumes.
//   ctb_extract <file.yaml> <phase|-> <out.ctbm>
#include "cantera_b200/MechanismExtract.h"
#include "cantera/base/global.h"
#include <iostream>

int main(int argc, char** argv)
{
    if (argc < 4) {
        std::cerr << "usage: ctb_extract file.yaml phase out.ctbm [datadir...]\n";
        return 2;
    }
    try {
        for (int i = 4; i < argc; i++) {
            Cantera::addDataDirectory(argv[i]);
        }
        std::string phase = argv[2];
        if (phase == "-") phase = "";
        auto sol = Cantera::newSolution(argv[1], phase, "none");
        auto m = cantera_b200::tablesFromSolution(*sol);
        m.save(argv[3]);
        std::cout << argv[3] << ": K=" << m.n_species << " R=" << m.n_reactions << "\n";
    } catch (std::exception& e) {
        std::cerr << e.what() << "\n";
        return 1;
    }
    return 0;
}
