// stringCheck <input> <output> -- drop-in for the reference executable built from stringCheck.cpp (main: :94-107):
// writes <output>.eds with every non-degenerate stretch bracketed.  C++ host code over the C ABI.
#include <cstdlib>
#include <iostream>
#include <string>

#include "edsbwt.h"

int main(int argc, char** argv) {
    if (argc != 3) {
        std::cerr << "usage: " << argv[0] << " input" << "output" << std::endl;
        std::cerr << "input is the full filename, output will have \".eds\" appended.\n";
        exit(1);
    }
    const std::string in = argv[1], out = std::string(argv[2]) + ".eds";
    std::cout << "stringCheck on " << in << ", output " << out << std::endl;
    const int rc = edsbwt_string_check(in.c_str(), out.c_str());
    if (rc == EDSBWT_E_FORMAT) {
        std::cerr << "\n\nError: the input contains character Z, which is used by this tool to represent an empty string. Please change EMPTY"
                  << "_CHAR in Parameters.h or change the character in your input.\n\n\n";
        exit(1);
    }
    if (rc) { std::cerr << "Error: cannot open file " << in << " or file " << out << std::endl; exit(1); }
    std::cout << "Done.\n";
    return 0;
}
