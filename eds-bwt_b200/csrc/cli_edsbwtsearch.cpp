// EDSBWTsearch <indexBase> <patternsFile> -- drop-in for the reference executable built from mainEDS-BWT.cpp
// (C++ host code over the C ABI; the batched search runs on the GPU, there is no CPU search path).
#include "cli_common.hpp"
int main(int argc, char** argv) { return edsbwt_cli::run_cli(argc, argv, false); }
