// MOVE_EDSBWTSearch <indexBase> <patternsFile> -- drop-in for the reference executable built from mainMove_EDSBWT.cpp:
// same search with LF done through a balanced move table (built on the GPU from the pile files at load: the
// reference's <base>_MLF.aux is a Move-r private serialisation that cannot be read without Move-r).
#include "cli_common.hpp"
int main(int argc, char** argv) { return edsbwt_cli::run_cli(argc, argv, true); }
