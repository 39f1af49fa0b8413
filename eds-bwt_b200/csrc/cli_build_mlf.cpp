// build_MLF <index> <a>: builds the balanced move table of MOVE_EDSBWTSearch on the GPU and writes it next to the index
// (<index>_MLF.edsbwt), the counterpart of the reference's build_MLF (build_MLF.cpp:53-164: argv, "<index>_MLF.aux").
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>

#include "edsbwt.h"

int main(int argc, char** argv) {
    if (argc != 3) {                                    // build_MLF.cpp:55-60
        std::cerr << "usage: " << argv[0] << " input a" << std::endl;
        exit(1);
    }
    edsbwt_open_opts o;
    memset(&o, 0, sizeof o);
    o.seed_k = 1; o.locate_mode = EDSBWT_LOCATE_WALK; o.verify = 1; o.move_a = atoi(argv[2]) > 1 ? atoi(argv[2]) : 2;
    edsbwt_index* idx = nullptr;
    int rc = edsbwt_index_open_ex(argv[1], getenv("EDSBWT_DEVICE") ? atoi(getenv("EDSBWT_DEVICE")) : 0, &o, &idx);
    if (!rc) rc = edsbwt_move_table_build(idx);
    edsbwt_info info;
    if (!rc) rc = edsbwt_index_info(idx, &info);
    if (!rc && (info.from_file & 2u)) std::cerr << "a valid " << argv[1] << "_MLF.edsbwt exists (a = " << info.move_a << "): rewritten" << std::endl;
    if (!rc) rc = edsbwt_move_table_save(idx, nullptr);
    if (rc) { std::cerr << "build_MLF: " << edsbwt_strerror(rc) << " (" << edsbwt_last_error() << ")" << std::endl; exit(1); }
    std::cout << "runs: " << info.move_runs << " input intervals: " << info.move_intervals << " a: " << info.move_a
              << " balancing rounds: " << info.move_rounds << std::endl;
    std::cout << "M_LF written to " << argv[1] << "_MLF.edsbwt" << std::endl;
    edsbwt_index_close(idx);
    return 0;
}
