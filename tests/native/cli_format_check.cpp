// CPU check of the executables' hand-written report formatter (csrc/cli_common.hpp put_found) against the printf format of
// the reference's line (EDSBWTsearch.cpp:746).  Built and run by tests/test_tools.py.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <random>

#include "edsbwt.h"
#include "../../eds-bwt_b200/csrc/cli_common.hpp"

int main() {
    std::mt19937_64 r(1);
    const uint32_t special[] = {0, 1, 9, 10, 99, 100, 999999999u, 1000000000u, 4294967295u};
    for (int t = 0; t < 300000; ++t) {
        const uint32_t a = t < 81 ? special[t / 9] : (uint32_t)(r() >> (r() % 33 + 31));
        const uint32_t b = t < 81 ? special[t % 9] : (uint32_t)(r() >> (r() % 33 + 31));
        char want[160], got[96];
        snprintf(want, sizeof want, "Found %u starting positions of occurrences of pattern number %u\n", a, b);
        char* e = edsbwt_cli::put_found(got, a, b);
        *e = 0;
        if (strcmp(got, want) != 0 || e - got > 80) { printf("MISMATCH %u %u\n[%s][%s]\n", a, b, got, want); return 1; }
    }
    puts("format ok");
    return 0;
}
