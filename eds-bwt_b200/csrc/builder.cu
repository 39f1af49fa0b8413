// placeholder replaced below
#include "internal.cuh"
namespace edsbwt {
void build_index_files(const uint8_t*, const uint64_t*, uint32_t, const uint8_t*, const std::string&, int) {
    EDS_FAIL(EDSBWT_E_UNSUPPORTED, "index builder not compiled in");
}
}
