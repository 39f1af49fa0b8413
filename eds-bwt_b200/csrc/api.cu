// C ABI of libedsbwt_cuda.so -- compute entry points (include/edsbwt.h).  No CPU fallback anywhere: every
// function that computes needs a CUDA device and fails with EDSBWT_E_CUDA otherwise.
#include "internal.cuh"

using namespace edsbwt;

namespace {
int require_device(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) { set_error("no CUDA device available (%s): libedsbwt_cuda has no CPU path", e == cudaSuccess ? "device count 0" : cudaGetErrorString(e)); return EDSBWT_E_CUDA; }
    if (device < 0 || device >= n) { set_error("device %d out of range (have %d)", device, n); return EDSBWT_E_ARG; }
    return EDSBWT_OK;
}
template <class F> int guarded(F&& f) {
    try { f(); return EDSBWT_OK; }
    catch (const Failure& e) { return e.code; }
    catch (const std::bad_alloc&) { set_error("host out of memory"); return EDSBWT_E_IO; }
    catch (const std::exception& e) { set_error("exception: %s", e.what()); return EDSBWT_E_ARG; }
}
}  // namespace

extern "C" {

int edsbwt_index_open_ex(const char* base, int device, const edsbwt_open_opts* opts, edsbwt_index** out) {
    if (!base || !out) return EDSBWT_E_ARG;
    *out = nullptr;
    int rc = require_device(device);
    if (rc) return rc;
    edsbwt_open_opts o{}; o.seed_k = -1; o.locate_mode = EDSBWT_LOCATE_SA; o.verify = 1;
    if (opts) o = *opts;
    edsbwt_index* h = new edsbwt_index;
    rc = guarded([&] {
        h->ix.device = device;
        load_host_index(base, o.verify != 0, h->ix.host);
        build_device_index(h->ix, o);
    });
    if (rc) { delete h; return rc; }
    *out = h;
    return EDSBWT_OK;
}
int edsbwt_index_open(const char* base, int device, edsbwt_index** out) { return edsbwt_index_open_ex(base, device, nullptr, out); }

void edsbwt_index_close(edsbwt_index* idx) {
    if (!idx) return;
    cudaSetDevice(idx->ix.device);
    delete idx;
}

int edsbwt_index_info(const edsbwt_index* idx, edsbwt_info* info) {
    if (!idx || !info) return EDSBWT_E_ARG;
    const HostIndex& h = idx->ix.host;
    memset(info, 0, sizeof *info);
    info->n = h.n; info->n_text = h.n_text; info->n_seg = h.n_seg; info->sigma = h.sigma;
    memcpy(info->alphabet, h.alphabet, 8);
    for (uint32_t i = 0; i < h.sigma * h.sigma && i < 25; ++i) info->table_occ[i] = h.table_occ[i];
    info->seed_k = idx->ix.seed_k; info->locate_mode = (uint32_t)idx->ix.locate_mode;
    info->device_bytes = idx->ix.device_bytes(); info->seed_intervals = idx->ix.seed_intervals;
    return EDSBWT_OK;
}

int edsbwt_set_stats(edsbwt_index* idx, int collect_counters, uint32_t canonical_k) {
    if (!idx) return EDSBWT_E_ARG;
    idx->ix.collect_counters = collect_counters; idx->ix.canonical_k = canonical_k ? canonical_k : 12;
    return EDSBWT_OK;
}
int edsbwt_last_stats(const edsbwt_index* idx, edsbwt_stats* st) {
    if (!idx || !st) return EDSBWT_E_ARG;
    *st = idx->ix.last;
    return EDSBWT_OK;
}

int edsbwt_search_device(edsbwt_index* idx, const uint8_t* d_pat_bytes, const uint64_t* d_pat_off, uint32_t npat,
                         uint64_t total_bytes, uint32_t max_len, int locate, int lf_variant, edsbwt_result** res) {
    if (!idx || !res || (npat && (!d_pat_off))) return EDSBWT_E_ARG;
    *res = nullptr;
    edsbwt_result* r = new edsbwt_result; r->device = idx->ix.device;
    int rc = guarded([&] { run_search(idx->ix, d_pat_bytes, d_pat_off, npat, total_bytes, max_len, locate != 0, lf_variant, r->out); });
    if (rc) { cudaSetDevice(idx->ix.device); delete r; return rc; }
    *res = r;
    return EDSBWT_OK;
}
int edsbwt_result_sizes(const edsbwt_result* res, uint32_t* npat, uint64_t* nocc) {
    if (!res) return EDSBWT_E_ARG;
    if (npat) *npat = res->out.npat;
    if (nocc) *nocc = res->out.nocc;
    return EDSBWT_OK;
}
int edsbwt_result_device_ptrs(const edsbwt_result* res, const uint32_t** d_counts, const edsbwt_occ** d_occ) {
    if (!res) return EDSBWT_E_ARG;
    if (d_counts) *d_counts = res->out.counts.p;
    if (d_occ) *d_occ = res->out.occ.p;
    return EDSBWT_OK;
}
int edsbwt_result_fetch(const edsbwt_result* res, uint32_t* counts, edsbwt_occ* occ) {
    if (!res) return EDSBWT_E_ARG;
    return guarded([&] {
        CUDA_OK(cudaSetDevice(res->device));
        if (counts && res->out.npat) CUDA_OK(cudaMemcpy(counts, res->out.counts.p, (size_t)res->out.npat * 4, cudaMemcpyDeviceToHost));
        if (occ && res->out.nocc) CUDA_OK(cudaMemcpy(occ, res->out.occ.p, (size_t)res->out.nocc * sizeof(edsbwt_occ), cudaMemcpyDeviceToHost));
    });
}
void edsbwt_result_free(edsbwt_result* res) {
    if (!res) return;
    cudaSetDevice(res->device);
    delete res;
}

int edsbwt_search(edsbwt_index* idx, const uint8_t* pat_bytes, const uint64_t* pat_off, uint32_t npat, int locate, int lf_variant,
                  uint32_t* counts, edsbwt_occ** occ, uint64_t* nocc) {
    if (!idx || (npat && (!pat_off || !counts))) return EDSBWT_E_ARG;
    if (locate && (!occ || !nocc)) return EDSBWT_E_ARG;
    if (occ) *occ = nullptr;
    if (nocc) *nocc = 0;
    if (npat == 0) return EDSBWT_OK;
    edsbwt_result* r = nullptr;
    int rc = guarded([&] {
        CUDA_OK(cudaSetDevice(idx->ix.device));
        uint64_t total = pat_off[npat] - pat_off[0];
        uint32_t max_len = 0;
        for (uint32_t i = 0; i < npat; ++i) { uint64_t l = pat_off[i + 1] - pat_off[i]; if (l > 0xffffffffull) EDS_FAIL(EDSBWT_E_ARG, "pattern %u too long", i); if (l > max_len) max_len = (uint32_t)l; }
        DevBuf<uint8_t> d_bytes; DevBuf<uint64_t> d_off;
        d_bytes.ensure(total + 16); d_off.ensure((size_t)npat + 1);
        std::vector<uint64_t> rel((size_t)npat + 1);
        for (uint32_t i = 0; i <= npat; ++i) rel[i] = pat_off[i] - pat_off[0];
        if (total) CUDA_OK(cudaMemcpy(d_bytes.p, pat_bytes + pat_off[0], total, cudaMemcpyHostToDevice));
        CUDA_OK(cudaMemcpy(d_off.p, rel.data(), rel.size() * 8, cudaMemcpyHostToDevice));
        int rc2 = edsbwt_search_device(idx, d_bytes.p, d_off.p, npat, total, max_len, locate, lf_variant, &r);
        if (rc2) throw Failure{rc2};
        CUDA_OK(cudaMemcpy(counts, r->out.counts.p, (size_t)npat * 4, cudaMemcpyDeviceToHost));
        if (locate) {
            *nocc = r->out.nocc;
            if (r->out.nocc) {
                *occ = (edsbwt_occ*)malloc((size_t)r->out.nocc * sizeof(edsbwt_occ));
                if (!*occ) EDS_FAIL(EDSBWT_E_IO, "host out of memory for %llu occurrences", (unsigned long long)r->out.nocc);
                CUDA_OK(cudaMemcpy(*occ, r->out.occ.p, (size_t)r->out.nocc * sizeof(edsbwt_occ), cudaMemcpyDeviceToHost));
            }
        }
    });
    edsbwt_result_free(r);
    if (rc && occ && *occ) { free(*occ); *occ = nullptr; }
    return rc;
}

int edsbwt_gather_bench(int device, uint64_t buffer_bytes, uint64_t n_access, int iters, double* gb_per_s) {
    if (!gb_per_s) return EDSBWT_E_ARG;
    int rc = require_device(device);
    if (rc) return rc;
    return guarded([&] { *gb_per_s = gather_bench(device, buffer_bytes, n_access, iters); });
}

int edsbwt_build_from_collection(const uint8_t* text, const uint64_t* str_off, uint32_t nstr, const uint8_t* seg_first,
                                 const char* out_base, int device) {
    if (!str_off || !seg_first || !out_base || nstr == 0) return EDSBWT_E_ARG;
    int rc = require_device(device);
    if (rc) return rc;
    return guarded([&] { build_index_files(text, str_off, nstr, seg_first, out_base, device); });
}
int edsbwt_build_from_eds(const char* eds_path, const char* out_base, int device) {
    if (!eds_path || !out_base) return EDSBWT_E_ARG;
    int rc = require_device(device);
    if (rc) return rc;
    return guarded([&] {
        std::vector<uint8_t> letters, seg_first; std::vector<uint64_t> off;
        parse_eds_text(read_file(eds_path), letters, off, seg_first);
        build_index_files(letters.data(), off.data(), (uint32_t)seg_first.size(), seg_first.data(), out_base, device);
    });
}

}  // extern "C"
