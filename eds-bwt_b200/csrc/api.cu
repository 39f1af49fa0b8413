// C ABI of libedsbwt_cuda.so -- compute entry points (include/edsbwt.h).  No CPU fallback anywhere: every
// function that computes needs a CUDA device and fails with EDSBWT_E_CUDA otherwise.
#include <algorithm>
#include <functional>
#include <thread>

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
    catch (const OccOverflow&) { set_error("more than 2^32 occurrences in one search call; split the batch"); return EDSBWT_E_UNSUPPORTED; }
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
    edsbwt_index* h = nullptr;
    rc = guarded([&] {
        h = new edsbwt_index;
        Index& ix = h->ix;
        ix.device = device; ix.base = base; ix.cache_mode = o.cache;
        ix.move_a_explicit = o.move_a > 0; ix.move_a = o.move_a > 0 ? (uint32_t)o.move_a : 8u;
        if (o.cache && image_cache_load(ix, o)) {          // <base>.edsbwt written by an earlier open of these very files
            ix.image_from_file = true;
            init_index_runtime(ix, o);
            if (o.build_move) build_move_table(ix);
        } else {
            load_host_index(base, o.verify != 0, ix.host);
            build_device_index(ix, o);
            if (o.cache) image_cache_save(ix);
        }
    });
    if (rc) { cudaSetDevice(device); delete h; return rc; }
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
    if (h.sigma <= 5) for (uint32_t i = 0; i < h.sigma * h.sigma; ++i) info->table_occ[i] = h.table_occ[i];
    for (uint32_t i = 0; i < h.sigma * h.sigma && i < 64; ++i) info->table_occ_wide[i] = h.table_occ[i];
    info->engine = (uint32_t)idx->ix.engine;
    info->seed_k = idx->ix.seed_k; info->locate_mode = (uint32_t)idx->ix.locate_mode;
    info->device_bytes = idx->ix.device_bytes(); info->seed_intervals = idx->ix.seed_intervals;
    info->from_file = (idx->ix.image_from_file ? 1u : 0u) | (idx->ix.move_from_file ? 2u : 0u);
    if (const MoveTable* mt = idx->ix.move) { info->move_runs = mt->r_runs; info->move_intervals = mt->r; info->move_a = mt->a; info->move_rounds = mt->rounds; }
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
    return edsbwt_search_device_at(idx, d_pat_bytes, d_pat_off, npat, total_bytes, max_len, locate, lf_variant, 0, res);
}
int edsbwt_search_device_at(edsbwt_index* idx, const uint8_t* d_pat_bytes, const uint64_t* d_pat_off, uint32_t npat,
                            uint64_t total_bytes, uint32_t max_len, int locate, int lf_variant, uint32_t pat_base, edsbwt_result** res) {
    if (!idx || !res || (npat && (!d_pat_off))) return EDSBWT_E_ARG;
    *res = nullptr;
    edsbwt_result* r = nullptr;
    int rc = guarded([&] {
        r = new edsbwt_result; r->device = idx->ix.device; r->pool = idx->ix.pool; r->outp = r->pool->get();
        r->outp->want_alive = false;
        run_search(idx->ix, d_pat_bytes, d_pat_off, npat, total_bytes, max_len, locate != 0, lf_variant, *r->outp, pat_base);
    });
    if (rc) { cudaSetDevice(idx->ix.device); if (r) { if (r->outp) r->pool->put(r->outp); delete r; } return rc; }
    *res = r;
    return EDSBWT_OK;
}
int edsbwt_result_sizes(const edsbwt_result* res, uint32_t* npat, uint64_t* nocc) {
    if (!res) return EDSBWT_E_ARG;
    if (npat) *npat = res->outp->npat;
    if (nocc) *nocc = res->outp->nocc;
    return EDSBWT_OK;
}
int edsbwt_result_device_ptrs(const edsbwt_result* res, const uint32_t** d_counts, const edsbwt_occ** d_occ) {
    if (!res) return EDSBWT_E_ARG;
    if (d_counts) *d_counts = res->outp->counts.p;
    if (d_occ) *d_occ = res->outp->occ.p;
    return EDSBWT_OK;
}
int edsbwt_result_fetch(const edsbwt_result* res, uint32_t* counts, edsbwt_occ* occ) {
    if (!res) return EDSBWT_E_ARG;
    return guarded([&] {
        CUDA_OK(cudaSetDevice(res->device));
        if (counts && res->outp->npat) CUDA_OK(cudaMemcpy(counts, res->outp->counts.p, (size_t)res->outp->npat * 4, cudaMemcpyDeviceToHost));
        if (occ && res->outp->nocc) CUDA_OK(cudaMemcpy(occ, res->outp->occ.p, (size_t)res->outp->nocc * sizeof(edsbwt_occ), cudaMemcpyDeviceToHost));
    });
}
void edsbwt_result_free(edsbwt_result* res) {
    if (!res) return;
    cudaSetDevice(res->device);
    res->pool->put(res->outp);
    delete res;
}

// Pipelined host-buffer search: H2D of chunk i+1 | search of chunk i | D2H of chunk i-1 on three streams.
// occ may be null (count-only or sizing pass).  *nocc receives the total number of occurrences even when it
// exceeds occ_cap (then EDSBWT_E_NOSPACE is returned and only the first occ_cap tuples were copied).
// streams / events / double buffers of the host-buffer searches, created on first use
static Pipe& pipe_of(Index& ix) {
    if (!ix.pipe) {
        ix.pipe = new Pipe;
        CUDA_OK(cudaStreamCreateWithFlags(&ix.pipe->in, cudaStreamNonBlocking));
        CUDA_OK(cudaStreamCreateWithFlags(&ix.pipe->outs, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) CUDA_OK(cudaEventCreateWithFlags(&ix.pipe->ev_in[i], cudaEventDisableTiming));
        for (int i = 0; i < Pipe::NRES; ++i) CUDA_OK(cudaEventCreateWithFlags(&ix.pipe->ev_out[i], cudaEventDisableTiming));
    }
    return *ix.pipe;
}

// One batch in the three forms the host-buffer entry points accept: bytes + offsets, equal-length bytes, 2-bit words.
struct HostBatch {
    const uint8_t* bytes = nullptr; const uint64_t* off = nullptr;      // off == nullptr: fixed_len characters per pattern
    uint32_t fixed_len = 0;
    const uint64_t* words = nullptr; const uint32_t* invalid = nullptr; // packed: stride = ceil(fixed_len / 32) words per pattern
    uint32_t npat = 0;
    uint32_t stride() const { return (fixed_len + 31) / 32; }
    uint64_t off_of(uint32_t i) const { return off ? off[i] : (uint64_t)i * fixed_len; }
};

// Drains the pipeline streams when a call leaves by an error path, so that no DMA is still reading or writing the caller's
// buffers after the call has returned (and the next call finds the slots idle).
struct PipeDrain {
    Index& ix; Pipe& P; bool armed = true;
    ~PipeDrain() { if (armed) { cudaStreamSynchronize(P.in); cudaStreamSynchronize(ix.stream); cudaStreamSynchronize(P.outs); } }
};

// pat_off == nullptr: npat patterns of fixed_len bytes back to back (no offset array crosses PCIe); words != nullptr: the
// caller's own 2-bit words (a quarter of the bytes).  grow != nullptr: *grow is a malloc'ed occurrence buffer of occ_cap
// tuples that is enlarged as needed (edsbwt_search) instead of failing with EDSBWT_E_NOSPACE.
static int search_into_impl(edsbwt_index* idx, const HostBatch& hb, int locate, int lf_variant,
                            uint32_t* counts, edsbwt_occ* occ, uint64_t occ_cap, uint64_t* nocc, edsbwt_occ** grow = nullptr) {
    const uint32_t npat = hb.npat;
    Index& ix = idx->ix;
    std::lock_guard<std::mutex> call_lock(ix.call_mu);
    CUDA_OK(cudaSetDevice(ix.device));
    Pipe& P = pipe_of(ix);
    PipeDrain drain{ix, P};
    // chunk schedule: equal chunks, then a tapered tail (halves down to a quarter chunk) -- the input copies are the long
    // pole of the pipeline, so what the batch pays on top of them is the search + copy-out of the LAST chunk only.
    // Packed batches are cut at multiples of 32 patterns (their invalid bitmap is addressed by 32-bit words).
    const bool packed = hb.words != nullptr;
    auto r32 = [&](uint32_t v) { return packed ? (uint32_t)std::min<uint64_t>(((uint64_t)v + 31) & ~31ull, 0xffffffe0ull) : v; };
    const uint32_t chunk = ix.chunk_patterns ? std::max<uint32_t>(r32(ix.chunk_patterns), 1u) : npat;
    std::vector<uint32_t> cb{0};
    // ... and a ramp at the head (an eighth, an eighth, a quarter, a half of a chunk): with 2-bit words the copy-OUT is the long
    // pole (204 MB of counts and tuples against 81 MB of words per 10 M patterns), and it cannot start before the first chunk
    // has been copied in and searched
    uint32_t ramp = (packed && npat > 2 * (uint64_t)chunk && chunk >= 1024) ? 0 : 4;
    static const uint32_t ramp_div[4] = {8, 8, 4, 2};
    for (uint32_t rem = npat; rem;) {
        uint32_t take;
        if (ramp < 4 && rem > chunk) take = std::max<uint32_t>(r32(chunk / ramp_div[ramp++]), 32u);
        else if (rem > chunk + chunk / 2) take = chunk;
        else if (npat > chunk && rem > chunk / 4 + 1) take = std::min(rem, std::max<uint32_t>(r32((rem + 1) / 2), 1u));
        else take = rem;
        cb.push_back(cb.back() + take); rem -= take;
    }
    const uint32_t nchunks = (uint32_t)cb.size() - 1;
    uint64_t total_occ = 0; bool nospace = false;
    edsbwt_stats acc{};
    auto issue_in = [&](uint32_t c) {
        const uint32_t b = cb[c], e = cb[c + 1];
        const int sl = c & 1;
        if (packed) {
            const size_t nw = (size_t)(e - b) * hb.stride();
            P.words[sl].ensure(nw + 1); P.invalid[sl].ensure((size_t)(e - b + 31) / 32 + 1);
            if (nw) CUDA_OK(cudaMemcpyAsync(P.words[sl].p, hb.words + (size_t)b * hb.stride(), nw * 8, cudaMemcpyHostToDevice, P.in));
            if (hb.invalid) CUDA_OK(cudaMemcpyAsync(P.invalid[sl].p, hb.invalid + b / 32, (size_t)((e - b + 31) / 32) * 4, cudaMemcpyHostToDevice, P.in));
        } else {
            const uint64_t nb = hb.off_of(e) - hb.off_of(b);
            P.bytes[sl].ensure(nb + 16);
            if (nb) CUDA_OK(cudaMemcpyAsync(P.bytes[sl].p, hb.bytes + hb.off_of(b), nb, cudaMemcpyHostToDevice, P.in));
            if (hb.off) { P.off[sl].ensure((size_t)(e - b) + 1); CUDA_OK(cudaMemcpyAsync(P.off[sl].p, hb.off + b, ((size_t)(e - b) + 1) * 8, cudaMemcpyHostToDevice, P.in)); }
        }
        CUDA_OK(cudaEventRecord(P.ev_in[sl], P.in));
    };
    // search of patterns [lo, hi) of chunk [b, ..) whose input sits in slot sl; on an occurrence overflow the range is halved
    // (results go to slot rs: Pipe::NRES of them, so that the search is not held up by a copy-out that lags two chunks behind)
    std::function<void(int, int, uint32_t, uint32_t, uint32_t)> search_range = [&](int sl, int rs, uint32_t b, uint32_t lo, uint32_t hi) {
        PatHint hint{};
        hint.after = P.outs;          // the copy-out below waits for the locate kernels on the device, not here
        const uint8_t* d_bytes = nullptr; const uint64_t* d_off = nullptr; uint64_t nbytes = 0;
        if (packed) { hint.fixed_len = hb.fixed_len; hint.d_words = P.words[sl].p + (size_t)(lo - b) * hb.stride(); hint.d_invalid = hb.invalid ? P.invalid[sl].p + (lo - b) / 32 : nullptr; }
        else {
            // offsets stay absolute: bias the byte pointer instead of rebasing npat offsets on the host
            d_bytes = (const uint8_t*)((uintptr_t)P.bytes[sl].p - (uintptr_t)hb.off_of(b));
            nbytes = hb.off_of(hi) - hb.off_of(lo);
            if (hb.off) d_off = P.off[sl].p + (lo - b);
            else { hint.fixed_len = hb.fixed_len; hint.fixed_base = hb.off_of(lo); }
        }
        try {
            run_search(ix, d_bytes, d_off, hi - lo, nbytes, 0, locate != 0, lf_variant, P.res[rs], lo, &hint);
        } catch (const OccOverflow&) {
            uint32_t mid = lo + (hi - lo) / 2;
            if (packed) mid = b + (((mid - b) + 31) & ~31u);
            if (mid <= lo || mid >= hi) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "patterns %u..%u alone have more than %llu occurrences", lo, hi - 1, (unsigned long long)ix.occ_limit);
            search_range(sl, rs, b, lo, mid);
            CUDA_OK(cudaStreamSynchronize(P.outs));              // the slot's result buffers are reused by the second half
            search_range(sl, rs, b, mid, hi);
            return;
        }
        const edsbwt_stats& st = ix.last;
        acc.I += st.I; acc.S += st.S; acc.E += st.E; acc.L += st.L; acc.O += st.O; acc.steps += st.steps; acc.kernel_launches += st.kernel_launches;
        acc.lf_intervals += st.lf_intervals; acc.lf_kernel_ms += st.lf_kernel_ms; acc.locate_kernel_ms += st.locate_kernel_ms; acc.total_ms += st.total_ms;
        acc.frontier_peak = std::max(acc.frontier_peak, st.frontier_peak); acc.occ_blocks += st.occ_blocks; acc.overflow_patterns += st.overflow_patterns;
        acc.levelsync_patterns += st.levelsync_patterns; acc.search_kernel_ms += st.search_kernel_ms;
        // run_search returned with the sizes known and P.outs waiting for the chunk's last kernels: copy out on that stream
        CUDA_OK(cudaMemcpyAsync(counts + lo, P.res[rs].counts.p, (size_t)(hi - lo) * 4, cudaMemcpyDeviceToHost, P.outs));
        const uint64_t k = P.res[rs].nocc;
        if (locate && k) {
            if (grow && total_occ + k > occ_cap) {               // edsbwt_search: enlarge the malloc'ed buffer (copies into the old one must land first)
                CUDA_OK(cudaStreamSynchronize(P.outs));
                const uint64_t ncap = std::max<uint64_t>(occ_cap * 2, total_occ + k);
                edsbwt_occ* nb = (edsbwt_occ*)realloc(*grow, (size_t)ncap * sizeof(edsbwt_occ));
                if (!nb) EDS_FAIL(EDSBWT_E_IO, "host out of memory for %llu occurrences", (unsigned long long)ncap);
                *grow = occ = nb; occ_cap = ncap;
            }
            if (occ && total_occ + k <= occ_cap) CUDA_OK(cudaMemcpyAsync(occ + total_occ, P.res[rs].occ.p, (size_t)k * sizeof(edsbwt_occ), cudaMemcpyDeviceToHost, P.outs));
            else if (occ && total_occ < occ_cap) { CUDA_OK(cudaMemcpyAsync(occ + total_occ, P.res[rs].occ.p, (size_t)(occ_cap - total_occ) * sizeof(edsbwt_occ), cudaMemcpyDeviceToHost, P.outs)); nospace = true; }
            else nospace = true;
            total_occ += k;
        }
    };
    issue_in(0);
    for (uint32_t c = 0; c < nchunks; ++c) {
        const uint32_t b = cb[c], e = cb[c + 1];
        const int sl = c & 1, rs = (int)(c % Pipe::NRES);
        if (c + 1 < nchunks) issue_in(c + 1);                       // overlaps the search below
        if (c >= (uint32_t)Pipe::NRES) CUDA_OK(cudaEventSynchronize(P.ev_out[rs]));     // result buffers of chunk c - NRES are free again
        CUDA_OK(cudaStreamWaitEvent(ix.stream, P.ev_in[sl], 0));
        search_range(sl, rs, b, b, e);
        CUDA_OK(cudaEventRecord(P.ev_out[rs], P.outs));
    }
    CUDA_OK(cudaStreamSynchronize(P.outs));
    drain.armed = false;
    ix.last = acc;
    if (nocc) *nocc = total_occ;
    return (locate && nospace && (occ || occ_cap)) ? EDSBWT_E_NOSPACE : EDSBWT_OK;
}

int edsbwt_search_into(edsbwt_index* idx, const uint8_t* pat_bytes, const uint64_t* pat_off, uint32_t npat, int locate, int lf_variant,
                       uint32_t* counts, edsbwt_occ* occ, uint64_t occ_cap, uint64_t* nocc) {
    if (!idx || (npat && (!pat_off || !counts))) return EDSBWT_E_ARG;
    if (nocc) *nocc = 0;
    if (npat == 0) return EDSBWT_OK;
    int rc2 = EDSBWT_OK;
    HostBatch hb; hb.bytes = pat_bytes; hb.off = pat_off; hb.npat = npat;
    int rc = guarded([&] { rc2 = search_into_impl(idx, hb, locate, lf_variant, counts, occ, occ_cap, nocc); });
    return rc ? rc : rc2;
}

int edsbwt_search_into_fixed(edsbwt_index* idx, const uint8_t* pat_bytes, uint32_t pat_len, uint32_t npat, int locate, int lf_variant,
                             uint32_t* counts, edsbwt_occ* occ, uint64_t occ_cap, uint64_t* nocc) {
    if (!idx || (npat && (!pat_bytes || !counts || !pat_len))) return EDSBWT_E_ARG;
    if (nocc) *nocc = 0;
    if (npat == 0) return EDSBWT_OK;
    int rc2 = EDSBWT_OK;
    HostBatch hb; hb.bytes = pat_bytes; hb.fixed_len = pat_len; hb.npat = npat;
    int rc = guarded([&] { rc2 = search_into_impl(idx, hb, locate, lf_variant, counts, occ, occ_cap, nocc); });
    return rc ? rc : rc2;
}

int edsbwt_search_into_packed(edsbwt_index* idx, const uint64_t* pat_words, const uint32_t* invalid, uint32_t pat_len, uint32_t npat,
                              int locate, int lf_variant, uint32_t* counts, edsbwt_occ* occ, uint64_t occ_cap, uint64_t* nocc) {
    if (!idx || (npat && (!pat_words || !counts || !pat_len))) return EDSBWT_E_ARG;
    if (nocc) *nocc = 0;
    if (npat == 0) return EDSBWT_OK;
    int rc2 = EDSBWT_OK;
    HostBatch hb; hb.words = pat_words; hb.invalid = invalid; hb.fixed_len = pat_len; hb.npat = npat;
    int rc = guarded([&] { rc2 = search_into_impl(idx, hb, locate, lf_variant, counts, occ, occ_cap, nocc); });
    return rc ? rc : rc2;
}

int edsbwt_pack_patterns(const edsbwt_index* idx, const uint8_t* pat_bytes, uint32_t pat_len, uint32_t npat, int nthreads,
                         uint64_t* words, uint32_t* invalid, uint32_t* n_invalid) {
    if (!idx) return EDSBWT_E_ARG;
    return edsbwt_pack_patterns_alphabet(idx->ix.host.alphabet, idx->ix.host.sigma, pat_bytes, pat_len, npat, nthreads, words, invalid, n_invalid);
}

int edsbwt_search(edsbwt_index* idx, const uint8_t* pat_bytes, const uint64_t* pat_off, uint32_t npat, int locate, int lf_variant,
                  uint32_t* counts, edsbwt_occ** occ, uint64_t* nocc) {
    if (!idx || (npat && (!pat_off || !counts))) return EDSBWT_E_ARG;
    if (locate && (!occ || !nocc)) return EDSBWT_E_ARG;
    if (occ) *occ = nullptr;
    if (nocc) *nocc = 0;
    if (npat == 0) return EDSBWT_OK;
    if (!locate) return edsbwt_search_into(idx, pat_bytes, pat_off, npat, 0, lf_variant, counts, nullptr, 0, nullptr);
    // occurrences are malloc'ed for the caller: a guessed capacity that the chunk loop enlarges when a chunk does not fit
    // (nothing is searched twice)
    const uint64_t cap = (uint64_t)npat * 2 + 1024;
    edsbwt_occ* buf = (edsbwt_occ*)malloc((size_t)cap * sizeof(edsbwt_occ));
    if (!buf) { set_error("host out of memory for %llu occurrences", (unsigned long long)cap); return EDSBWT_E_IO; }
    uint64_t got = 0; int rc2 = EDSBWT_OK;
    HostBatch hb; hb.bytes = pat_bytes; hb.off = pat_off; hb.npat = npat;
    int rc = guarded([&] { rc2 = search_into_impl(idx, hb, 1, lf_variant, counts, buf, cap, &got, &buf); });
    if (rc || rc2) { free(buf); return rc ? rc : rc2; }
    if (got == 0) { free(buf); buf = nullptr; }
    *occ = buf; *nocc = got;
    return EDSBWT_OK;
}

int edsbwt_search_trace(edsbwt_index* idx, const uint8_t* pat_bytes, const uint64_t* pat_off, uint32_t npat, int lf_variant,
                        uint32_t* counts, uint32_t* interrupted) {
    if (!idx || (npat && (!pat_off || !counts || !interrupted))) return EDSBWT_E_ARG;
    if (npat == 0) return EDSBWT_OK;
    return guarded([&] {
        Index& ix = idx->ix;
        CUDA_OK(cudaSetDevice(ix.device));
        const uint64_t nb = pat_off[npat] - pat_off[0];
        const uint32_t max_len = 0;
        DevBuf<uint8_t> d_bytes; DevBuf<uint64_t> d_off; d_bytes.ensure(nb + 16); d_off.ensure((size_t)npat + 1);
        if (nb) CUDA_OK(cudaMemcpy(d_bytes.p, pat_bytes + pat_off[0], nb, cudaMemcpyHostToDevice));
        CUDA_OK(cudaMemcpy(d_off.p, pat_off, ((size_t)npat + 1) * 8, cudaMemcpyHostToDevice));
        SearchOut out; out.want_alive = true;
        run_search(ix, (const uint8_t*)((uintptr_t)d_bytes.p - (uintptr_t)pat_off[0]), d_off.p, npat, nb, max_len, false, lf_variant, out);
        CUDA_OK(cudaMemcpy(counts, out.counts.p, (size_t)npat * 4, cudaMemcpyDeviceToHost));
        CUDA_OK(cudaMemcpy(interrupted, out.alive.p, (size_t)npat * 4, cudaMemcpyDeviceToHost));
        for (uint32_t i = 0; i < npat; ++i) {
            // alive = longest suffix with a non-empty frontier; the reference's loop index when it gives up is m - alive,
            // except that its initial interval of an absent letter survives until the first step (posSymb = m-1)
            const uint32_t m = (uint32_t)(pat_off[i + 1] - pat_off[i]), alive = interrupted[i];
            uint32_t pos = 0;
            if (m >= 2 && counts[i] == 0 && alive < m) { pos = m - alive; if (pos > m - 1) pos = m - 1; }
            interrupted[i] = pos;
        }
    });
}

int edsbwt_set_engine(edsbwt_index* idx, int engine) {
    if (!idx || engine < EDSBWT_ENGINE_FUSED || engine > EDSBWT_ENGINE_ITEMS) return EDSBWT_E_ARG;
    std::lock_guard<std::mutex> lock(idx->ix.mu);
    if (idx->ix.wide && engine == EDSBWT_ENGINE_ITEMS) {
        set_error("the item kernel reads 2-bit pattern words: an alphabet of %u letters is searched by the step-loop kernel or the level-synchronous engine", idx->ix.host.sigma - 1);
        return EDSBWT_E_UNSUPPORTED;
    }
    idx->ix.engine = engine;
    return EDSBWT_OK;
}

int edsbwt_set_row_order(edsbwt_index* idx, int order) {
    if (!idx || (order != EDSBWT_ROWS_GROUPED && order != EDSBWT_ROWS_REFERENCE)) return EDSBWT_E_ARG;
    idx->ix.row_order = order;
    return EDSBWT_OK;
}

static const char* csv_header(int move) { return move ? "#Pat\t$_i\tD[i]\tS_j\tS_j[r] \n" : "#Pat\t$_i\tD[i]\tS_j\tS_j[r]\n"; }

int edsbwt_result_write_csv(edsbwt_index* idx, const edsbwt_result* res, const char* path, int move, int append, uint64_t* bytes) {
    if (!idx || !res || !path) return EDSBWT_E_ARG;
    if (bytes) *bytes = 0;
    FILE* f = fopen(path, append ? "ab" : "wb");
    if (!f) { set_error("ERROR opening file %s to write output", path); return EDSBWT_E_IO; }
    if (!append) fputs(csv_header(move), f);
    int rc = guarded([&] {
        Index& ix = idx->ix;
        CUDA_OK(cudaSetDevice(ix.device));
        std::lock_guard<std::mutex> lock(ix.mu);
        uint64_t w = write_csv_device(ix, res->outp->occ.p, res->outp->nocc, 1, f, ix.stream, nullptr);   // tuples already carry pat_base
        if (bytes) *bytes = w;
    });
    if (fclose(f) != 0 && !rc) { set_error("error closing %s", path); rc = EDSBWT_E_IO; }
    return rc;
}

// search + locate + ordered rows + device formatting, chunk by chunk: H2D of chunk c+1 overlaps the search of chunk c
int edsbwt_search_to_csv(edsbwt_index* idx, const uint8_t* pat_bytes, const uint64_t* pat_off, uint32_t npat, int lf_variant,
                         uint32_t* counts, const char* csv_path, uint64_t* nocc) {
    if (!idx || !csv_path || (npat && (!pat_off || !counts))) return EDSBWT_E_ARG;
    if (nocc) *nocc = 0;
    if (npat == 0) return EDSBWT_OK;
    FILE* f = fopen(csv_path, "ab");
    if (!f) { set_error("ERROR opening file %s to write output", csv_path); return EDSBWT_E_IO; }
    int rc = guarded([&] {
        Index& ix = idx->ix;
        std::lock_guard<std::mutex> call_lock(ix.call_mu);
        CUDA_OK(cudaSetDevice(ix.device));
        Pipe& P = pipe_of(ix);
        const int saved_order = ix.row_order;
        ix.row_order = EDSBWT_ROWS_REFERENCE;
        struct Restore { Index& ix; int o; ~Restore() { ix.row_order = o; } } restore{ix, saved_order};
        const uint32_t chunk = ix.chunk_patterns ? ix.chunk_patterns : npat;
        const uint32_t nchunks = (npat + chunk - 1) / chunk;
        uint64_t total = 0, launches = 0;
        edsbwt_stats acc{};
        PipeDrain drain{ix, P};
        auto issue_in = [&](uint32_t c) {
            const uint32_t b = c * chunk, e = (uint32_t)std::min<uint64_t>((uint64_t)b + chunk, npat);
            const uint64_t nb = pat_off[e] - pat_off[b];
            const int sl = c & 1;
            P.bytes[sl].ensure(nb + 16); P.off[sl].ensure((size_t)(e - b) + 1);
            if (nb) CUDA_OK(cudaMemcpyAsync(P.bytes[sl].p, pat_bytes + pat_off[b], nb, cudaMemcpyHostToDevice, P.in));
            CUDA_OK(cudaMemcpyAsync(P.off[sl].p, pat_off + b, ((size_t)(e - b) + 1) * 8, cudaMemcpyHostToDevice, P.in));
            CUDA_OK(cudaEventRecord(P.ev_in[sl], P.in));
        };
        // patterns [lo, hi) of the chunk starting at b (input in slot sl); a range with more than 2^32 occurrences (short patterns
        // on a large index) is halved until it fits -- the reference handles such files too, one pattern at a time
        std::function<void(int, uint32_t, uint32_t, uint32_t)> search_range = [&](int sl, uint32_t b, uint32_t lo, uint32_t hi) {
            const uint8_t* d_bytes = (const uint8_t*)((uintptr_t)P.bytes[sl].p - (uintptr_t)pat_off[b]);
            try {
                run_search(ix, d_bytes, P.off[sl].p + (lo - b), hi - lo, pat_off[hi] - pat_off[lo], 0, true, lf_variant, P.res[sl], lo);
            } catch (const OccOverflow&) {
                const uint32_t mid = lo + (hi - lo) / 2;
                if (mid <= lo) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "pattern %u alone has more than %llu occurrences", lo, (unsigned long long)ix.occ_limit);
                search_range(sl, b, lo, mid);
                search_range(sl, b, mid, hi);
                return;
            }
            const edsbwt_stats st = ix.last;
            acc.L += st.L; acc.O += st.O; acc.kernel_launches += st.kernel_launches; acc.total_ms += st.total_ms;
            acc.search_kernel_ms += st.search_kernel_ms; acc.locate_kernel_ms += st.locate_kernel_ms; acc.overflow_patterns += st.overflow_patterns;
            CUDA_OK(cudaMemcpyAsync(counts + lo, P.res[sl].counts.p, (size_t)(hi - lo) * 4, cudaMemcpyDeviceToHost, ix.stream));
            {
                std::lock_guard<std::mutex> lock(ix.mu);
                write_csv_device(ix, P.res[sl].occ.p, P.res[sl].nocc, 1, f, ix.stream, &launches);
            }
            CUDA_OK(cudaStreamSynchronize(ix.stream));
            total += P.res[sl].nocc;
        };
        issue_in(0);
        for (uint32_t c = 0; c < nchunks; ++c) {
            const uint32_t b = c * chunk, e = (uint32_t)std::min<uint64_t>((uint64_t)b + chunk, npat);
            const int sl = c & 1;
            if (c + 1 < nchunks) issue_in(c + 1);                   // the other slot is free (chunk c-1 fully written): overlaps the search below
            CUDA_OK(cudaStreamWaitEvent(ix.stream, P.ev_in[sl], 0));
            search_range(sl, b, b, e);
        }
        drain.armed = false;
        acc.kernel_launches += launches;
        ix.last = acc;
        if (nocc) *nocc = total;
    });
    if (fclose(f) != 0 && !rc) { set_error("error closing %s", csv_path); rc = EDSBWT_E_IO; }
    return rc;
}

int edsbwt_move_table_build(edsbwt_index* idx) {
    if (!idx) return EDSBWT_E_ARG;
    return guarded([&] { std::lock_guard<std::mutex> lock(idx->ix.mu); build_move_table(idx->ix); });
}
int edsbwt_move_table_save(edsbwt_index* idx, const char* path) {
    if (!idx) return EDSBWT_E_ARG;
    int rc = edsbwt_move_table_build(idx);
    if (rc) return rc;
    return guarded([&] { std::lock_guard<std::mutex> lock(idx->ix.mu); move_cache_save(idx->ix, path ? std::string(path) : idx->ix.base + "_MLF.edsbwt"); });
}
int edsbwt_move_table_fetch(edsbwt_index* idx, uint32_t* p, uint32_t* q, uint32_t* dest, uint8_t* sym, uint32_t cap, uint32_t* r) {
    if (!idx || !r) return EDSBWT_E_ARG;
    int rc = edsbwt_move_table_build(idx);
    if (rc) return rc;
    return guarded([&] {
        const MoveTable& mt = *idx->ix.move;
        *r = mt.r;
        if (!p && !q && !dest && !sym) return;
        if (cap < mt.r) EDS_FAIL(EDSBWT_E_NOSPACE, "move table has %u intervals, caller arrays hold %u", mt.r, cap);
        CUDA_OK(cudaSetDevice(idx->ix.device));
        std::vector<uint4> h(mt.r);
        CUDA_OK(cudaMemcpy(h.data(), mt.T.p, (size_t)mt.r * sizeof(uint4), cudaMemcpyDeviceToHost));
        for (uint32_t i = 0; i < mt.r; ++i) {
            if (p) p[i] = h[i].x;
            if (q) q[i] = h[i].y;
            if (dest) dest[i] = h[i].z;
            if (sym) sym[i] = idx->ix.host.alphabet[h[i].w];
        }
    });
}

int edsbwt_host_alloc(void** p, uint64_t bytes) {
    if (!p) return EDSBWT_E_ARG;
    *p = nullptr;
    return guarded([&] { CUDA_OK(cudaHostAlloc(p, bytes ? bytes : 1, cudaHostAllocPortable)); });
}
void edsbwt_host_free(void* p) { if (p) cudaFreeHost(p); }
int edsbwt_set_chunk(edsbwt_index* idx, uint32_t patterns_per_chunk) {
    if (!idx) return EDSBWT_E_ARG;
    idx->ix.chunk_patterns = patterns_per_chunk;
    return EDSBWT_OK;
}

int edsbwt_gather_bench(int device, uint64_t buffer_bytes, uint64_t n_access, int iters, double* gb_per_s) {
    if (!gb_per_s) return EDSBWT_E_ARG;
    int rc = require_device(device);
    if (rc) return rc;
    return guarded([&] { *gb_per_s = gather_bench(device, buffer_bytes, n_access, iters); });
}

int edsbwt_gather_bench_ex(int device, uint64_t buffer_bytes, uint64_t n_access, int iters, int granule_bytes, double* gb_per_s) {
    if (!gb_per_s) return EDSBWT_E_ARG;
    int rc = require_device(device);
    if (rc) return rc;
    return guarded([&] { *gb_per_s = gather_bench(device, buffer_bytes, n_access, iters, granule_bytes); });
}

int edsbwt_gather_bench_cfg(int device, const edsbwt_gather_cfg* cfg, double* gb_per_s, double* req_per_s) {
    if (!cfg || !gb_per_s) return EDSBWT_E_ARG;
    int rc = require_device(device);
    if (rc) return rc;
    return guarded([&] {
        GatherCfg c; c.buffer_bytes = cfg->buffer_bytes; c.n_access = cfg->n_access; c.iters = cfg->iters; c.granule = cfg->granule_bytes; c.ilp = cfg->ilp;
        c.ctas_per_sm = cfg->ctas_per_sm; c.dependent = cfg->dependent; c.mix_bytes = cfg->mix_bytes; c.mix_frac = cfg->mix_frac;
        *gb_per_s = gather_bench_cfg(device, c, req_per_s);
    });
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
