// Multi-GPU search inside ONE process: patterns sharded in contiguous blocks over the GPUs of the box, index
// replicated on every GPU, results gathered on the first device with NCCL -- ncclSend / ncclRecv inside one
// ncclGroupStart/End per chunk (a gather-v of the per-pattern counts and of the 20-byte occurrence tuples), issued from this
// C++ code on per-device communication streams.  Every shard is searched in chunks, and the gather of chunk c travels over
// NVLink while chunk c + 1 is being searched, so the collective hides behind the kernels instead of following them.
//
// Replaces the reference's single object on a single thread (mainEDS-BWT.cpp:66-70: `num_threads = 1`, one EDSBWT
// constructed per run); the executables use it when EDSBWT_DEVICES asks for more than one GPU (cli_common.hpp).
// There is no exchange inside the search itself: patterns are independent.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <thread>

#include "internal.cuh"

using namespace edsbwt;

namespace {

// NCCL is bound at run time: the library stays loadable on a box without libnccl (single-GPU use), and a process that already
// holds an NCCL (PyTorch brings its own libnccl.so.2) shares it instead of loading a second copy.
struct NcclApi {
    void* so = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};
NcclApi& nccl_api() {
    static NcclApi api = [] {
        NcclApi a;
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) { a.so = dlopen(name, RTLD_NOW | RTLD_GLOBAL); if (a.so) break; }
        if (!a.so) return a;
        a.CommInitAll = (decltype(a.CommInitAll))dlsym(a.so, "ncclCommInitAll");
        a.CommDestroy = (decltype(a.CommDestroy))dlsym(a.so, "ncclCommDestroy");
        a.GroupStart = (decltype(a.GroupStart))dlsym(a.so, "ncclGroupStart");
        a.GroupEnd = (decltype(a.GroupEnd))dlsym(a.so, "ncclGroupEnd");
        a.Send = (decltype(a.Send))dlsym(a.so, "ncclSend");
        a.Recv = (decltype(a.Recv))dlsym(a.so, "ncclRecv");
        a.GetErrorString = (decltype(a.GetErrorString))dlsym(a.so, "ncclGetErrorString");
        a.ok = a.CommInitAll && a.CommDestroy && a.GroupStart && a.GroupEnd && a.Send && a.Recv && a.GetErrorString;
        return a;
    }();
    return api;
}
#define NCCL_OK(expr) do { ncclResult_t r__ = (expr); if (r__ != ncclSuccess) { \
    ::edsbwt::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, nccl_api().GetErrorString(r__)); \
    throw ::edsbwt::Failure{EDSBWT_E_CUDA}; } } while (0)

template <class F> int guarded_multi(F&& f) {
    try { f(); return EDSBWT_OK; }
    catch (const Failure& e) { return e.code; }
    catch (const OccOverflow&) { set_error("a chunk of a shard has more than 2^32 occurrences; raise chunks_per_gpu"); return EDSBWT_E_UNSUPPORTED; }
    catch (const std::bad_alloc&) { set_error("host out of memory"); return EDSBWT_E_IO; }
    catch (const std::exception& e) { set_error("exception: %s", e.what()); return EDSBWT_E_ARG; }
}

}  // namespace

struct edsbwt_multi {
    std::vector<int> dev;
    std::vector<edsbwt_index*> idx;
    std::vector<ncclComm_t> comm;                 // empty: no NCCL (one GPU, or libnccl missing -> peer copies)
    std::vector<cudaStream_t> cstream, hstream;   // per device: gather stream, host-to-device copy stream
    std::mutex mu;                                // one multi-GPU call at a time
    ~edsbwt_multi() {
        for (size_t g = 0; g < dev.size(); ++g) {
            cudaSetDevice(dev[g]);
            if (g < comm.size() && comm[g]) nccl_api().CommDestroy(comm[g]);
            if (g < cstream.size() && cstream[g]) cudaStreamDestroy(cstream[g]);
            if (g < hstream.size() && hstream[g]) cudaStreamDestroy(hstream[g]);
            if (g < idx.size() && idx[g]) edsbwt_index_close(idx[g]);
        }
    }
};

struct edsbwt_multi_result {
    int root = 0;
    uint32_t npat = 0; uint64_t nocc = 0;
    uint32_t* d_counts = nullptr;                 // [npat] on the root device
    struct Seg { uint32_t pat_lo, pat_hi; uint64_t nocc; edsbwt_occ* d_occ; };
    std::vector<Seg> seg;                         // ascending pattern order: shard-major, chunk-minor
    double search_ms = 0, gather_ms = 0, wall_ms = 0; uint64_t gather_bytes = 0; int nccl = 0;
    ~edsbwt_multi_result() {
        cudaSetDevice(root);
        if (d_counts) cudaFree(d_counts);
        for (auto& s : seg) if (s.d_occ) cudaFree(s.d_occ);
    }
};

extern "C" {

int edsbwt_multi_open(const char* base, const int* devices, int ndev, const edsbwt_open_opts* opts, edsbwt_multi** out) {
    if (!base || !out || ndev < 0 || (ndev > 0 && !devices)) return EDSBWT_E_ARG;
    *out = nullptr;
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess || have == 0) { set_error("no CUDA device available: libedsbwt_cuda has no CPU path"); return EDSBWT_E_CUDA; }
    edsbwt_multi* m = new edsbwt_multi;
    if (ndev == 0) for (int d = 0; d < have; ++d) m->dev.push_back(d);
    else m->dev.assign(devices, devices + ndev);
    const int G = (int)m->dev.size();
    for (int g = 0; g < G; ++g) if (m->dev[g] < 0 || m->dev[g] >= have || std::count(m->dev.begin(), m->dev.end(), m->dev[g]) != 1) {
        set_error("device list: %d is out of range or repeated (have %d devices)", m->dev[g], have); delete m; return EDSBWT_E_ARG; }
    m->idx.assign(G, nullptr);
    // every GPU builds its own HBM image from the same files, side by side
    std::vector<int> rcs(G, 0); std::vector<std::string> errs(G);
    {
        std::vector<std::thread> th;
        for (int g = 0; g < G; ++g) th.emplace_back([&, g] {
            rcs[g] = edsbwt_index_open_ex(base, m->dev[g], opts, &m->idx[g]);
            if (rcs[g]) errs[g] = edsbwt_last_error();
        });
        for (auto& t : th) t.join();
    }
    for (int g = 0; g < G; ++g) if (rcs[g]) { set_error("device %d: %s", m->dev[g], errs[g].c_str()); int rc = rcs[g]; delete m; return rc; }
    int rc = guarded_multi([&] {
        m->cstream.assign(G, nullptr); m->hstream.assign(G, nullptr);
        for (int g = 0; g < G; ++g) {
            CUDA_OK(cudaSetDevice(m->dev[g]));
            CUDA_OK(cudaStreamCreateWithFlags(&m->cstream[g], cudaStreamNonBlocking));
            CUDA_OK(cudaStreamCreateWithFlags(&m->hstream[g], cudaStreamNonBlocking));
            for (int h = 0; h < G; ++h) if (h != g) {             // peer access: the fallback copies and NCCL's P2P transport
                int can = 0; cudaDeviceCanAccessPeer(&can, m->dev[g], m->dev[h]);
                if (can) { cudaError_t e = cudaDeviceEnablePeerAccess(m->dev[h], 0); if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CUDA_OK(e); cudaGetLastError(); }
            }
        }
        const char* p2p = getenv("EDSBWT_MULTI_P2P");             // 1: cudaMemcpyPeerAsync instead of NCCL (A/B, and boxes without libnccl)
        if (G > 1 && !(p2p && p2p[0] == '1') && nccl_api().ok) {
            m->comm.assign(G, nullptr);
            NCCL_OK(nccl_api().CommInitAll(m->comm.data(), G, m->dev.data()));
        }
    });
    if (rc) { delete m; return rc; }
    *out = m;
    return EDSBWT_OK;
}

void edsbwt_multi_close(edsbwt_multi* m) { delete m; }

int edsbwt_multi_info(const edsbwt_multi* m, int* ndev, int* devices, int devices_cap, int* nccl) {
    if (!m) return EDSBWT_E_ARG;
    if (ndev) *ndev = (int)m->dev.size();
    if (devices) for (int g = 0; g < (int)m->dev.size() && g < devices_cap; ++g) devices[g] = m->dev[g];
    if (nccl) *nccl = m->comm.empty() ? 0 : 1;
    return EDSBWT_OK;
}

edsbwt_index* edsbwt_multi_handle(edsbwt_multi* m, int i) { return (m && i >= 0 && i < (int)m->idx.size()) ? m->idx[i] : nullptr; }

int edsbwt_search_multi(edsbwt_multi* m, const uint8_t* pat_bytes, const uint64_t* pat_off, uint32_t fixed_len, uint32_t npat,
                        int locate, int lf_variant, int chunks_per_gpu, edsbwt_multi_result** res) {
    if (!m || !res || (npat && (!pat_bytes || (!pat_off && !fixed_len)))) return EDSBWT_E_ARG;
    *res = nullptr;
    const int G = (int)m->dev.size();
    const uint32_t C = (uint32_t)std::max(1, std::min(chunks_per_gpu <= 0 ? 4 : chunks_per_gpu, 64));
    std::lock_guard<std::mutex> call_lock(m->mu);
    edsbwt_multi_result* R = new edsbwt_multi_result;
    R->root = m->dev[0]; R->npat = npat; R->nccl = m->comm.empty() ? 0 : 1;
    auto off_of = [&](uint64_t i) -> uint64_t { return pat_off ? pat_off[i] : i * fixed_len; };
    // shard g = patterns [lo(g), lo(g+1)), chunk c of it = [clo(g,c), clo(g,c+1))
    auto lo = [&](int g) { return (uint32_t)(((uint64_t)npat * g) / G); };
    auto clo = [&](int g, uint32_t c) { const uint64_t a = lo(g), n = lo(g + 1) - a; return (uint32_t)(a + (n * c) / C); };

    struct Shared {
        std::mutex mu; std::condition_variable cv;
        std::vector<uint32_t> done;           // per device: chunks searched so far
        std::vector<int> rc; std::vector<std::string> err;
    } sh;
    sh.done.assign(G, 0); sh.rc.assign(G, 0); sh.err.assign(G, "");
    // per device: two result slots (chunk c uses slot c & 1) + the event that says "the gather of this slot's chunk has left"
    std::vector<SearchOut> slots((size_t)G * 2);
    std::vector<cudaEvent_t> ev_sent((size_t)G * 2, nullptr);
    std::vector<uint64_t> chunk_nocc((size_t)G * C, 0);
    std::vector<double> search_ms(G, 0.0);
    const auto t_wall0 = std::chrono::steady_clock::now();

    int rc = guarded_multi([&] {
        CUDA_OK(cudaSetDevice(R->root));
        if (npat) CUDA_OK(cudaMalloc(&R->d_counts, (size_t)npat * 4));
        for (int g = 0; g < G; ++g) { CUDA_OK(cudaSetDevice(m->dev[g])); for (int s = 0; s < 2; ++s) CUDA_OK(cudaEventCreateWithFlags(&ev_sent[g * 2 + s], cudaEventDisableTiming)); }
        std::vector<uint32_t> sent(G, 0);      // chunks of device g whose gather has been issued (read by its worker)
        std::vector<std::thread> workers;
        for (int g = 0; g < G; ++g) workers.emplace_back([&, g] {
            int wrc = guarded_multi([&] {
                Index& ix = m->idx[g]->ix;
                CUDA_OK(cudaSetDevice(m->dev[g]));
                DevBuf<uint8_t> d_bytes[2]; DevBuf<uint64_t> d_off[2];
                cudaEvent_t ev_in[2];
                for (int s = 0; s < 2; ++s) CUDA_OK(cudaEventCreateWithFlags(&ev_in[s], cudaEventDisableTiming));
                auto issue_in = [&](uint32_t c) {
                    const uint32_t b = clo(g, c), e = clo(g, c + 1);
                    const uint64_t nb = off_of(e) - off_of(b);
                    const int s = c & 1;
                    d_bytes[s].ensure(nb + 16);
                    if (nb) CUDA_OK(cudaMemcpyAsync(d_bytes[s].p, pat_bytes + off_of(b), nb, cudaMemcpyHostToDevice, m->hstream[g]));
                    if (pat_off) { d_off[s].ensure((size_t)(e - b) + 1); CUDA_OK(cudaMemcpyAsync(d_off[s].p, pat_off + b, ((size_t)(e - b) + 1) * 8, cudaMemcpyHostToDevice, m->hstream[g])); }
                    CUDA_OK(cudaEventRecord(ev_in[s], m->hstream[g]));
                };
                issue_in(0);
                for (uint32_t c = 0; c < C; ++c) {
                    const uint32_t b = clo(g, c), e = clo(g, c + 1);
                    const int s = c & 1;
                    if (c + 1 < C) issue_in(c + 1);
                    if (c >= 2) {                                 // slot s still holds chunk c - 2 until its gather has been issued and has left
                        std::unique_lock<std::mutex> l(sh.mu);
                        sh.cv.wait(l, [&] { return sent[g] >= c - 1 || sh.rc[0] < 0; });
                        l.unlock();
                        CUDA_OK(cudaEventSynchronize(ev_sent[g * 2 + s]));
                    }
                    CUDA_OK(cudaStreamWaitEvent(ix.stream, ev_in[s], 0));
                    PatHint hint{};
                    const uint8_t* db = (const uint8_t*)((uintptr_t)d_bytes[s].p - (uintptr_t)off_of(b));
                    if (!pat_off) { hint.fixed_len = fixed_len; hint.fixed_base = off_of(b); }
                    run_search(ix, db, pat_off ? d_off[s].p : nullptr, e - b, off_of(e) - off_of(b), 0, locate != 0, lf_variant, slots[g * 2 + s], b, &hint);
                    search_ms[g] += ix.last.total_ms;
                    chunk_nocc[(size_t)g * C + c] = slots[g * 2 + s].nocc;
                    { std::lock_guard<std::mutex> l(sh.mu); sh.done[g] = c + 1; }
                    sh.cv.notify_all();
                }
                for (int s = 0; s < 2; ++s) cudaEventDestroy(ev_in[s]);
                CUDA_OK(cudaStreamSynchronize(m->hstream[g]));
            });
            if (wrc) {
                std::lock_guard<std::mutex> l(sh.mu);
                sh.rc[g] = wrc; sh.err[g] = edsbwt_last_error(); sh.done[g] = C; sh.rc[0] = sh.rc[0] ? sh.rc[0] : wrc;
            }
            sh.cv.notify_all();
        });
        // ---- coordinator: when every device has searched chunk c, its gather is issued (asynchronously) and the devices move on
        bool failed = false;
        try {
            R->seg.assign((size_t)G * C, edsbwt_multi_result::Seg{0, 0, 0, nullptr});
            cudaEvent_t g0, g1;
            CUDA_OK(cudaSetDevice(R->root));
            CUDA_OK(cudaEventCreate(&g0)); CUDA_OK(cudaEventCreate(&g1));
            bool first = true;
            for (uint32_t c = 0; c < C && !failed; ++c) {
                {
                    std::unique_lock<std::mutex> l(sh.mu);
                    sh.cv.wait(l, [&] { for (int g = 0; g < G; ++g) if (sh.done[g] < c + 1) return false; return true; });
                    for (int g = 0; g < G; ++g) if (sh.rc[g]) failed = true;
                }
                if (failed) break;
                CUDA_OK(cudaSetDevice(R->root));
                for (int g = 0; g < G; ++g) {
                    auto& sg = R->seg[(size_t)g * C + c];
                    sg.pat_lo = clo(g, c); sg.pat_hi = clo(g, c + 1); sg.nocc = locate ? chunk_nocc[(size_t)g * C + c] : 0;
                    if (sg.nocc) CUDA_OK(cudaMalloc(&sg.d_occ, (size_t)sg.nocc * sizeof(edsbwt_occ)));
                }
                if (first) { CUDA_OK(cudaEventRecord(g0, m->cstream[0])); first = false; }
                const bool use_nccl = !m->comm.empty();
                if (use_nccl) NCCL_OK(nccl_api().GroupStart());
                for (int g = 0; g < G; ++g) {
                    const auto& sg = R->seg[(size_t)g * C + c];
                    SearchOut& so = slots[g * 2 + (c & 1)];
                    const size_t cb = (size_t)(sg.pat_hi - sg.pat_lo) * 4, ob = (size_t)sg.nocc * sizeof(edsbwt_occ);
                    if (g == 0 || !use_nccl) {
                        // the root's own shard (and every shard when NCCL is not in use) moves with copy engines: device-to-device on
                        // the root, peer copies over NVLink otherwise
                        if (cb) CUDA_OK(cudaMemcpyPeerAsync(R->d_counts + sg.pat_lo, R->root, so.counts.p, m->dev[g], cb, m->cstream[g]));
                        if (ob) CUDA_OK(cudaMemcpyPeerAsync(sg.d_occ, R->root, so.occ.p, m->dev[g], ob, m->cstream[g]));
                    } else {
                        if (cb) { NCCL_OK(nccl_api().Send(so.counts.p, cb, ncclUint8, 0, m->comm[g], m->cstream[g]));
                                  NCCL_OK(nccl_api().Recv(R->d_counts + sg.pat_lo, cb, ncclUint8, g, m->comm[0], m->cstream[0])); }
                        if (ob) { NCCL_OK(nccl_api().Send(so.occ.p, ob, ncclUint8, 0, m->comm[g], m->cstream[g]));
                                  NCCL_OK(nccl_api().Recv(sg.d_occ, ob, ncclUint8, g, m->comm[0], m->cstream[0])); }
                    }
                    if (g) R->gather_bytes += cb + ob;
                    R->nocc += sg.nocc;
                }
                if (use_nccl) NCCL_OK(nccl_api().GroupEnd());
                for (int g = 0; g < G; ++g) { CUDA_OK(cudaSetDevice(m->dev[g])); CUDA_OK(cudaEventRecord(ev_sent[g * 2 + (c & 1)], m->cstream[g])); }
                { std::lock_guard<std::mutex> l(sh.mu); for (int g = 0; g < G; ++g) sent[g] = c + 1; }
                sh.cv.notify_all();
            }
            for (int g = 0; g < G; ++g) { CUDA_OK(cudaSetDevice(m->dev[g])); CUDA_OK(cudaStreamSynchronize(m->cstream[g])); }
            CUDA_OK(cudaSetDevice(R->root));
            if (!first) { CUDA_OK(cudaEventRecord(g1, m->cstream[0])); CUDA_OK(cudaEventSynchronize(g1)); float ms = 0; CUDA_OK(cudaEventElapsedTime(&ms, g0, g1)); R->gather_ms = ms; }
            cudaEventDestroy(g0); cudaEventDestroy(g1);
        } catch (...) {
            { std::lock_guard<std::mutex> l(sh.mu); if (!sh.rc[0]) sh.rc[0] = EDSBWT_E_CUDA; for (int g = 0; g < G; ++g) sent[g] = C; }
            sh.cv.notify_all();
            for (auto& t : workers) t.join();
            throw;
        }
        { std::lock_guard<std::mutex> l(sh.mu); for (int g = 0; g < G; ++g) sent[g] = C; }
        sh.cv.notify_all();
        for (auto& t : workers) t.join();
        for (int g = 0; g < G; ++g) if (sh.rc[g]) { set_error("device %d: %s", m->dev[g], sh.err[g].c_str()); throw Failure{sh.rc[g]}; }
    });
    for (int g = 0; g < G; ++g) { cudaSetDevice(m->dev[g]); cudaStreamSynchronize(m->cstream[g]); for (int s = 0; s < 2; ++s) if (ev_sent[g * 2 + s]) cudaEventDestroy(ev_sent[g * 2 + s]); }
    for (int g = 0; g < G; ++g) { cudaSetDevice(m->dev[g]); slots[g * 2].counts.release(); slots[g * 2].occ.release(); slots[g * 2 + 1].counts.release(); slots[g * 2 + 1].occ.release(); }
    if (rc) { delete R; return rc; }
    R->search_ms = *std::max_element(search_ms.begin(), search_ms.end());
    R->wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_wall0).count();
    *res = R;
    return EDSBWT_OK;
}

int edsbwt_multi_result_info(const edsbwt_multi_result* r, uint32_t* npat, uint64_t* nocc, uint32_t* nseg, double* search_ms, double* gather_ms,
                             double* wall_ms, uint64_t* gather_bytes) {
    if (!r) return EDSBWT_E_ARG;
    if (npat) *npat = r->npat;
    if (nocc) *nocc = r->nocc;
    if (nseg) *nseg = (uint32_t)r->seg.size();
    if (search_ms) *search_ms = r->search_ms;
    if (gather_ms) *gather_ms = r->gather_ms;
    if (wall_ms) *wall_ms = r->wall_ms;
    if (gather_bytes) *gather_bytes = r->gather_bytes;
    return EDSBWT_OK;
}

int edsbwt_multi_result_device_ptrs(const edsbwt_multi_result* r, const uint32_t** d_counts, const edsbwt_occ** seg_occ, uint64_t* seg_nocc,
                                    uint32_t* seg_pat_lo, uint32_t seg_cap) {
    if (!r) return EDSBWT_E_ARG;
    if (d_counts) *d_counts = r->d_counts;
    for (uint32_t i = 0; i < r->seg.size() && i < seg_cap; ++i) {
        if (seg_occ) seg_occ[i] = r->seg[i].d_occ;
        if (seg_nocc) seg_nocc[i] = r->seg[i].nocc;
        if (seg_pat_lo) seg_pat_lo[i] = r->seg[i].pat_lo;
    }
    return EDSBWT_OK;
}

int edsbwt_multi_result_fetch(const edsbwt_multi_result* r, uint32_t* counts, edsbwt_occ* occ) {
    if (!r) return EDSBWT_E_ARG;
    return guarded_multi([&] {
        CUDA_OK(cudaSetDevice(r->root));
        if (counts && r->npat) CUDA_OK(cudaMemcpy(counts, r->d_counts, (size_t)r->npat * 4, cudaMemcpyDeviceToHost));
        if (occ) { uint64_t o = 0; for (const auto& s : r->seg) { if (s.nocc) CUDA_OK(cudaMemcpy(occ + o, s.d_occ, (size_t)s.nocc * sizeof(edsbwt_occ), cudaMemcpyDeviceToHost)); o += s.nocc; } }
    });
}

int edsbwt_multi_result_write_csv(edsbwt_multi* m, const edsbwt_multi_result* r, const char* path, int move, int append, uint64_t* bytes) {
    if (!m || !r || !path) return EDSBWT_E_ARG;
    if (bytes) *bytes = 0;
    if (!append) { int rc = edsbwt_write_csv(path, nullptr, 0, 0, move, 0); if (rc) return rc; }
    FILE* f = fopen(path, "ab");
    if (!f) { set_error("ERROR opening file %s to write output", path); return EDSBWT_E_IO; }
    int rc = guarded_multi([&] {
        Index& ix = m->idx[0]->ix;
        CUDA_OK(cudaSetDevice(ix.device));
        std::lock_guard<std::mutex> lock(ix.mu);
        uint64_t w = 0;
        for (const auto& s : r->seg) if (s.nocc) w += write_csv_device(ix, s.d_occ, s.nocc, 1, f, ix.stream, nullptr);   // tuples carry batch-global pattern numbers
        if (bytes) *bytes = w;
    });
    if (fclose(f) != 0 && !rc) { set_error("error closing %s", path); rc = EDSBWT_E_IO; }
    return rc;
}

void edsbwt_multi_result_free(edsbwt_multi_result* r) { delete r; }

// ---- one process per GPU: the gather over peer memory -------------------------------------------------------------------------
// The root process allocates an arena in its HBM and exports it (CUDA IPC); every other process maps it and, as soon as a chunk
// of its shard has been searched, pushes the chunk's counts and tuples straight to their place in the arena with asynchronous
// peer copies over NVLink (copy engines, not SMs), on a stream of its own: the transfer of chunk c runs under the search of
// chunk c + 1, nobody exchanges sizes first and the root posts nothing.  Rank r owns a fixed region of the arena and fills it
// front to back, so every destination is known locally.
int edsbwt_device_alloc(int device, uint64_t bytes, void** d_ptr) {
    if (!d_ptr) return EDSBWT_E_ARG;
    *d_ptr = nullptr;
    return guarded_multi([&] { CUDA_OK(cudaSetDevice(device)); CUDA_OK(cudaMalloc(d_ptr, bytes ? bytes : 1)); });
}
void edsbwt_device_free(int device, void* d_ptr) { if (d_ptr) { cudaSetDevice(device); cudaFree(d_ptr); } }
int edsbwt_ipc_export(int device, const void* d_ptr, uint8_t handle[64]) {
    if (!d_ptr || !handle) return EDSBWT_E_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    return guarded_multi([&] { CUDA_OK(cudaSetDevice(device)); cudaIpcMemHandle_t h; CUDA_OK(cudaIpcGetMemHandle(&h, const_cast<void*>(d_ptr))); memcpy(handle, &h, 64); });
}
int edsbwt_ipc_open(int device, const uint8_t handle[64], void** d_ptr) {
    if (!d_ptr || !handle) return EDSBWT_E_ARG;
    *d_ptr = nullptr;
    return guarded_multi([&] { CUDA_OK(cudaSetDevice(device)); cudaIpcMemHandle_t h; memcpy(&h, handle, 64);
                               CUDA_OK(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess)); });
}
int edsbwt_ipc_close(int device, void* d_ptr) {
    if (!d_ptr) return EDSBWT_OK;
    return guarded_multi([&] { CUDA_OK(cudaSetDevice(device)); CUDA_OK(cudaIpcCloseMemHandle(d_ptr)); });
}
int edsbwt_result_push(edsbwt_index* idx, const edsbwt_result* res, uint32_t* d_counts_dst, edsbwt_occ* d_occ_dst) {
    if (!idx || !res) return EDSBWT_E_ARG;
    return guarded_multi([&] {
        Index& ix = idx->ix;
        CUDA_OK(cudaSetDevice(ix.device));
        if (!ix.push) CUDA_OK(cudaStreamCreateWithFlags(&ix.push, cudaStreamNonBlocking));
        const SearchOut& o = *res->outp;           // the search has completed (run_search returns with its stream idle)
        if (d_counts_dst && o.npat) CUDA_OK(cudaMemcpyAsync(d_counts_dst, o.counts.p, (size_t)o.npat * 4, cudaMemcpyDefault, ix.push));
        if (d_occ_dst && o.nocc) CUDA_OK(cudaMemcpyAsync(d_occ_dst, o.occ.p, (size_t)o.nocc * sizeof(edsbwt_occ), cudaMemcpyDefault, ix.push));
    });
}
int edsbwt_push_wait(edsbwt_index* idx) {
    if (!idx) return EDSBWT_E_ARG;
    return guarded_multi([&] { if (idx->ix.push) { CUDA_OK(cudaSetDevice(idx->ix.device)); CUDA_OK(cudaStreamSynchronize(idx->ix.push)); } });
}

}  // extern "C"
