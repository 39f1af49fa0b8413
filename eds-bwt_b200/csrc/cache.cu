// Persisted device tables ("our own documented on-disk format next to _MLF.aux").
//
//   <base>_MLF.edsbwt   the move table of MOVE_EDSBWTSearch -- what the reference's build_MLF writes as <base>_MLF.aux
//                       (build_MLF.cpp:153-161) and MOVE_EDSBWT::retrieve_MLF reads back (MOVE_EDSBWTSearch.cpp:186-192).  Move-r's
//                       own serialisation is private to that library, so this is our format, not a re-implementation of theirs.
//   <base>.edsbwt       the whole HBM image of a handle (Occ blocks, link records, SA table, seed levels, and the move table
//                       when built): opt-in (edsbwt_open_opts.cache), because reading 20 GB back is only worth it when the file
//                       sits in the page cache.
//
// File layout (little-endian, both files):
//   [0, 1024)        CacheHeader: magic "EDSBWTC1", kind, fingerprint of the source files, every scalar of the index
//   [1024, ...)      nsec x CacheSec {name[24], offset, bytes}
//   4096-aligned     the sections: raw device arrays, byte for byte
// A file is used only if its fingerprint equals the one computed from the index files at open: sizes, modification times and
// 16 sampled 64-KB windows of <base>_info.aux, <base>.bitvector, <base>_bwt_<p>.aux and <base>_bv_<p>.aux (FNV-1a).  Anything
// else -- truncated file, other version, other shape -- makes the open fall back to building, silently and correctly.
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <thread>

#include "internal.cuh"

namespace edsbwt {

namespace {

struct CacheHeader {
    char     magic[8];
    uint32_t version, kind, nsec, pad0;
    uint64_t fingerprint;
    uint32_t n, n_text, n_seg, sigma;
    uint8_t  alphabet[8];
    uint32_t pile_start[9], fsi, nletters, Cl[5];
    uint32_t table_occ[64];
    uint32_t seed_k, locate_mode;
    uint64_t seed_intervals, seed_niv[17];
    uint32_t has_move, move_r, move_r_runs, move_a, move_rounds, selbase[5];
    uint8_t  reserved[1024 - 8 - 16 - 8 - 16 - 8 - 64 - 256 - 8 - 8 - 136 - 40];
};
static_assert(sizeof(CacheHeader) == 1024, "cache header layout");
struct CacheSec { char name[24]; uint64_t off, bytes; };
static_assert(sizeof(CacheSec) == 40, "cache section layout");
constexpr uint32_t CACHE_VERSION = 1, KIND_IMAGE = 1, KIND_MOVE = 2;
constexpr size_t CHUNK = 8u << 20;

struct Job { uint64_t off, bytes; uint8_t* dev; };

uint64_t fnv(uint64_t h, const void* p, size_t n) {
    const uint8_t* b = (const uint8_t*)p;
    for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 0x100000001b3ull; }
    return h;
}

// device <-> file, `dir` = 0: file -> device, 1: device -> file.  Chunks of 8 MB go through pinned bounce buffers on several
// threads (page-cache reads are memcpy-bound per thread; PCIe wants ~40 GB/s)
void bulk_transfer(int fd, const std::vector<Job>& jobs, int device, int dir) {
    struct Piece { uint64_t off; size_t len; uint8_t* dev; };
    std::vector<Piece> pieces;
    for (const Job& j : jobs) for (uint64_t o = 0; o < j.bytes; o += CHUNK) pieces.push_back({j.off + o, (size_t)std::min<uint64_t>(CHUNK, j.bytes - o), j.dev + o});
    if (pieces.empty()) return;
    std::atomic<size_t> next{0};
    std::atomic<int> failed{0};
    std::string err; std::mutex emu;
    size_t nt_ = std::min<size_t>(8, pieces.size());
    nt_ = std::min<size_t>(nt_, std::max(1u, std::thread::hardware_concurrency()));
    const int nt = (int)std::max<size_t>(1, nt_);
    auto worker = [&] {
        try {
            CUDA_OK(cudaSetDevice(device));
            cudaStream_t s; CUDA_OK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
            uint8_t* buf[2] = {nullptr, nullptr};
            CUDA_OK(cudaHostAlloc((void**)&buf[0], CHUNK, cudaHostAllocDefault));
            CUDA_OK(cudaHostAlloc((void**)&buf[1], CHUNK, cudaHostAllocDefault));
            cudaEvent_t ev[2]; CUDA_OK(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming)); CUDA_OK(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
            bool used[2] = {false, false};
            Piece pending[2];
            auto io = [&](const Piece& p, uint8_t* b) {
                size_t done = 0;
                while (done < p.len) {
                    const ssize_t k = dir == 0 ? pread(fd, b + done, p.len - done, (off_t)(p.off + done)) : pwrite(fd, b + done, p.len - done, (off_t)(p.off + done));
                    if (k <= 0) EDS_FAIL(EDSBWT_E_IO, "cache file: short %s at offset %llu", dir == 0 ? "read" : "write", (unsigned long long)(p.off + done));
                    done += (size_t)k;
                }
            };
            for (int turn = 0;; turn ^= 1) {
                const size_t i = next.fetch_add(1);
                if (used[turn]) {                                 // the buffer's previous copy must have finished
                    CUDA_OK(cudaEventSynchronize(ev[turn]));
                    if (dir == 1) io(pending[turn], buf[turn]);
                    used[turn] = false;
                }
                if (i >= pieces.size() || failed.load()) {
                    const int other = turn ^ 1;
                    if (used[other]) { CUDA_OK(cudaEventSynchronize(ev[other])); if (dir == 1) io(pending[other], buf[other]); used[other] = false; }
                    break;
                }
                const Piece& p = pieces[i];
                if (dir == 0) { io(p, buf[turn]); CUDA_OK(cudaMemcpyAsync(p.dev, buf[turn], p.len, cudaMemcpyHostToDevice, s)); }
                else CUDA_OK(cudaMemcpyAsync(buf[turn], p.dev, p.len, cudaMemcpyDeviceToHost, s));
                CUDA_OK(cudaEventRecord(ev[turn], s));
                pending[turn] = p; used[turn] = true;
            }
            CUDA_OK(cudaStreamSynchronize(s));
            cudaEventDestroy(ev[0]); cudaEventDestroy(ev[1]);
            cudaFreeHost(buf[0]); cudaFreeHost(buf[1]); cudaStreamDestroy(s);
        } catch (const Failure& f) {
            std::lock_guard<std::mutex> l(emu);
            if (!failed.exchange(f.code ? f.code : EDSBWT_E_IO)) err = edsbwt_last_error();
        }
    };
    std::vector<std::thread> th;
    for (int t = 0; t < nt; ++t) th.emplace_back(worker);
    for (auto& t : th) t.join();
    if (failed.load()) { set_error("%s", err.c_str()); throw Failure{failed.load()}; }
}

struct Writer {
    int fd = -1; std::string tmp, final_path; std::vector<CacheSec> secs; std::vector<Job> jobs; uint64_t cursor = 0;
    void open(const std::string& path, uint32_t max_sec) {
        final_path = path; tmp = path + ".tmp." + std::to_string((long long)getpid());
        fd = ::open(tmp.c_str(), O_CREAT | O_TRUNC | O_WRONLY, 0644);
        if (fd < 0) EDS_FAIL(EDSBWT_E_IO, "cannot create %s: %s", tmp.c_str(), strerror(errno));
        cursor = (sizeof(CacheHeader) + (uint64_t)max_sec * sizeof(CacheSec) + 4095) & ~4095ull;
    }
    void add(const char* name, const void* dev, uint64_t bytes) {
        CacheSec s{}; snprintf(s.name, sizeof s.name, "%s", name); s.off = cursor; s.bytes = bytes;
        secs.push_back(s);
        if (bytes) jobs.push_back({cursor, bytes, (uint8_t*)dev});
        cursor = (cursor + bytes + 4095) & ~4095ull;
    }
    void finish(CacheHeader& h, int device) {
        h.nsec = (uint32_t)secs.size();
        bulk_transfer(fd, jobs, device, 1);
        if (pwrite(fd, secs.data(), secs.size() * sizeof(CacheSec), sizeof(CacheHeader)) != (ssize_t)(secs.size() * sizeof(CacheSec)) ||
            pwrite(fd, &h, sizeof h, 0) != (ssize_t)sizeof h) EDS_FAIL(EDSBWT_E_IO, "cannot write %s", tmp.c_str());
        if (::close(fd) != 0) { fd = -1; EDS_FAIL(EDSBWT_E_IO, "cannot close %s", tmp.c_str()); }
        fd = -1;
        if (rename(tmp.c_str(), final_path.c_str()) != 0) EDS_FAIL(EDSBWT_E_IO, "cannot rename %s", tmp.c_str());
    }
    ~Writer() { if (fd >= 0) { ::close(fd); unlink(tmp.c_str()); } }
};

struct Reader {
    int fd = -1; CacheHeader h{}; std::vector<CacheSec> secs; uint64_t size = 0;
    bool open(const std::string& path, uint32_t kind, uint64_t fingerprint) {
        fd = ::open(path.c_str(), O_RDONLY);
        if (fd < 0) return false;
        struct stat sb;
        if (fstat(fd, &sb) != 0) return false;
        size = (uint64_t)sb.st_size;
        if (size < sizeof h || pread(fd, &h, sizeof h, 0) != (ssize_t)sizeof h) return false;
        if (memcmp(h.magic, "EDSBWTC1", 8) != 0 || h.version != CACHE_VERSION || h.kind != kind || h.fingerprint != fingerprint || h.nsec > 4096) return false;
        secs.resize(h.nsec);
        if (h.nsec && pread(fd, secs.data(), h.nsec * sizeof(CacheSec), sizeof h) != (ssize_t)(h.nsec * sizeof(CacheSec))) return false;
        for (const CacheSec& s : secs) if (s.off + s.bytes > size) return false;          // truncated
        return true;
    }
    const CacheSec* find(const char* name) const { for (const CacheSec& s : secs) if (strncmp(s.name, name, sizeof s.name) == 0) return &s; return nullptr; }
    ~Reader() { if (fd >= 0) ::close(fd); }
};

void fill_header_common(CacheHeader& h, const Index& ix, uint32_t kind, uint64_t fp) {
    memset(&h, 0, sizeof h);
    memcpy(h.magic, "EDSBWTC1", 8); h.version = CACHE_VERSION; h.kind = kind; h.fingerprint = fp;
    const HostIndex& H = ix.host;
    h.n = H.n; h.n_text = H.n_text; h.n_seg = H.n_seg; h.sigma = H.sigma;
    memcpy(h.alphabet, H.alphabet, 8); memcpy(h.pile_start, H.pile_start, sizeof h.pile_start); memcpy(h.table_occ, H.table_occ, sizeof h.table_occ);
    h.fsi = ix.view.fsi; h.nletters = ix.view.nletters; memcpy(h.Cl, ix.view.Cl, sizeof h.Cl);
    h.seed_k = ix.seed_k; h.locate_mode = (uint32_t)ix.locate_mode; h.seed_intervals = ix.seed_intervals;
}
void add_move(Writer& w, CacheHeader& h, const MoveTable& mt, uint32_t n, uint32_t n_text) {
    h.has_move = 1; h.move_r = mt.r; h.move_r_runs = mt.r_runs; h.move_a = mt.a; h.move_rounds = mt.rounds;
    memcpy(h.selbase, mt.view.selbase, sizeof h.selbase);
    w.add("move.T", mt.T.p, ((uint64_t)mt.r + 2) * sizeof(uint4));
    w.add("move.occL", mt.occL.p, ((uint64_t)mt.r / OCC_ROWS + 1) * 64);
    w.add("move.sel", mt.sel.p, ((uint64_t)mt.view.selbase[4] + 1) * 4);
    w.add("move.coarse", mt.coarse.p, ((uint64_t)n / OCC_ROWS + 3) * 4);
    w.add("move.row_run0", mt.row_run0.p, ((uint64_t)n_text + 1) * 4);
}
// -> jobs that fill a fresh MoveTable from the file; false when a section is missing or has the wrong size
bool plan_move(const Reader& r, MoveTable& mt, uint32_t n, uint32_t n_text, std::vector<Job>& jobs) {
    const CacheHeader& h = r.h;
    if (!h.has_move || h.move_r == 0 || h.move_r > n) return false;
    mt.r = h.move_r; mt.r_runs = h.move_r_runs; mt.a = h.move_a; mt.rounds = h.move_rounds;
    struct Want { const char* name; uint64_t bytes; void** dst; } wants[5];
    mt.T.exact((size_t)mt.r + 2); mt.occL.exact(((size_t)mt.r / OCC_ROWS + 1) * 4); mt.sel.exact((size_t)h.selbase[4] + 1);
    mt.coarse.exact((size_t)n / OCC_ROWS + 3); mt.row_run0.exact((size_t)n_text + 1);
    wants[0] = {"move.T", ((uint64_t)mt.r + 2) * sizeof(uint4), (void**)&mt.T.p};
    wants[1] = {"move.occL", ((uint64_t)mt.r / OCC_ROWS + 1) * 64, (void**)&mt.occL.p};
    wants[2] = {"move.sel", ((uint64_t)h.selbase[4] + 1) * 4, (void**)&mt.sel.p};
    wants[3] = {"move.coarse", ((uint64_t)n / OCC_ROWS + 3) * 4, (void**)&mt.coarse.p};
    wants[4] = {"move.row_run0", ((uint64_t)n_text + 1) * 4, (void**)&mt.row_run0.p};
    for (const Want& w : wants) {
        const CacheSec* s = r.find(w.name);
        if (!s || s->bytes != w.bytes) return false;
        jobs.push_back({s->off, s->bytes, (uint8_t*)*w.dst});
    }
    MoveView& mv = mt.view;
    mv.T = mt.T.p; mv.r = mt.r; mv.occL = mt.occL.p; mv.sel = mt.sel.p; mv.coarse = mt.coarse.p; mv.row_run0 = mt.row_run0.p;
    memcpy(mv.selbase, h.selbase, sizeof mv.selbase);
    return true;
}

}  // namespace

uint64_t source_fingerprint(const std::string& base) {
    uint64_t h = 0xcbf29ce484222325ull;
    std::vector<std::string> files = {base + "_info.aux", base + ".bitvector"};
    for (int p = 0; p < 8; ++p) { files.push_back(base + "_bwt_" + std::to_string(p) + ".aux"); files.push_back(base + "_bv_" + std::to_string(p) + ".aux"); }
    std::vector<uint8_t> buf(65536);
    for (const std::string& f : files) {
        struct stat sb;
        if (stat(f.c_str(), &sb) != 0) { h = fnv(h, "absent", 6); continue; }
        const uint64_t meta[3] = {(uint64_t)sb.st_size, (uint64_t)sb.st_mtim.tv_sec, (uint64_t)sb.st_mtim.tv_nsec};
        h = fnv(h, meta, sizeof meta);
        const int fd = ::open(f.c_str(), O_RDONLY);
        if (fd < 0) continue;
        const uint64_t size = (uint64_t)sb.st_size, nwin = size <= 16 * 65536 ? (size + 65535) / 65536 : 16;
        for (uint64_t w = 0; w < nwin; ++w) {
            const uint64_t off = nwin == 16 && size > 16 * 65536 ? (w == 15 ? size - 65536 : (size / 16) * w) : w * 65536;
            const ssize_t k = pread(fd, buf.data(), buf.size(), (off_t)off);
            if (k > 0) h = fnv(h, buf.data(), (size_t)k);
        }
        ::close(fd);
    }
    return h;
}

// ---- move table: <base>_MLF.edsbwt ------------------------------------------------------------------------------
void move_cache_save(Index& ix, const std::string& path) {
    if (!ix.move) EDS_FAIL(EDSBWT_E_ARG, "no move table to save");
    CUDA_OK(cudaSetDevice(ix.device));
    CUDA_OK(cudaStreamSynchronize(ix.stream));
    Writer w; w.open(path, 8);
    CacheHeader h; fill_header_common(h, ix, KIND_MOVE, source_fingerprint(ix.base));
    add_move(w, h, *ix.move, ix.view.n, ix.view.n_text);
    w.finish(h, ix.device);
}
bool move_cache_load(Index& ix, const std::string& path) {
    Reader r;
    if (!r.open(path, KIND_MOVE, source_fingerprint(ix.base))) return false;
    if (r.h.n != ix.view.n || r.h.n_text != ix.view.n_text) return false;
    if (ix.move_a_explicit && r.h.move_a != ix.move_a) return false;          // the caller asked for another balancing parameter
    CUDA_OK(cudaSetDevice(ix.device));
    std::unique_ptr<MoveTable> mt(new MoveTable);
    std::vector<Job> jobs;
    if (!plan_move(r, *mt, ix.view.n, ix.view.n_text, jobs)) return false;
    bulk_transfer(r.fd, jobs, ix.device, 0);
    ix.move = mt.release();
    ix.move_from_file = true;
    return true;
}

// ---- whole image: <base>.edsbwt ---------------------------------------------------------------------------------
void image_cache_save(Index& ix) {
    if (ix.wide) return;          // the image format describes the 2-bit layout only: wide-alphabet indexes are rebuilt at every open
    CUDA_OK(cudaSetDevice(ix.device));
    CUDA_OK(cudaStreamSynchronize(ix.stream));
    Writer w; w.open(ix.base + ".edsbwt", 16 + 4 * 17);
    CacheHeader h; fill_header_common(h, ix, KIND_IMAGE, source_fingerprint(ix.base));
    const IndexView& v = ix.view;
    w.add("occ", ix.occ.p, ((uint64_t)v.n / OCC_ROWS + 1) * 64);
    w.add("eof_flat", ix.eof_flat.p, (uint64_t)v.n_text * 4);
    w.add("str_seg", ix.str_seg.p, (uint64_t)v.n_text * 4);
    w.add("seg_first", ix.seg_first.p, ((uint64_t)v.n_seg + 2) * 4);
    w.add("seg_low", ix.seg_low.p, ((uint64_t)v.n_seg + 1) * 4);
    w.add("link_rec", ix.link_rec.p, (uint64_t)v.n_text * 8);
    w.add("str_info", ix.str_info.p, (uint64_t)v.n_text * 8);
    if (ix.locate_mode == EDSBWT_LOCATE_SA) w.add("sa4", ix.sa4.p, (uint64_t)v.n * 16);
    for (uint32_t L = 1; L <= ix.seed_k; ++L) {
        const SeedLevel* lv = ix.seed[L];
        const uint64_t nkeys = 1ull << (2 * L);
        char nm[24];
        h.seed_niv[L] = lv->n_iv;
        snprintf(nm, sizeof nm, "seed.%u.off", L); w.add(nm, lv->off.p, (nkeys + 1) * 4);
        snprintf(nm, sizeof nm, "seed.%u.iv", L); w.add(nm, lv->iv.p, (uint64_t)lv->n_iv * 8);
        if (lv->cnt.p) { snprintf(nm, sizeof nm, "seed.%u.cnt", L); w.add(nm, lv->cnt.p, nkeys * 4); }
        if (lv->first.p) { snprintf(nm, sizeof nm, "seed.%u.first", L); w.add(nm, lv->first.p, nkeys * 16); }
    }
    if (ix.move) add_move(w, h, *ix.move, v.n, v.n_text);
    w.finish(h, ix.device);
}

// fills ix (host scalars, device arrays, views, seed levels, move table) from <base>.edsbwt; false: not usable, build instead.
// The runtime part of the handle (stream, workspace, lut) is set up by the caller afterwards (init_index_runtime).
bool image_cache_load(Index& ix, const edsbwt_open_opts& o) {
    Reader r;
    if (!r.open(ix.base + ".edsbwt", KIND_IMAGE, source_fingerprint(ix.base))) return false;
    const CacheHeader& h = r.h;
    if (h.sigma < 2 || h.sigma > 5 || h.seed_k < 1 || h.seed_k > 16) return false;
    if (o.seed_k >= 0 && (uint32_t)std::max(1, std::min(o.seed_k, 14)) != h.seed_k) return false;
    const int want_mode = o.locate_mode == EDSBWT_LOCATE_WALK ? EDSBWT_LOCATE_WALK : EDSBWT_LOCATE_SA;
    if ((int)h.locate_mode != want_mode) return false;
    CUDA_OK(cudaSetDevice(ix.device));
    HostIndex& H = ix.host;
    H.n = h.n; H.n_text = h.n_text; H.n_seg = h.n_seg; H.sigma = h.sigma;
    memcpy(H.alphabet, h.alphabet, 8); memcpy(H.pile_start, h.pile_start, sizeof h.pile_start); memcpy(H.table_occ, h.table_occ, sizeof h.table_occ);
    IndexView& v = ix.view;
    v.n = h.n; v.n_text = h.n_text; v.n_seg = h.n_seg; v.fsi = h.fsi; v.nletters = h.nletters; memcpy(v.Cl, h.Cl, sizeof h.Cl);
    std::vector<Job> jobs;
    bool ok = true;
    auto want = [&](const char* name, uint64_t bytes, auto& buf, size_t elems) {
        const CacheSec* s = r.find(name);
        if (!s || s->bytes != bytes) { ok = false; return; }
        buf.exact(elems ? elems : 1);
        if (bytes) jobs.push_back({s->off, bytes, (uint8_t*)buf.p});
    };
    want("occ", ((uint64_t)v.n / OCC_ROWS + 1) * 64, ix.occ, ((size_t)v.n / OCC_ROWS + 1) * 4);
    want("eof_flat", (uint64_t)v.n_text * 4, ix.eof_flat, (size_t)v.n_text + 1);
    want("str_seg", (uint64_t)v.n_text * 4, ix.str_seg, (size_t)v.n_text + 1);
    want("seg_first", ((uint64_t)v.n_seg + 2) * 4, ix.seg_first, (size_t)v.n_seg + 3);
    want("seg_low", ((uint64_t)v.n_seg + 1) * 4, ix.seg_low, (size_t)v.n_seg + 2);
    want("link_rec", (uint64_t)v.n_text * 8, ix.link_rec, (size_t)v.n_text + 1);
    want("str_info", (uint64_t)v.n_text * 8, ix.str_info, (size_t)v.n_text + 1);
    if (want_mode == EDSBWT_LOCATE_SA) want("sa4", (uint64_t)v.n * 16, ix.sa4, (size_t)v.n + 1);
    if (!ok) return false;
    ix.seed.assign(h.seed_k + 1, nullptr);
    std::vector<const uint2*> ivp(17, nullptr);
    for (uint32_t L = 1; L <= h.seed_k && ok; ++L) {
        SeedLevel* lv = new SeedLevel; ix.seed[L] = lv;
        const uint64_t nkeys = 1ull << (2 * L);
        char nm[24];
        lv->n_iv = h.seed_niv[L];
        snprintf(nm, sizeof nm, "seed.%u.off", L); want(nm, (nkeys + 1) * 4, lv->off, nkeys + 1);
        snprintf(nm, sizeof nm, "seed.%u.iv", L); want(nm, (uint64_t)lv->n_iv * 8, lv->iv, lv->n_iv + 1);
        snprintf(nm, sizeof nm, "seed.%u.cnt", L); if (r.find(nm)) want(nm, nkeys * 4, lv->cnt, nkeys);
        snprintf(nm, sizeof nm, "seed.%u.first", L); if (r.find(nm)) want(nm, nkeys * 16, lv->first, nkeys);
        ivp[L] = lv->iv.p;
    }
    if (!ok || !ix.seed[h.seed_k]->first.p) { for (SeedLevel* s : ix.seed) delete s; ix.seed.clear(); return false; }
    std::unique_ptr<MoveTable> mt;
    if (h.has_move && (!ix.move_a_explicit || h.move_a == ix.move_a)) { mt.reset(new MoveTable); if (!plan_move(r, *mt, v.n, v.n_text, jobs)) mt.reset(); }
    bulk_transfer(r.fd, jobs, ix.device, 0);
    v.occ = ix.occ.p; v.eof_flat = ix.eof_flat.p; v.str_seg = ix.str_seg.p; v.seg_first = ix.seg_first.p; v.seg_low = ix.seg_low.p;
    v.link_rec = ix.link_rec.p; v.str_info = ix.str_info.p; v.sa4 = want_mode == EDSBWT_LOCATE_SA ? ix.sa4.p : nullptr;
    ix.seed_k = h.seed_k; ix.locate_mode = want_mode; ix.seed_intervals = h.seed_intervals;
    ix.seed_ivp.ensure(17);
    CUDA_OK(cudaMemcpy(ix.seed_ivp.p, ivp.data(), 17 * sizeof(const uint2*), cudaMemcpyHostToDevice));
    if (mt) { ix.move = mt.release(); ix.move_from_file = true; }
    return true;
}

}  // namespace edsbwt
