// Host-side file formats of the EDS-BWT index: info.aux, pile files, sdsl bit_vector and
// rrr_vector<63> serialisations, the pattern file and the output csv.  Pure C++ (no CUDA).
#include <cerrno>
#include <cstdarg>
#include <fstream>
#include <new>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "internal.cuh"

namespace edsbwt {

static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap);
}
const char* last_error_cstr() { return g_err; }

std::vector<uint8_t> read_file(const std::string& path) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) EDS_FAIL(EDSBWT_E_IO, "Error opening %s: %s", path.c_str(), strerror(errno));
    struct stat sb;
    if (fstat(fileno(f), &sb) != 0 || !S_ISREG(sb.st_mode)) { fclose(f); EDS_FAIL(EDSBWT_E_IO, "Error opening %s: not a regular file", path.c_str()); }
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    if (sz < 0) { fclose(f); EDS_FAIL(EDSBWT_E_IO, "Error reading %s: %s", path.c_str(), strerror(errno)); }
    std::vector<uint8_t> b((size_t)sz);
    if (sz && fread(b.data(), 1, (size_t)sz, f) != (size_t)sz) { fclose(f); EDS_FAIL(EDSBWT_E_IO, "short read on %s", path.c_str()); }
    fclose(f);
    return b;
}
void write_file(const std::string& path, const void* data, size_t len) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) EDS_FAIL(EDSBWT_E_IO, "Error opening %s for writing: %s", path.c_str(), strerror(errno));
    if (len && fwrite(data, 1, len, f) != len) { fclose(f); EDS_FAIL(EDSBWT_E_IO, "short write on %s", path.c_str()); }
    if (fclose(f) != 0) EDS_FAIL(EDSBWT_E_IO, "Error writing %s: %s", path.c_str(), strerror(errno));
}

// ---------------------------------------------------------------------------------------------------
// sdsl int_vector helpers
// ---------------------------------------------------------------------------------------------------
namespace {
struct Cursor {
    const std::vector<uint8_t>& b; size_t off = 0;
    explicit Cursor(const std::vector<uint8_t>& v) : b(v) {}
    void need(size_t k) const { if (off + k > b.size()) EDS_FAIL(EDSBWT_E_FORMAT, "sdsl structure truncated (need %zu bytes at %zu of %zu)", k, off, b.size()); }
    uint64_t u64() { need(8); uint64_t v; memcpy(&v, &b[off], 8); off += 8; return v; }
    uint8_t u8() { need(1); return b[off++]; }
};
struct Packed {            // int_vector<0> (has_width) or int_vector<1>
    uint64_t nbits = 0; unsigned width = 1; std::vector<uint64_t> w;
    void load(Cursor& c, bool has_width) {
        nbits = c.u64(); width = has_width ? c.u8() : 1;
        size_t nw = (size_t)((nbits + 63) / 64);
        c.need(nw * 8);
        w.assign(nw + 1, 0);
        if (nw) memcpy(w.data(), &c.b[c.off], nw * 8);
        c.off += nw * 8;
    }
    uint64_t bits_at(uint64_t pos, unsigned len) const {
        if (len == 0) return 0;
        uint64_t lo = w[pos >> 6] >> (pos & 63);
        unsigned got = 64 - (unsigned)(pos & 63);
        if (got < len) lo |= w[(pos >> 6) + 1] << got;
        return len == 64 ? lo : (lo & ((uint64_t(1) << len) - 1));
    }
    uint64_t at(uint64_t j) const { return bits_at(j * width, width); }
    uint64_t count() const { return width ? nbits / width : 0; }
};
uint64_t g_binom[64][64];
bool g_binom_init = false;
void init_binom() {
    if (g_binom_init) return;
    for (int a = 0; a < 64; ++a) { g_binom[a][0] = 1; for (int b = 1; b < 64; ++b) g_binom[a][b] = a == 0 ? 0 : g_binom[a - 1][b - 1] + g_binom[a - 1][b]; }
    g_binom_init = true;
}
inline unsigned space_for_bt(unsigned k) { uint64_t c = g_binom[63][k]; return c == 1 ? 0 : 64 - (unsigned)__builtin_clzll(c); }
inline unsigned hi_bit(uint64_t x) { return x ? 63 - (unsigned)__builtin_clzll(x) : 0; }
struct BitWriter {
    std::vector<uint64_t> w; uint64_t nbits = 0;
    void put(uint64_t v, unsigned len) {
        if (!len) return;
        size_t wi = (size_t)(nbits >> 6); unsigned sh = (unsigned)(nbits & 63);
        if (w.size() < wi + 2) w.resize(wi + 2, 0);
        w[wi] |= v << sh;
        if (sh + len > 64) w[wi + 1] |= v >> (64 - sh);
        nbits += len;
    }
};
void append_u64(std::vector<uint8_t>& o, uint64_t v) { size_t s = o.size(); o.resize(s + 8); memcpy(&o[s], &v, 8); }
void append_words(std::vector<uint8_t>& o, const std::vector<uint64_t>& w, uint64_t nbits) {
    size_t nw = (size_t)((nbits + 63) / 64);
    for (size_t i = 0; i < nw; ++i) append_u64(o, i < w.size() ? w[i] : 0);
}
void append_int_vector(std::vector<uint8_t>& o, const std::vector<uint64_t>& vals, unsigned width) {
    BitWriter bw;
    for (uint64_t v : vals) bw.put(v, width);
    append_u64(o, (uint64_t)vals.size() * width);
    o.push_back((uint8_t)width);
    append_words(o, bw.w, (uint64_t)vals.size() * width);
}
}  // namespace

// sdsl bit_vector (App. A.3): u64 size in bits + ceil(size/64) u64 words, LSB first.
std::vector<uint8_t> bitvector_decode(const std::vector<uint8_t>& file) {
    Cursor c(file); Packed p; p.load(c, false);
    std::vector<uint8_t> out((size_t)p.nbits);
    for (uint64_t i = 0; i < p.nbits; ++i) out[(size_t)i] = (p.w[i >> 6] >> (i & 63)) & 1;
    return out;
}
std::vector<uint8_t> bitvector_encode(const uint8_t* bits01, size_t nbits) {
    std::vector<uint64_t> w((nbits + 63) / 64, 0);
    for (size_t i = 0; i < nbits; ++i) if (bits01[i]) w[i >> 6] |= uint64_t(1) << (i & 63);
    std::vector<uint8_t> o; append_u64(o, nbits); append_words(o, w, nbits);
    return o;
}

// sdsl rrr_vector<63,int_vector<>,32> (App. A.4).  Inversion semantics as probed against the reference's
// own prebuilt binary: an inverted superblock stores m_bt = 63 - popcount per block while the block code
// (m_btnr) stays the combinatorial rank of the ORIGINAL block.
std::vector<uint8_t> rrr_decode(const std::vector<uint8_t>& file) {
    init_binom();
    Cursor c(file);
    uint64_t m_size = c.u64();
    Packed bt, btnr, btnrp, rk, inv;
    bt.load(c, true); btnr.load(c, false); btnrp.load(c, true); rk.load(c, true); inv.load(c, false);
    std::vector<uint8_t> out((size_t)m_size, 0);
    uint64_t nreal = (m_size + 62) / 63, pos = 0, ones = 0;
    if (bt.count() < nreal) EDS_FAIL(EDSBWT_E_FORMAT, "rrr_vector: m_bt has %llu entries, need %llu", (unsigned long long)bt.count(), (unsigned long long)nreal);
    for (uint64_t j = 0; j < nreal; ++j) {
        if (j % 32 == 0) pos = btnrp.at(j / 32);
        unsigned k = (unsigned)bt.at(j);
        unsigned len = space_for_bt(k);
        uint64_t nr = btnr.bits_at(pos, len);
        pos += len;
        if (inv.bits_at(j / 32, 1)) k = 63 - k;
        for (unsigned p = 0; p < 63 && k > 0; ++p) {
            uint64_t cb = g_binom[62 - p][k];
            if (nr >= cb) { nr -= cb; --k; uint64_t i = 63 * j + p; if (i < m_size) { out[(size_t)i] = 1; ++ones; } }
        }
    }
    if (rk.count() && rk.at(rk.count() - 1) != ones)
        EDS_FAIL(EDSBWT_E_FORMAT, "rrr_vector: decoded %llu ones, m_rank says %llu", (unsigned long long)ones, (unsigned long long)rk.at(rk.count() - 1));
    return out;
}

std::vector<uint8_t> rrr_encode(const uint8_t* bits, size_t m_size) {
    init_binom();
    const size_t nblocks = m_size / 63 + 1;                 // incl. the dummy block when m_size % 63 == 0
    const size_t nsb = (nblocks + 31) / 32;
    const size_t nreal = (m_size + 62) / 63;
    std::vector<uint64_t> blk(nblocks, 0), bt(nblocks, 0);
    for (size_t i = 0; i < m_size; ++i) if (bits[i]) blk[i / 63] |= uint64_t(1) << (i % 63);
    uint64_t total_bits = 0, total_ones = 0;
    for (size_t j = 0; j < nblocks; ++j) { bt[j] = (uint64_t)__builtin_popcountll(blk[j]); total_bits += space_for_bt((unsigned)bt[j]); total_ones += bt[j]; }
    std::vector<uint64_t> btnrp(nsb, 0), rank(nsb + ((m_size % (32 * 63)) > 0 ? 1 : 0), 0), invert(nsb, 0);
    BitWriter code; uint64_t sum_rank = 0; bool inv = false;
    for (size_t i = 0; i < nreal; ++i) {
        const bool full = (i + 1) * 63 <= m_size;
        if (i % 32 == 0) {
            btnrp[i / 32] = code.nbits; rank[i / 32] = sum_rank; inv = false;
            if (full && i + 32 <= nblocks) {
                unsigned gt = 0;
                for (size_t j = i; j < i + 32; ++j) gt += bt[j] > 31;
                if (gt > 16) { invert[i / 32] = 1; for (size_t j = i; j < i + 32; ++j) bt[j] = 63 - bt[j]; inv = true; }
            }
        }
        unsigned x = (unsigned)bt[i], k = inv ? 63 - x : x;
        sum_rank += k;
        unsigned sp = space_for_bt(x);
        if (sp) {
            uint64_t nr = 0, b = blk[i]; unsigned kk = k;
            for (unsigned p = 0; b; ++p, b >>= 1) if (b & 1) { nr += g_binom[62 - p][kk]; --kk; }
            code.put(nr, sp);
        }
    }
    if (!rank.empty()) rank.back() = sum_rank;
    std::vector<uint8_t> o;
    append_u64(o, m_size);
    append_int_vector(o, bt, 6);
    uint64_t nb = total_bits > 64 ? total_bits : 64;
    append_u64(o, nb); append_words(o, code.w, nb);
    append_int_vector(o, btnrp, hi_bit(total_bits) + 1);
    append_int_vector(o, rank, hi_bit(total_ones) + 1);
    { std::vector<uint64_t> w((nsb + 63) / 64, 0); for (size_t s = 0; s < nsb; ++s) if (invert[s]) w[s >> 6] |= uint64_t(1) << (s & 63);
      append_u64(o, nsb); append_words(o, w, nsb); }
    return o;
}

// ---------------------------------------------------------------------------------------------------
// EDS text -> string collection (eds_to_fasta.cpp:55-140)
// ---------------------------------------------------------------------------------------------------
void parse_eds_text(const std::vector<uint8_t>& text, std::vector<uint8_t>& letters, std::vector<uint64_t>& str_off,
                    std::vector<uint8_t>& seg_first) {
    letters.clear(); str_off.clear(); seg_first.clear();
    size_t i0 = 0;
    while (i0 < text.size() && (text[i0] == ' ' || text[i0] == '\n' || text[i0] == '\r' || text[i0] == '\t')) ++i0;
    if (i0 >= text.size() || text[i0] != '{') EDS_FAIL(EDSBWT_E_FORMAT, "Error: the string does not start with {.");
    for (size_t i = i0; i < text.size(); ++i) {
        uint8_t ch = text[i];
        if (ch == '{') { str_off.push_back(letters.size()); seg_first.push_back(1); }
        else if (ch == ',') { str_off.push_back(letters.size()); seg_first.push_back(0); }
        else if (ch == '}' || ch == 'E' || ch == ' ' || ch == '\n' || ch == '\r' || ch == '\t') {}
        else if (ch == 'Z' || ch == '#') EDS_FAIL(EDSBWT_E_FORMAT, "Error: the input contains character %c, which is reserved (EMPTY_CHAR / TERMINATE_CHAR in Parameters.h)", ch);
        else letters.push_back(ch);
    }
    str_off.push_back(letters.size());
}

// ---------------------------------------------------------------------------------------------------
// index files -> HostIndex   (EDSBWT::recoverInfo, EDSBWTsearch.cpp:1557-1656; ctor :150-168)
// ---------------------------------------------------------------------------------------------------
void load_host_index(const std::string& base, bool verify, HostIndex& h) {
    std::vector<uint8_t> info = read_file(base + "_info.aux");
    if (info.size() < 10) EDS_FAIL(EDSBWT_E_FORMAT, "%s_info.aux is too short", base.c_str());
    memcpy(&h.n, &info[0], 4); memcpy(&h.n_text, &info[4], 4); h.sigma = info[8];
    if (h.sigma < 2) EDS_FAIL(EDSBWT_E_FORMAT, "alphabet of size %u", h.sigma);
    if (h.sigma > 8) EDS_FAIL(EDSBWT_E_ALPHABET, "alphabet of size %u: this build packs at most 7 letters + end-marker (3-bit Occ blocks)", h.sigma);
    size_t need = 9 + h.sigma + 4ull * h.n_text + 4ull * h.sigma * h.sigma;
    if (info.size() != need) EDS_FAIL(EDSBWT_E_FORMAT, "%s_info.aux has %zu bytes, expected %zu", base.c_str(), info.size(), need);
    memcpy(h.alphabet, &info[9], h.sigma);
    if (h.alphabet[0] != '#') EDS_FAIL(EDSBWT_E_FORMAT, "alphabet[0] is 0x%02x, expected the end-marker '#'", h.alphabet[0]);
    size_t off = 9 + h.sigma;
    h.eof_flat.resize(h.n_text);
    if (h.n_text) memcpy(h.eof_flat.data(), &info[off], 4ull * h.n_text);
    off += 4ull * h.n_text;
    memcpy(h.table_occ, &info[off], 4ull * h.sigma * h.sigma);
    h.ebwt.clear(); h.ebwt.reserve(h.n);
    uint64_t eof_seen = 0;
    for (uint32_t p = 0; p < h.sigma; ++p) {
        h.pile_start[p] = (uint32_t)h.ebwt.size();
        std::vector<uint8_t> pile = read_file(base + "_bwt_" + std::to_string(p) + ".aux");
        uint64_t row_sum = 0;
        for (uint32_t c = 0; c < h.sigma; ++c) row_sum += h.table_occ[p * h.sigma + c];
        if (pile.size() != row_sum) EDS_FAIL(EDSBWT_E_FORMAT, "%s_bwt_%u.aux has %zu bytes, tableOcc row sums to %llu", base.c_str(), p, pile.size(), (unsigned long long)row_sum);
        if (verify) {
            // _bv_<p>.aux must be the '#' bitmap of the pile (da_to_everything.cpp:171,219,234-236)
            std::vector<uint8_t> bv = rrr_decode(read_file(base + "_bv_" + std::to_string(p) + ".aux"));
            if (bv.size() != pile.size()) EDS_FAIL(EDSBWT_E_FORMAT, "%s_bv_%u.aux has %zu bits, pile has %zu rows", base.c_str(), p, bv.size(), pile.size());
            uint32_t cnt[256] = {0};
            for (size_t i = 0; i < pile.size(); ++i) {
                cnt[pile[i]]++;
                if ((pile[i] == '#') != (bv[i] != 0)) EDS_FAIL(EDSBWT_E_FORMAT, "%s_bv_%u.aux disagrees with the pile at row %zu", base.c_str(), p, i);
            }
            for (uint32_t c = 0; c < h.sigma; ++c)
                if (cnt[h.alphabet[c]] != h.table_occ[p * h.sigma + c]) EDS_FAIL(EDSBWT_E_FORMAT, "tableOcc[%u][%u] = %u but pile %u holds %u", p, c, h.table_occ[p * h.sigma + c], p, cnt[h.alphabet[c]]);
        }
        eof_seen += h.table_occ[p * h.sigma + 0];
        h.ebwt.insert(h.ebwt.end(), pile.begin(), pile.end());
    }
    h.pile_start[h.sigma] = (uint32_t)h.ebwt.size();
    if (h.ebwt.size() != h.n) EDS_FAIL(EDSBWT_E_FORMAT, "piles hold %zu rows, info says %u", h.ebwt.size(), h.n);
    if (eof_seen != h.n_text) EDS_FAIL(EDSBWT_E_FORMAT, "tableOcc counts %llu end-markers, info says %u strings", (unsigned long long)eof_seen, h.n_text);
    if (h.pile_start[1] != h.n_text) EDS_FAIL(EDSBWT_E_FORMAT, "pile 0 has %u rows, expected one per string (%u)", h.pile_start[1], h.n_text);
    h.seg_first_bit = bitvector_decode(read_file(base + ".bitvector"));
    if (h.seg_first_bit.size() != h.n_text) EDS_FAIL(EDSBWT_E_FORMAT, "%s.bitvector has %zu bits, expected %u", base.c_str(), h.seg_first_bit.size(), h.n_text);
    if (h.n_text && !h.seg_first_bit[0]) EDS_FAIL(EDSBWT_E_FORMAT, ".bitvector: string 0 does not open a segment");
    h.n_seg = 0;
    for (uint8_t b : h.seg_first_bit) h.n_seg += b;
}

}  // namespace edsbwt

// ---------------------------------------------------------------------------------------------------
// C ABI: host-only entry points
// ---------------------------------------------------------------------------------------------------
using namespace edsbwt;
namespace edsbwt { const char* last_error_cstr(); }

namespace {
// read-only view of a regular file (the pattern file can be hundreds of MB: no copy through a buffer)
struct MappedFile {
    const uint8_t* p = nullptr; size_t n = 0;
    explicit MappedFile(const char* path) {
        const int fd = open(path, O_RDONLY);
        if (fd < 0) EDS_FAIL(EDSBWT_E_IO, "Error opening %s: %s", path, strerror(errno));
        struct stat sb;
        if (fstat(fd, &sb) != 0 || !S_ISREG(sb.st_mode)) { close(fd); EDS_FAIL(EDSBWT_E_IO, "Error opening %s: not a regular file", path); }
        n = (size_t)sb.st_size;
        if (n) {
            void* m = mmap(nullptr, n, PROT_READ, MAP_PRIVATE, fd, 0);
            if (m == MAP_FAILED) { const int e = errno; close(fd); EDS_FAIL(EDSBWT_E_IO, "Error mapping %s: %s", path, strerror(e)); }
            p = (const uint8_t*)m;
            madvise(m, n, MADV_SEQUENTIAL);
        }
        close(fd);
    }
    ~MappedFile() { if (p) munmap((void*)p, n); }
    MappedFile(const MappedFile&) = delete;
    MappedFile& operator=(const MappedFile&) = delete;
};
// no exception may cross the C ABI (same mapping as api.cu's guard)
template <class F> int host_guard(F&& f) {
    try { return f(); }
    catch (const Failure& e) { return e.code; }
    catch (const std::bad_alloc&) { set_error("host out of memory"); return EDSBWT_E_IO; }
    catch (const std::exception& e) { set_error("exception: %s", e.what()); return EDSBWT_E_ARG; }
}
}  // namespace

extern "C" {

const char* edsbwt_last_error(void) { return edsbwt::last_error_cstr(); }
const char* edsbwt_version(void) { return "edsbwt-b200 0.1 (sm_100a)"; }
void edsbwt_free(void* p) { free(p); }
const char* edsbwt_strerror(int code) {
    switch (code) {
        case EDSBWT_OK: return "ok";
        case EDSBWT_E_IO: return "I/O error";
        case EDSBWT_E_FORMAT: return "inconsistent index files";
        case EDSBWT_E_ALPHABET: return "alphabet not supported";
        case EDSBWT_E_CUDA: return "CUDA error / no device";
        case EDSBWT_E_ARG: return "bad argument";
        case EDSBWT_E_UNSUPPORTED: return "unsupported";
        case EDSBWT_E_NOSPACE: return "occurrence buffer too small";
        default: return "unknown error";
    }
}

int edsbwt_read_patterns(const char* path, uint8_t** bytes, uint64_t** off, uint32_t* npat) {
    if (!path || !bytes || !off || !npat) return EDSBWT_E_ARG;
    return host_guard([&]() -> int {
        const MappedFile f(path);
        // std::getline semantics (EDSBWTsearch.cpp:207): '\n' separated, a final '\n' adds no line.  Two memchr sweeps (count,
        // then copy) into exactly sized buffers: a 10 M-line file is read in a fraction of a second
        const uint8_t* p = f.p; const size_t n = f.n;
        size_t nl = 0;
        for (const uint8_t* q = p; n && (q = (const uint8_t*)memchr(q, '\n', (size_t)(p + n - q))) != nullptr; ++q) ++nl;
        const size_t nlines = nl + ((n && p[n - 1] != '\n') ? 1 : 0);
        if (nlines > 0xffffffffull) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "%s holds %zu patterns, at most 2^32-1 per batch", path, nlines);
        uint8_t* pb = (uint8_t*)malloc(n - nl ? n - nl : 1); uint64_t* po = (uint64_t*)malloc((nlines + 1) * 8);
        if (!pb || !po) { free(pb); free(po); set_error("host out of memory"); return EDSBWT_E_IO; }
        size_t w = 0, k = 0;
        for (size_t i = 0; i < n;) {
            const uint8_t* e = (const uint8_t*)memchr(p + i, '\n', n - i);
            const size_t j = e ? (size_t)(e - p) : n;
            po[k++] = w;
            memcpy(pb + w, p + i, j - i); w += j - i;
            i = j + 1;
        }
        po[k] = w;
        *bytes = pb; *off = po; *npat = (uint32_t)nlines;
        return EDSBWT_OK;
    });
}

// stringCheck (stringCheck.cpp:11-84): brackets every non-degenerate stretch of an EDS text, drops <...> comments and
// refuses the character the tool chain reserves for the empty word (EMPTY_CHAR 'Z', Parameters.h:34).  Same decisions,
// in the same order, as the reference's get()/peek() loop -- including its treatment of a trailing newline as a letter.
int edsbwt_string_check(const char* in_path, const char* out_path) {
    if (!in_path || !out_path) return EDSBWT_E_ARG;
    const uint8_t EMPTY = 'Z';
    return host_guard([&]() -> int {
        const std::vector<uint8_t> f = read_file(in_path);
        std::vector<uint8_t> o; o.reserve(f.size() + f.size() / 8 + 16);
        const size_t n = f.size();
        auto reserved = [&]() -> int {
            set_error("the input contains character %c, which is used by this tool to represent an empty string", EMPTY);
            return EDSBWT_E_FORMAT;
        };
        size_t i = 0;
        if (n) {
            const uint8_t c = f[i++];
            if (c == EMPTY) return reserved();
            if (c != '{') o.push_back('{');
            o.push_back(c);
        }
        while (i < n) {
            uint8_t c = f[i++];
            if (c == EMPTY) return reserved();
            const int next = i < n ? (int)f[i] : -1;
            if (c == '}' && next != '{' && next != -1) { o.push_back('}'); o.push_back('{'); }
            else if (c != '}' && next == '{') { o.push_back(c); o.push_back('}'); }
            else if (c == '<') {
                while (i < n && c != '>') c = f[i++];            // the comment, through its closing '>' (the reference never ends without one)
                if (i < n && f[i] == '{') o.push_back('}');
            }
            else if (next == -1 && c != '}') { o.push_back(c); o.push_back('}'); }
            else o.push_back(c);
        }
        write_file(out_path, o.data(), o.size());
        return EDSBWT_OK;
    });
}

int edsbwt_write_csv(const char* path, const edsbwt_occ* occ, uint64_t nocc, uint32_t pat_base, int move, int append) {
    if (!path || (nocc && !occ)) return EDSBWT_E_ARG;
    return host_guard([&]() -> int {
        std::vector<char> buf; buf.reserve(1 << 20);          // filled to at most 1 MiB: no allocation while the file is open
        FILE* f = fopen(path, append ? "ab" : "wb");
        if (!f) { set_error("ERROR opening file %s to write output", path); return EDSBWT_E_IO; }
        bool ok = true;
        if (!append) ok = fputs(move ? "#Pat\t$_i\tD[i]\tS_j\tS_j[r] \n" : "#Pat\t$_i\tD[i]\tS_j\tS_j[r]\n", f) >= 0;
        char line[96];
        for (uint64_t i = 0; i < nocc && ok; ++i) {
            int k = snprintf(line, sizeof line, "%u\t%u\t%u\t%u\t%u\n", occ[i].pat + pat_base + 1, occ[i].str, occ[i].seg, occ[i].idx_in_seg, occ[i].off);
            buf.insert(buf.end(), line, line + k);
            if (buf.size() > (1 << 20) - 128) { ok = fwrite(buf.data(), 1, buf.size(), f) == buf.size(); buf.clear(); }
        }
        if (ok && !buf.empty()) ok = fwrite(buf.data(), 1, buf.size(), f) == buf.size();
        const int werr = errno;
        if (fclose(f) != 0 && ok) { ok = false; }
        if (!ok) { set_error("ERROR writing %s: %s", path, strerror(werr ? werr : errno)); return EDSBWT_E_IO; }
        return EDSBWT_OK;
    });
}

}  // extern "C"
