// TEST INFRASTRUCTURE (oracle) -- never included by the product path.
//
// Private stand-in for <sdsl/bit_vectors.hpp> so that the UNMODIFIED reference sources
// (/root/reference/{mainEDS-BWT,EDSBWTsearch,Sorting}.cpp) compile here without sdsl-lite
// (build 2).  Only what the reference's search path uses is provided:
//
//   bit_vector + load_from_file        EDSBWTsearch.cpp:158      (<base>.bitvector)
//   rank_support_v<1>                  EDSBWTsearch.cpp:162,1291,1322
//   bit_vector::select_1_type          EDSBWTsearch.cpp:167-168,1293-1296,1323
//   rrr_vector<> + ::rank_1_type       EDSBWTsearch.cpp:1598-1603,1268-1269,1313-1314,1343,1617
//
// The rrr_vector here reads the REAL sdsl rrr_vector<63,int_vector<>,32> serialisation
// (inversion semantics as probed against the prebuilt reference binary:
// an inverted superblock stores m_bt = 63 - popcount, the block code is that of the
// original block) and decodes it to a plain bitmap with an O(1) sampled rank, so the same
// index files feed the prebuilt binary and this build.
#pragma once
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <string>
#include <vector>

namespace sdsl {

namespace shim_detail {
inline uint64_t binom(unsigned n, unsigned k) {
    static uint64_t tab[64][64];
    static bool init = false;
    if (!init) {
        for (unsigned a = 0; a < 64; ++a) {
            tab[a][0] = 1;
            for (unsigned b = 1; b < 64; ++b) tab[a][b] = (a == 0) ? 0 : tab[a - 1][b - 1] + tab[a - 1][b];
        }
        init = true;
    }
    return tab[n][k];
}
inline unsigned space_for_bt(unsigned k) {
    uint64_t c = binom(63, k);
    if (c == 1) return 0;
    return 64 - __builtin_clzll(c);
}
struct plain_bits {
    uint64_t nbits = 0;
    std::vector<uint64_t> w;
    std::vector<uint64_t> cum;   // ones before each word
    bool get(uint64_t i) const { return (w[i >> 6] >> (i & 63)) & 1; }
    void set(uint64_t i) { w[i >> 6] |= (uint64_t(1) << (i & 63)); }
    void build_rank() {
        cum.assign(w.size() + 1, 0);
        for (size_t j = 0; j < w.size(); ++j) cum[j + 1] = cum[j] + __builtin_popcountll(w[j]);
    }
    uint64_t rank1(uint64_t i) const {   // ones in [0,i)
        uint64_t r = cum[i >> 6];
        if (i & 63) r += __builtin_popcountll(w[i >> 6] & ((uint64_t(1) << (i & 63)) - 1));
        return r;
    }
};
inline uint64_t rd64(std::istream& in) { uint64_t v = 0; in.read(reinterpret_cast<char*>(&v), 8); return v; }
inline uint8_t rd8(std::istream& in) { uint8_t v = 0; in.read(reinterpret_cast<char*>(&v), 1); return v; }
struct packed {
    uint64_t nbits = 0; unsigned width = 1; std::vector<uint64_t> w;
    void load(std::istream& in, bool has_width) {
        nbits = rd64(in);
        width = has_width ? rd8(in) : 1;
        w.assign((nbits + 63) / 64 + 1, 0);
        in.read(reinterpret_cast<char*>(w.data()), ((nbits + 63) / 64) * 8);
    }
    uint64_t bits_at(uint64_t pos, unsigned len) const {
        if (len == 0) return 0;
        uint64_t lo = w[pos >> 6] >> (pos & 63);
        unsigned got = 64 - (pos & 63);
        if (got < len) lo |= w[(pos >> 6) + 1] << got;
        return len == 64 ? lo : (lo & ((uint64_t(1) << len) - 1));
    }
    uint64_t at(uint64_t j) const { return bits_at(j * width, width); }
    uint64_t count() const { return width ? nbits / width : 0; }
};
}  // namespace shim_detail

class bit_vector {
public:
    shim_detail::plain_bits b;
    typedef uint64_t size_type;
    uint64_t size() const { return b.nbits; }
    bool operator[](uint64_t i) const { return b.get(i); }
    void load(std::istream& in) {
        b.nbits = shim_detail::rd64(in);
        b.w.assign((b.nbits + 63) / 64, 0);
        in.read(reinterpret_cast<char*>(b.w.data()), b.w.size() * 8);
        b.build_rank();
    }
    class select_1_type {
        const bit_vector* v = nullptr;
        std::vector<uint64_t> pos;
    public:
        select_1_type() {}
        explicit select_1_type(const bit_vector* bv) : v(bv) {
            for (uint64_t i = 0; i < bv->size(); ++i) if ((*bv)[i]) pos.push_back(i);
        }
        uint64_t operator()(uint64_t k) const { return pos.at(k - 1); }   // position of the k-th one (k >= 1)
    };
};

template <uint8_t t_b = 1>
class rank_support_v {
    const bit_vector* v = nullptr;
public:
    rank_support_v() {}
    explicit rank_support_v(const bit_vector* bv) : v(bv) {}
    uint64_t operator()(uint64_t i) const { return v->b.rank1(i); }
    uint64_t size() const { return v->size(); }
};

template <uint16_t t_bs = 63, class t_rac = void, uint16_t t_k = 32>
class rrr_vector {
public:
    shim_detail::plain_bits b;
    uint64_t size() const { return b.nbits; }
    void load(std::istream& in) {
        using namespace shim_detail;
        uint64_t m_size = rd64(in);
        packed bt, btnr, btnrp, rk, inv;
        bt.load(in, true); btnr.load(in, false); btnrp.load(in, true); rk.load(in, true); inv.load(in, false);
        b.nbits = m_size;
        b.w.assign((m_size + 63) / 64 + 1, 0);
        uint64_t nreal = (m_size + 62) / 63, pos = 0;
        for (uint64_t j = 0; j < nreal; ++j) {
            if (j % 32 == 0) pos = btnrp.at(j / 32);
            unsigned k = (unsigned)bt.at(j);
            unsigned len = space_for_bt(k);
            uint64_t nr = btnr.bits_at(pos, len);
            pos += len;
            if (inv.bits_at(j / 32, 1)) k = 63 - k;
            for (unsigned p = 0; p < 63 && k > 0; ++p) {
                uint64_t c = binom(62 - p, k);
                if (nr >= c) {
                    nr -= c; --k;
                    uint64_t i = 63 * j + p;
                    if (i < m_size) b.set(i);
                }
            }
        }
        b.w.resize((m_size + 63) / 64);
        b.build_rank();
    }
    class rank_1_type {
        const rrr_vector* v = nullptr;
    public:
        rank_1_type() {}
        explicit rank_1_type(const rrr_vector* rv) : v(rv) {}
        uint64_t operator()(uint64_t i) const { return v->b.rank1(i); }
        uint64_t size() const { return v->size(); }
    };
};

template <class T>
bool load_from_file(T& v, const std::string& file) {
    std::ifstream in(file, std::ios::binary);
    if (!in) return false;
    v.load(in);
    return true;
}

}  // namespace sdsl
