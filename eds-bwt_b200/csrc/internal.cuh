// Internal declarations of libedsbwt_cuda.so (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/edsbwt.h"

namespace edsbwt {

// ---------------------------------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
struct Failure { int code; };
struct OccOverflow {};      // a search found more occurrences than Index::occ_limit: the chunked callers halve the chunk and retry
#define EDS_FAIL(code, ...) do { ::edsbwt::set_error(__VA_ARGS__); throw ::edsbwt::Failure{code}; } while (0)
#define CUDA_OK(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) { \
    ::edsbwt::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); \
    throw ::edsbwt::Failure{EDSBWT_E_CUDA}; } } while (0)

// ---------------------------------------------------------------------------------------------------
// growable device buffer
// ---------------------------------------------------------------------------------------------------
template <class T>
struct DevBuf {
    T* p = nullptr; size_t cap = 0;
    DevBuf() {}
    DevBuf(const DevBuf&) = delete; DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    // grow (contents NOT preserved unless keep = true)
    void ensure(size_t n, cudaStream_t s = 0, bool keep = false, size_t keep_n = 0) {
        if (n <= cap) return;
        size_t ncap = n + n / 4 + 64;
        T* q = nullptr;
        CUDA_OK(cudaMalloc(&q, ncap * sizeof(T)));
        if (keep_n > cap) keep_n = cap;           // a cursor that ran past the end (staging overflow) keeps what exists
        if (keep && p && keep_n) { CUDA_OK(cudaMemcpyAsync(q, p, keep_n * sizeof(T), cudaMemcpyDeviceToDevice, s)); CUDA_OK(cudaStreamSynchronize(s)); }
        if (p) cudaFree(p);
        p = q; cap = ncap;
    }
    // static index arrays: exactly n elements, no growth slack (the seed table and the locate table are gigabytes)
    void exact(size_t n) {
        if (n <= cap) return;
        T* q = nullptr;
        CUDA_OK(cudaMalloc(&q, n * sizeof(T)));
        if (p) cudaFree(p);
        p = q; cap = n;
    }
    size_t bytes() const { return cap * sizeof(T); }
};

// ---------------------------------------------------------------------------------------------------
// Occ layout (DESIGN.md "HBM layout"): one 64-byte block per 128 eBWT rows = 4 x uint4, lane j holds
//   .x  number of occurrences of letter j (alphabet index j+1) in rows [0, 128*blk)
//   .y  bit-plane 0 (low bit of the 2-bit letter code) of rows 128*blk + 32*j .. +31
//   .z  bit-plane 1
//   .w  end-marker mask of those 32 rows ('#' rows have planes 0)
// so that a 4-lane group fetches one block with ONE coalesced 128-bit load per lane and the rank of any
// letter and of '#' (rank_# = pos - sum of letter ranks) come out of the same 64 bytes.
// ---------------------------------------------------------------------------------------------------
constexpr uint32_t OCC_ROWS = 128;
constexpr int FUSED_THREADS = 128;
constexpr uint32_t FUSED_BIG_CAP = 192;      // 192 slots x 128 threads x 8 B = 192 KB of shared memory: one CTA per SM

struct IndexView {
    const uint4* occ;            // (n/128 + 1) blocks
    uint32_t n, n_text, n_seg, fsi;      // fsi = first_symbol_index = last string of segment 1
    uint32_t Cl[5];              // Cl[c] = global row of the first suffix starting with letter c; Cl[4] = n
    uint32_t nletters;
    const uint32_t* eof_flat;    // EOF_ID flattened in pile order == order of '#' rows in the eBWT
    const uint32_t* str_seg;     // string -> 1-based segment        (rank on .bitvector)
    const uint32_t* seg_first;   // [1..n_seg+1] first string of segment (select), [n_seg+1] = n_text
    const uint32_t* seg_low;     // [1..n_seg] lowest segment reachable from s through segments holding ""
    const uint4*    sa4;         // row -> (string, offset in string, segment, index in segment); null in WALK mode
    const uint2*    link_rec;    // per end-marker rank r (order of '#' rows in the eBWT): pile-0 row range [x,y] of all
                                 // strings of the segments reachable from string eof_flat[r] (x > y: none) -- the whole
                                 // preceding_dollars_finder + empty-string closure of link() in one 8-byte gather
    const uint2*    str_info;    // string -> (1-based segment, index inside the segment)
    // wide alphabets (5..7 letters + end-marker, e.g. {#,A,C,G,N,T}): 3-bit letter codes, 96-byte Occ blocks (below).
    // wide == 0: everything above, Cl[] included, is the 2-bit layout.
    uint32_t wide;
    uint32_t Clw[9];             // wide: Clw[c] = first row of the pile of letter c, n for c >= nletters
};

#ifdef __CUDACC__
// rank of letter c and of '#' in rows [0,pos), computed by a 4-lane group (lanes hold lane4 = 0..3).
// Every lane of the warp must call it (full-mask shuffles).
__device__ __forceinline__ uint4 occ_load(const uint4* __restrict__ occ, uint32_t blk, int lane4) {
    return __ldg(occ + (size_t)blk * 4 + lane4);
}
__device__ __forceinline__ void rank4_from(uint4 v, uint32_t pos, int c, int lane4, uint32_t& rc, uint32_t& rd) {
    int below = (int)(pos & (OCC_ROWS - 1)) - 32 * lane4;
    uint32_t lm = below <= 0 ? 0u : (below >= 32 ? 0xffffffffu : ((1u << below) - 1u));
    uint32_t m = ((c & 1) ? v.y : ~v.y) & ((c & 2) ? v.z : ~v.z) & ~v.w & lm;
    uint32_t pc = __popc(m) + (lane4 == c ? v.x : 0u);
    uint32_t pd = __popc(v.w & lm) - v.x;                     // wraps; the 4-lane sum is exact mod 2^32
    pc += __shfl_xor_sync(0xffffffffu, pc, 1); pc += __shfl_xor_sync(0xffffffffu, pc, 2);
    pd += __shfl_xor_sync(0xffffffffu, pd, 1); pd += __shfl_xor_sync(0xffffffffu, pd, 2);
    rc = pc;
    rd = pd + (pos & ~(OCC_ROWS - 1));
}
// single-thread variants (index build, LF-walk locate)
__device__ __forceinline__ uint32_t occ_symbol(const uint4* __restrict__ occ, uint32_t row) {   // 0 = '#', 1+c = letter c
    uint4 v = __ldg(occ + (size_t)(row >> 7) * 4 + ((row >> 5) & 3));
    uint32_t b = row & 31;
    if ((v.w >> b) & 1) return 0;
    return 1 + (((v.y >> b) & 1) | (((v.z >> b) & 1) << 1));
}
__device__ __forceinline__ uint32_t occ_rank1(const uint4* __restrict__ occ, uint32_t pos, int c) {
    uint32_t blk = pos >> 7, off = pos & 127;
    const uint4* b = occ + (size_t)blk * 4;
    uint32_t r = __ldg(&b[c].x);
    for (int j = 0; j < 4; ++j) {
        int below = (int)off - 32 * j;
        if (below <= 0) break;
        uint4 v = __ldg(&b[j]);
        uint32_t lm = below >= 32 ? 0xffffffffu : ((1u << below) - 1u);
        r += __popc(((c & 1) ? v.y : ~v.y) & ((c & 2) ? v.z : ~v.z) & ~v.w & lm);
    }
    return r;
}
__device__ __forceinline__ uint32_t occ_rank_eof(const uint4* __restrict__ occ, uint32_t pos) {
    uint32_t blk = pos >> 7, off = pos & 127;
    const uint4* b = occ + (size_t)blk * 4;
    uint32_t r = blk * OCC_ROWS;
    for (int j = 0; j < 4; ++j) {
        uint4 v = __ldg(&b[j]);
        r -= v.x;
        int below = (int)off - 32 * j;
        if (below > 0) r += __popc(v.w & (below >= 32 ? 0xffffffffu : ((1u << below) - 1u)));
    }
    return r;
}

// ---- wide alphabets (5..7 letters): one 96-byte block per 128 rows = 6 x uint4
//   [j], j = 0..3   rows 32j..32j+31: .x .y .z bit-planes 0..2 of the 3-bit letter code, .w end-marker mask ('#' rows have planes 0)
//   [4], [5]        occurrences of letters 0..3 / 4..7 in rows [0, 128*blk)
// A 4-lane group ranks any letter and '#' with one 128-bit + one 64-bit load per lane (lane j: entry j + the counts of letters
// 2j, 2j+1); the one-thread-per-pattern kernel fetches the whole block with three 256-bit loads.
constexpr uint32_t OCCW_U4 = 6;
__device__ __forceinline__ uint32_t occw_match(const uint4& v, int c) {
    return ((c & 1) ? v.x : ~v.x) & ((c & 2) ? v.y : ~v.y) & ((c & 4) ? v.z : ~v.z) & ~v.w;
}
__device__ __forceinline__ void rank4w_from(uint4 va, uint2 cn, uint32_t pos, int c, int lane4, uint32_t& rc, uint32_t& rd) {
    int below = (int)(pos & (OCC_ROWS - 1)) - 32 * lane4;
    uint32_t lm = below <= 0 ? 0u : (below >= 32 ? 0xffffffffu : ((1u << below) - 1u));
    uint32_t pc = __popc(occw_match(va, c) & lm) + (lane4 == (c >> 1) ? ((c & 1) ? cn.y : cn.x) : 0u);
    uint32_t pd = __popc(va.w & lm) - (cn.x + cn.y);          // wraps; the 4-lane sum is exact mod 2^32
    pc += __shfl_xor_sync(0xffffffffu, pc, 1); pc += __shfl_xor_sync(0xffffffffu, pc, 2);
    pd += __shfl_xor_sync(0xffffffffu, pd, 1); pd += __shfl_xor_sync(0xffffffffu, pd, 2);
    rc = pc;
    rd = pd + (pos & ~(OCC_ROWS - 1));
}
__device__ __forceinline__ uint32_t occw_symbol(const uint4* __restrict__ occ, uint32_t row) {   // 0 = '#', 1+c = letter c
    uint4 v = __ldg(occ + (size_t)(row >> 7) * OCCW_U4 + ((row >> 5) & 3));
    uint32_t b = row & 31;
    if ((v.w >> b) & 1) return 0;
    return 1 + (((v.x >> b) & 1) | (((v.y >> b) & 1) << 1) | (((v.z >> b) & 1) << 2));
}
__device__ __forceinline__ uint32_t occw_rank1(const uint4* __restrict__ occ, uint32_t pos, int c) {
    uint32_t blk = pos >> 7, off = pos & 127;
    const uint4* b = occ + (size_t)blk * OCCW_U4;
    uint32_t r = __ldg(reinterpret_cast<const uint32_t*>(b + 4) + c);
    for (int j = 0; j < 4; ++j) {
        int below = (int)off - 32 * j;
        if (below <= 0) break;
        uint4 v = __ldg(&b[j]);
        r += __popc(occw_match(v, c) & (below >= 32 ? 0xffffffffu : ((1u << below) - 1u)));
    }
    return r;
}
__device__ __forceinline__ uint32_t occw_rank_eof(const uint4* __restrict__ occ, uint32_t pos) {
    uint32_t blk = pos >> 7, off = pos & 127;
    const uint4* b = occ + (size_t)blk * OCCW_U4;
    const uint4 c0 = __ldg(&b[4]), c1 = __ldg(&b[5]);
    uint32_t r = blk * OCC_ROWS - (c0.x + c0.y + c0.z + c0.w + c1.x + c1.y + c1.z + c1.w);
    for (int j = 0; j < 4; ++j) {
        int below = (int)off - 32 * j;
        if (below <= 0) break;
        r += __popc(__ldg(&b[j]).w & (below >= 32 ? 0xffffffffu : ((1u << below) - 1u)));
    }
    return r;
}
// one LF step of a single row, either layout (index build, LF-walk locate): false when the row holds '#'
__device__ __forceinline__ bool lf_row_any(const IndexView& ix, uint32_t& row) {
    if (ix.wide) {
        const uint32_t s = occw_symbol(ix.occ, row);
        if (s == 0) return false;
        row = ix.Clw[s - 1] + occw_rank1(ix.occ, row, (int)s - 1);
    } else {
        const uint32_t s = occ_symbol(ix.occ, row);
        if (s == 0) return false;
        row = ix.Cl[s - 1] + occ_rank1(ix.occ, row, (int)s - 1);
    }
    return true;
}
__device__ __forceinline__ uint32_t rank_eof_any(const IndexView& ix, uint32_t row) {
    return ix.wide ? occw_rank_eof(ix.occ, row) : occ_rank_eof(ix.occ, row);
}

// ---- one-thread-per-pattern kernels (fused.cu, move.cu): whole 64-byte blocks in registers, 2-bit pattern windows
struct Blk { uint32_t w[16]; };      // w[4j] = count of letter j before the block, w[4j+1] plane0, w[4j+2] plane1, w[4j+3] '#' mask of rows 32j..32j+31

__device__ __forceinline__ void ld256(const void* p, uint32_t* r) {
    asm volatile("ld.global.nc.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(p));
}
__device__ __forceinline__ void load_blk(const uint4* __restrict__ occ, uint32_t blk, Blk& b) {
    const uint4* p = occ + (size_t)blk * 4;
    ld256(p, b.w); ld256(p + 2, b.w + 8);
}
// Pattern characters travel as 2-bit codes in a 64-bit window, 32 characters per fetch: n <= 32 bytes at p -> codes
// (character i at bits 2i, 2i+1).  Returns false when a byte is not a letter of the index alphabet.  One or two 128-bit
// loads when p is 16-byte aligned (every fixed-length batch whose length is a multiple of 16), 32-bit loads when 4-byte
// aligned, byte loads otherwise.  Reading the aligned 16-byte / 4-byte unit that holds the last valid byte stays inside
// the allocation (cudaMalloc granularity), so short tails need no special case.
__device__ __forceinline__ bool fetch_codes(const uint8_t* __restrict__ p, uint32_t n, const uint8_t* s_lut, uint64_t& codes) {
    uint32_t w[8];
    const uint32_t al = (uint32_t)(uintptr_t)p & 15u;
    if (al == 0) {
        const uint4 a = __ldg((const uint4*)p);
        const uint4 b = n > 16 ? __ldg((const uint4*)p + 1) : make_uint4(0, 0, 0, 0);
        w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    } else if ((al & 3u) == 0) {
        #pragma unroll
        for (int j = 0; j < 8; ++j) w[j] = 4u * j < n ? __ldg((const uint32_t*)p + j) : 0u;
    } else {
        #pragma unroll
        for (int j = 0; j < 8; ++j) {
            uint32_t v = 0;
            #pragma unroll
            for (int t = 0; t < 4; ++t) if (4u * j + t < n) v |= (uint32_t)__ldg(p + 4 * j + t) << (8 * t);
            w[j] = v;
        }
    }
    uint32_t lo = 0, hi = 0, bad = 0;                 // branch-free: codes of all 32 slots, invalid slots masked afterwards
    #pragma unroll
    for (int i = 0; i < 16; ++i) {
        const uint32_t c0 = s_lut[(w[i >> 2] >> (8 * (i & 3))) & 255u];
        const uint32_t c1 = s_lut[(w[4 + (i >> 2)] >> (8 * (i & 3))) & 255u];
        lo |= (c0 & 3u) << (2 * i); hi |= (c1 & 3u) << (2 * i);
        bad |= (c0 > 3u ? 1u : 0u) << i; bad |= (c1 > 3u ? 1u : 0u) << (16 + i);
    }
    const uint64_t cw = ((uint64_t)hi << 32) | lo;
    codes = n >= 32 ? cw : cw & ((1ull << (2 * n)) - 1);
    return (bad & (n >= 32 ? 0xffffffffu : ((1u << n) - 1u))) == 0;
}
// rank of letter c (x0/x1 = plane inversion masks of c) and number of '#' in rows [0,off) of the block, off in [0,128)
__device__ __forceinline__ void rank_in_blk(const Blk& b, uint32_t off, int c, uint32_t x0, uint32_t x1, uint32_t& rc, uint32_t& nmk) {
    uint32_t r = c == 0 ? b.w[0] : c == 1 ? b.w[4] : c == 2 ? b.w[8] : b.w[12];
    uint32_t k = 0;
    #pragma unroll
    for (int j = 0; j < 4; ++j) {
        int below = (int)off - 32 * j;
        uint32_t lm = below <= 0 ? 0u : (below >= 32 ? 0xffffffffu : ((1u << below) - 1u));
        uint32_t mk = b.w[4 * j + 3] & lm;
        r += __popc((b.w[4 * j + 1] ^ x0) & (b.w[4 * j + 2] ^ x1) & ~b.w[4 * j + 3] & lm);
        k += __popc(mk);
    }
    rc = r; nmk = k;
}
__device__ __forceinline__ uint32_t letters_before(const Blk& b) { return b.w[0] + b.w[4] + b.w[8] + b.w[12]; }
// mask of the bits below position `below` of a 32-bit word, below clamped to [0, 32] (shl.b32 clamps its shift amount)
__device__ __forceinline__ uint32_t low_mask(int below) {
    uint32_t r; const int bb = below < 0 ? 0 : below;
    asm("shl.b32 %0, 1, %1;" : "=r"(r) : "r"(bb));
    return r - 1u;
}
// LF of the row range [a, b) by letter c: rca / rcb = rank of c before a / b; rda / rdb = rank of '#' before a / b (both 0
// when the range holds no '#' and lies inside one block).  After the seed lookup nearly every range is one or a few rows
// inside ONE block: that case is a single sweep over the four 32-row words (letter mask computed once, '#' ranks only
// when a '#' is present).  nblk = Occ blocks fetched (1 or 2).
// [a, b) inside ONE block (already in registers): a single sweep over the four 32-row words
__device__ __forceinline__ void lf_same_block(const Blk& B, uint32_t a, uint32_t b, int c, uint32_t x0, uint32_t x1,
                                              uint32_t& rca, uint32_t& rcb, uint32_t& rda, uint32_t& rdb) {
    const int offA = (int)(a & 127u), offB = (int)(b & 127u);
    uint32_t r = (c & 2) ? ((c & 1) ? B.w[12] : B.w[8]) : ((c & 1) ? B.w[4] : B.w[0]);      // selects, not branches: lanes hold different letters
    uint32_t in = 0, hin = 0;
    #pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint32_t lmA = low_mask(offA - 32 * j), dm = low_mask(offB - 32 * j) & ~lmA;
        const uint32_t M = (B.w[4 * j + 1] ^ x0) & (B.w[4 * j + 2] ^ x1) & ~B.w[4 * j + 3];
        r += __popc(M & lmA); in += __popc(M & dm); hin |= B.w[4 * j + 3] & dm;
    }
    rca = r; rcb = r + in; rda = rdb = 0;
    if (hin) {
        uint32_t ka = 0, kin = 0;
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t lmA = low_mask(offA - 32 * j), dm = low_mask(offB - 32 * j) & ~lmA;
            ka += __popc(B.w[4 * j + 3] & lmA); kin += __popc(B.w[4 * j + 3] & dm);
        }
        rda = (a & ~127u) - letters_before(B) + ka; rdb = rda + kin;
    }
}
// B = block a >> 7, already in registers (the item kernel issues the fetch one stage before it consumes it)
__device__ __forceinline__ void lf_range_loaded(Blk& B, const uint4* __restrict__ occ, uint32_t a, uint32_t b, int c, uint32_t x0, uint32_t x1,
                                                uint32_t& rca, uint32_t& rcb, uint32_t& rda, uint32_t& rdb, uint32_t& nblk) {
    const uint32_t blkA = a >> 7, blkB = b >> 7;
    if (blkA == blkB) { lf_same_block(B, a, b, c, x0, x1, rca, rcb, rda, rdb); nblk = 1; }
    else {
        uint32_t ka, kb;
        rank_in_blk(B, a & 127u, c, x0, x1, rca, ka);
        rda = blkA * OCC_ROWS - letters_before(B) + ka;
        load_blk(occ, blkB, B);
        rank_in_blk(B, b & 127u, c, x0, x1, rcb, kb);
        rdb = blkB * OCC_ROWS - letters_before(B) + kb;
        nblk = 2;
    }
}
__device__ __forceinline__ void lf_range(const uint4* __restrict__ occ, uint32_t a, uint32_t b, int c, uint32_t x0, uint32_t x1,
                                         uint32_t& rca, uint32_t& rcb, uint32_t& rda, uint32_t& rdb, uint32_t& nblk) {
    Blk B;
    load_blk(occ, a >> 7, B);
    lf_range_loaded(B, occ, a, b, c, x0, x1, rca, rcb, rda, rdb, nblk);
}

// ---- the one-thread-per-pattern kernel on the wide layout: a whole 96-byte block in registers
struct BlkW { uint32_t w[24]; };     // w[4j..4j+3] = planes 0..2 + '#' mask of rows 32j..32j+31; w[16 + c] = count of letter c before the block
__device__ __forceinline__ void load_blk_w(const uint4* __restrict__ occ, uint32_t blk, BlkW& b) {
    const uint4* p = occ + (size_t)blk * OCCW_U4;
    ld256(p, b.w); ld256(p + 2, b.w + 8); ld256(p + 4, b.w + 16);
}
__device__ __forceinline__ uint32_t count_of_w(const BlkW& b, int c) {      // selects, not an indexed register array
    const uint32_t lo = (c & 2) ? ((c & 1) ? b.w[19] : b.w[18]) : ((c & 1) ? b.w[17] : b.w[16]);
    const uint32_t hi = (c & 2) ? ((c & 1) ? b.w[23] : b.w[22]) : ((c & 1) ? b.w[21] : b.w[20]);
    return (c & 4) ? hi : lo;
}
__device__ __forceinline__ uint32_t letters_before_w(const BlkW& b) {
    return b.w[16] + b.w[17] + b.w[18] + b.w[19] + b.w[20] + b.w[21] + b.w[22] + b.w[23];
}
// rank of letter c and number of '#' in rows [0, off) of the block
__device__ __forceinline__ void rank_in_blk_w(const BlkW& b, uint32_t off, int c, uint32_t& rc, uint32_t& nmk) {
    const uint32_t x0 = (c & 1) ? 0u : 0xffffffffu, x1 = (c & 2) ? 0u : 0xffffffffu, x2 = (c & 4) ? 0u : 0xffffffffu;
    uint32_t r = count_of_w(b, c), k = 0;
    #pragma unroll
    for (int j = 0; j < 4; ++j) {
        const uint32_t lm = low_mask((int)off - 32 * j);
        r += __popc((b.w[4 * j] ^ x0) & (b.w[4 * j + 1] ^ x1) & (b.w[4 * j + 2] ^ x2) & ~b.w[4 * j + 3] & lm);
        k += __popc(b.w[4 * j + 3] & lm);
    }
    rc = r; nmk = k;
}
// LF of the row range [a, b) by letter c on the wide layout (same contract as lf_range)
__device__ __forceinline__ void lf_range_w(const uint4* __restrict__ occ, uint32_t a, uint32_t b, int c,
                                           uint32_t& rca, uint32_t& rcb, uint32_t& rda, uint32_t& rdb, uint32_t& nblk) {
    BlkW B;
    const uint32_t blkA = a >> 7, blkB = b >> 7;
    load_blk_w(occ, blkA, B);
    uint32_t ka, kb;
    rank_in_blk_w(B, a & 127u, c, rca, ka);
    rda = blkA * OCC_ROWS - letters_before_w(B) + ka;
    nblk = 1;
    if (blkB != blkA) { load_blk_w(occ, blkB, B); nblk = 2; }
    rank_in_blk_w(B, b & 127u, c, rcb, kb);
    rdb = blkB * OCC_ROWS - letters_before_w(B) + kb;
}
// n <= 16 bytes at p -> 4-bit codes (character i at bits 4i..4i+3); false when a byte is not a letter of the index alphabet
__device__ __forceinline__ bool fetch_codes_w(const uint8_t* __restrict__ p, uint32_t n, const uint8_t* s_lut, uint64_t& codes) {
    uint32_t w[4];
    if (((uint32_t)(uintptr_t)p & 15u) == 0) { const uint4 a = __ldg((const uint4*)p); w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; }
    else {
        #pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint32_t v = 0;
            #pragma unroll
            for (int t = 0; t < 4; ++t) if (4u * j + t < n) v |= (uint32_t)__ldg(p + 4 * j + t) << (8 * t);
            w[j] = v;
        }
    }
    uint64_t cw = 0; uint32_t bad = 0;
    #pragma unroll
    for (int i = 0; i < 16; ++i) {
        const uint32_t c = s_lut[(w[i >> 2] >> (8 * (i & 3))) & 255u];
        cw |= (uint64_t)(c & 7u) << (4 * i);
        bad |= (c > 7u ? 1u : 0u) << i;
    }
    codes = n >= 16 ? cw : cw & ((1ull << (4 * n)) - 1);
    return (bad & (n >= 16 ? 0xffffu : ((1u << n) - 1u))) == 0;
}
#endif

// ---------------------------------------------------------------------------------------------------
// host-side index image (everything read from the reference's files)
// ---------------------------------------------------------------------------------------------------
struct HostIndex {
    uint32_t n = 0, n_text = 0, sigma = 0, n_seg = 0;
    uint8_t alphabet[8] = {0};
    std::vector<uint8_t> ebwt;            // piles concatenated (raw bytes)
    uint32_t pile_start[9] = {0};         // global row of the start of pile p, [sigma] = n
    std::vector<uint32_t> eof_flat;       // nText
    uint32_t table_occ[64] = {0};         // sigma x sigma
    std::vector<uint8_t> seg_first_bit;   // nText 0/1
};
void load_host_index(const std::string& base, bool verify, HostIndex& h);      // host_io.cpp

// sdsl formats (host_io.cpp)
std::vector<uint8_t> rrr_decode(const std::vector<uint8_t>& file);             // -> 0/1 per bit
std::vector<uint8_t> rrr_encode(const uint8_t* bits01, size_t nbits);          // sdsl rrr_vector<63> bytes
std::vector<uint8_t> bitvector_encode(const uint8_t* bits01, size_t nbits);
std::vector<uint8_t> bitvector_decode(const std::vector<uint8_t>& file);
std::vector<uint8_t> read_file(const std::string& path);
void write_file(const std::string& path, const void* data, size_t len);
// EDS text -> collection (eds_to_fasta.cpp:55-140)
void parse_eds_text(const std::vector<uint8_t>& text, std::vector<uint8_t>& letters, std::vector<uint64_t>& str_off,
                    std::vector<uint8_t>& seg_first);

// ---------------------------------------------------------------------------------------------------
// device index + search engine
// ---------------------------------------------------------------------------------------------------
struct SeedLevel { DevBuf<uint32_t> off, cnt; DevBuf<uint2> iv; DevBuf<uint4> first; size_t n_iv = 0; };     // iv[off[key] .. off[key+1]) = frontier (lo,hi) of key; cnt[key] = rows in it (levels < k)
struct SeedView { const uint32_t* off[17]; const uint2* iv[17]; const uint32_t* cnt[17];
                  const uint4* first; uint32_t first_level; };   // deepest level: {lo, hi of the first interval, offset, count} per key -- one gather instead of two
// plist entries of patterns no longer than the seed depth reference the seed table itself (no copy): y = PLIST_SEED | level << 27 | n
constexpr uint32_t PLIST_SEED = 0x80000000u, PLIST_N_MASK = 0x07ffffffu;
// ... and a final frontier of ONE interval travels inside its plist entry: x = first row, y = PLIST_INLINE; the interval's
// length is the pattern's count (items.cu; read by the locate kernels)
constexpr uint32_t PLIST_INLINE = 0x40000000u;

// Patterns as 2-bit codes (items.cu).  Pattern q of m characters occupies ceil(m/32) 64-bit words starting at word
//   q * stride                            (poff == nullptr: every pattern has fixed_len characters, stride = ceil(fixed_len/32))
//   ((poff[q] - poff[0]) >> 5) + q        (variable lengths: a closed form that never overlaps, no scan needed)
// word w holds characters [max(0, m - 32(w+1)), m - 32w), the last of them in bits 62..63, code = rank of the letter in
// the index alphabet.  invalid: bit q & 31 of word q >> 5 is set when pattern q holds a byte outside the alphabet (such a
// pattern has 0 occurrences); nullptr = all valid.
struct PackedPats {
    const uint64_t* words;
    const uint64_t* poff;
    const uint32_t* invalid;
    uint32_t fixed_len, stride;
};

struct StatsDev { unsigned long long I, S, E, L, O, lf_intervals, occ_blocks, overflow; };

struct Workspace;   // engine.cu
struct CsvScratch;  // csv.cu
void free_csv_scratch(CsvScratch* c);
struct SearchOut {
    DevBuf<uint32_t> counts; DevBuf<edsbwt_occ> occ; uint32_t npat = 0; uint64_t nocc = 0;
    DevBuf<uint32_t> alive; bool want_alive = false;    // alive[q] = longest suffix of pattern q whose frontier is non-empty
};
// host-buffer searches (edsbwt_search / edsbwt_search_into): chunks of the batch are copied in on one stream while the
// previous chunk is searched on the compute stream and the chunk before that is copied out on a third
struct Pipe {
    cudaStream_t in = nullptr, outs = nullptr;
    static constexpr int NRES = 2;      // result slots (4 were measured: no gain, profiles/r02_e2e_chunks.txt -- the copy-out is slow because it shares the memory system with the search kernel, not because the search waits for a slot)
    cudaEvent_t ev_in[2] = {nullptr, nullptr}, ev_out[NRES] = {nullptr, nullptr};
    DevBuf<uint8_t> bytes[2]; DevBuf<uint64_t> off[2];
    DevBuf<uint64_t> words[2]; DevBuf<uint32_t> invalid[2];      // packed batches (edsbwt_search_into_packed)
    SearchOut res[NRES];
    ~Pipe() { for (int i = 0; i < 2; ++i) if (ev_in[i]) cudaEventDestroy(ev_in[i]);
              for (int i = 0; i < NRES; ++i) if (ev_out[i]) cudaEventDestroy(ev_out[i]);
              if (in) cudaStreamDestroy(in); if (outs) cudaStreamDestroy(outs); }
};

// ---------------------------------------------------------------------------------------------------
// move table (MOVE_EDSBWTSearch variant; csrc/move.cu).  Input intervals = runs of the eBWT where every end-marker is a
// run of its own (da_to_everything.cpp:440-442), split further until a-balanced.  One 16-byte entry per interval:
//   .x p   first row of the interval          .y q   LF(p)
//   .z idx input interval containing q        .w c   symbol of the run: 0 = '#', 1 + letter code otherwise
// T[r] = {n, n, r, 0xff} is a sentinel.  L' (the run symbols) additionally lives in an Occ-block structure over RUN
// indices (rank of a letter / of '#' among the run heads) with a select array per letter.
// ---------------------------------------------------------------------------------------------------
struct MoveView {
    const uint4* T; uint32_t r;
    const uint4* occL;             // Occ blocks over L'
    const uint32_t* sel;           // sel[selbase[c] + k] = k-th run (0-based) whose symbol is letter c
    uint32_t selbase[5];           // [4] = number of letter runs
    const uint32_t* coarse;        // coarse[b] = number of intervals starting before row 128*b  (b = 0 .. n/128 + 1)
    const uint32_t* row_run0;      // interval containing pile-0 row i (M_LF_Dollar_Input_interval, MOVE_EDSBWTSearch.cpp:200-205)
};
struct MoveTable {
    DevBuf<uint4> T, occL; DevBuf<uint32_t> sel, coarse, row_run0;
    uint32_t r = 0, r_runs = 0, a = 8, rounds = 0;
    MoveView view{};
    uint64_t bytes() const { return T.bytes() + occL.bytes() + sel.bytes() + coarse.bytes() + row_run0.bytes(); }
};

// device buffers of finished searches are recycled: edsbwt_result_free hands them back instead of cudaFree (which
// synchronises the device and costs more than the locate kernel at cfg4 sizes)
struct ResultPool {
    std::mutex mu; std::vector<SearchOut*> free_list;
    SearchOut* get() { std::lock_guard<std::mutex> l(mu); if (free_list.empty()) return new SearchOut; SearchOut* r = free_list.back(); free_list.pop_back(); return r; }
    void put(SearchOut* r) { { std::lock_guard<std::mutex> l(mu); if (free_list.size() < 16) { free_list.push_back(r); return; } } delete r; }
    ~ResultPool() { for (SearchOut* r : free_list) delete r; }
};

struct Index {
    int device = 0;
    std::string base;     // <indexBase> of the reference's files (the persisted tables live next to them, cache.cu)
    int cache_mode = 0;   // edsbwt_open_opts.cache
    bool image_from_file = false, move_from_file = false, move_a_explicit = false;
    cudaStream_t stream = nullptr;
    std::mutex mu;        // one search (run_search / csv emission / move-table build) at a time on the handle's stream and workspace
    std::mutex call_mu;   // one host-buffer call at a time: the chunk pipeline (Pipe) is per handle; taken before mu, never after
    HostIndex host;       // ebwt bytes dropped after upload
    DevBuf<uint4> occ;
    DevBuf<uint32_t> eof_flat, str_seg, seg_first, seg_low;
    DevBuf<uint4> sa4;
    DevBuf<uint2> link_rec, str_info;
    DevBuf<const uint2*> seed_ivp;   // [17] device copy of the seed levels' interval arrays (locate of PLIST_SEED entries)
    DevBuf<uint8_t> lut;             // 256-byte byte -> code table on the device (0..3 letters, 4 end-marker, 255 other)
    IndexView view{};
    uint32_t seed_k = 0;
    int locate_mode = EDSBWT_LOCATE_SA;
    bool wide = false;               // 5..7 letters: 3-bit codes, 96-byte Occ blocks
    uint32_t key_radix = 4;          // seed key = sum of letter codes * radix^i (the number of letters when wide: 5^12 keys for {A,C,G,N,T})
    std::vector<SeedLevel*> seed;       // seed[L] for L = 1..seed_k (index 0 unused)
    uint64_t seed_intervals = 0;
    Workspace* ws = nullptr;
    Pipe* pipe = nullptr;
    cudaStream_t push = nullptr;        // peer pushes of finished results (edsbwt_result_push), created on first use
    MoveTable* move = nullptr;          // built at open (opts.build_move) or by the first EDSBWT_LF_MOVE search
    CsvScratch* csv = nullptr;          // scratch of the row-order sort and of the device csv formatter (csv.cu)
    DevBuf<uint32_t> eof_inv;           // string -> rank of its '#' row (inverse of eof_flat); built by the first ordered search
    int row_order = EDSBWT_ROWS_GROUPED;
    uint32_t move_a = 8;
    uint64_t l2_persist_bytes = 0;      // persisting-L2 carve-out granted for the Occ array
    std::shared_ptr<ResultPool> pool = std::make_shared<ResultPool>();
    uint32_t chunk_patterns = 1u << 21;     // patterns per pipeline chunk of the host-buffer searches
    uint64_t occ_limit = 0xfffffff0ull;     // occurrences one search may produce (32-bit offsets); EDSBWT_OCC_LIMIT lowers it (tests of the chunk-halving path)
    edsbwt_stats last{};
    int collect_counters = 0; uint32_t canonical_k = 12;
    int engine = EDSBWT_ENGINE_FUSED; uint32_t fused_cap = 13, fused_cap2 = FUSED_BIG_CAP;
    bool fused_cap_auto = true;         // no explicit fused_cap: a batch that overflows the default slots in > 0.1 % of its patterns switches later batches to 20
    uint64_t device_bytes() const;
    ~Index();
};

// engine.cu
struct Lut { uint8_t code[256]; };     // letter code 0..nletters-1, LUT_EOF = end-marker (4, the narrow kernels test for it; 8 for a wide alphabet), 255 = not in the alphabet
Lut make_lut(const HostIndex& h);
void build_occ_blocks(const uint8_t* d_bytes, uint32_t n, const Lut& lut, DevBuf<uint4>& occ, DevBuf<uint8_t>& cubtmp, cudaStream_t s,
                      uint32_t tot[4], unsigned int* bad_out);
void build_occ_blocks_wide(const uint8_t* d_bytes, uint32_t n, const Lut& lut, DevBuf<uint4>& occ, DevBuf<uint8_t>& cubtmp, cudaStream_t s,
                           uint32_t tot[8], unsigned int* bad_out);
void init_index_runtime(Index& ix, const edsbwt_open_opts& o);      // stream, workspace, engine options, byte -> code table
void build_device_index(Index& ix, const edsbwt_open_opts& o);
// cache.cu -- persisted tables: <base>.edsbwt (whole HBM image, opt-in) and <base>_MLF.edsbwt (move table)
uint64_t source_fingerprint(const std::string& base);
bool image_cache_load(Index& ix, const edsbwt_open_opts& o);
void image_cache_save(Index& ix);
bool move_cache_load(Index& ix, const std::string& path);
void move_cache_save(Index& ix, const std::string& path);
// what a call knows about its batch besides bytes + offsets
struct PatHint {
    uint32_t fixed_len = 0;              // > 0: every pattern has this many characters; pattern q starts at byte fixed_base + q * fixed_len of d_bytes (d_off may be null)
    uint64_t fixed_base = 0;
    const uint64_t* d_words = nullptr;   // the caller's own 2-bit words (PackedPats layout, stride ceil(fixed_len / 32)); d_bytes may then be null
    const uint32_t* d_invalid = nullptr; // with d_words: bitmap of patterns that hold a byte outside the alphabet (null: none)
    cudaStream_t after = nullptr;        // not null: run_search returns without waiting for its last kernels; `after` is made to wait for them instead
};
void run_search(Index& ix, const uint8_t* d_bytes, const uint64_t* d_off, uint32_t npat, uint64_t total_bytes,
                uint32_t max_len, bool locate, int lf_variant, SearchOut& out, uint32_t pat_base = 0, const PatHint* hint = nullptr);

// fused.cu -- the step-loop kernel: one thread per pattern (second pass over the patterns that outgrew the item kernel; A/B engine)
struct FusedOut {
    uint2*    plist;        // [npat] (first slot in staging, number of final intervals)
    uint32_t* counts;       // [npat] number of occurrences ("Found N")
    uint2*    staging;      // final (lo,hi) row intervals, contiguous per pattern
    uint32_t  staging_cap;
    uint32_t* cursor;       // slots used in staging
    uint32_t* ovf_count;    // number of patterns whose frontier did not fit
    uint32_t* ovf_ids;      // those patterns
    uint32_t* alive;        // optional [npat]: longest pattern suffix with a non-empty frontier (edsbwt_search_trace)
};
void launch_fused_search(const IndexView& ix, const SeedView& sv, uint32_t k, const uint8_t* d_bytes, const uint64_t* d_off,
                         const uint32_t* d_ids, uint32_t npat, const uint8_t* d_lut, uint32_t cap, const FusedOut& out, StatsDev* st, uint32_t kc, cudaStream_t s);

// items.cu -- the hot kernel: persistent warps, one item per lane per sweep, lanes refilled from a pool of patterns
void launch_pack_patterns(const uint8_t* d_bytes, const uint64_t* d_off, uint32_t fixed_len, uint64_t fixed_base, uint32_t stride, uint32_t npat,
                          const uint8_t* d_lut, uint64_t* words, uint64_t nwords, uint32_t* invalid, uint32_t* err, cudaStream_t s);
void launch_unpack_patterns(const uint64_t* words, uint32_t fixed_len, uint32_t stride, const uint32_t* ids, uint32_t nsub, uint64_t letters,
                            const uint32_t* invalid, uint8_t* bytes, cudaStream_t s);
void launch_items_search(const IndexView& ix, const SeedView& sv, uint32_t k, const PackedPats& pp, uint32_t npat, uint32_t cap, const FusedOut& out,
                         uint32_t* work, StatsDev* st, uint32_t kc, cudaStream_t s);

// move.cu
void build_move_table(Index& ix);
struct MoveOut {            // what the move search leaves for the common result path of run_search
    uint2* plist; uint32_t* counts; uint2* staging; uint32_t staging_cap; uint32_t* cursor; uint32_t* ovf_count; uint32_t* ovf_ids;
};
void launch_move_search(const IndexView& ix, const MoveView& mv, const SeedView& sv, uint32_t k, const uint8_t* d_bytes, const uint64_t* d_off,
                        const uint32_t* d_ids, uint32_t npat, const uint8_t* d_lut, uint32_t cap, const MoveOut& out, cudaStream_t s);
void launch_move_locate(const IndexView& ix, const MoveView& mv, const uint2* plist, const uint32_t* occ_ofs, const uint2* staging,
                        uint32_t npat, uint64_t nocc, uint32_t pat_base, edsbwt_occ* out, StatsDev* st, cudaStream_t s);
// csv.cu
uint64_t order_rows_reference(Index& ix, SearchOut& out, const uint32_t* d_occ_ofs, uint32_t npat, cudaStream_t s);   // -> kernels launched
uint64_t write_csv_device(Index& ix, const edsbwt_occ* d_occ, uint64_t nocc, uint32_t pat_add, FILE* f, cudaStream_t s, uint64_t* launches);   // -> bytes
// builder.cu
void build_index_files(const uint8_t* letters, const uint64_t* str_off, uint32_t nstr, const uint8_t* seg_first,
                       const std::string& out_base, int device);
// gather_bench.cu
struct GatherCfg { uint64_t buffer_bytes = 0, n_access = 0; int iters = 3, granule = 64, ilp = 4, ctas_per_sm = 0, dependent = 0;
                   uint64_t mix_bytes = 0; double mix_frac = 0; };
double gather_bench_cfg(int device, const GatherCfg& c, double* req_per_s);
double gather_bench(int device, uint64_t buffer_bytes, uint64_t n_access, int iters, int granule = 64);

}  // namespace edsbwt

struct edsbwt_index { edsbwt::Index ix; };
struct edsbwt_result { edsbwt::SearchOut* outp = nullptr; std::shared_ptr<edsbwt::ResultPool> pool; int device = 0; };
