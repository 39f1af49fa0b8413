// Device-resident EDS-BWT index + level-synchronous batched search engine (sm_100a).
//
// Data-parallel restatement of the reference's per-pattern loop:
//   frontier F   = all (pattern, [lo,hi]) row intervals of the whole batch, grouped by pattern, global 0-based rows
//   one step     = lf_step_kernel  : LF_c of every interval + '#' count          (updateIntervals, EDSBWTsearch.cpp:956-1157)
//                  enum_dollars     : strings whose start row lies in F           (dollars_in_interval, :1265-1281)
//                  sort + dol_runs  : maximal runs of reachable segments          (link/preceding_dollars_finder, :1212-1307)
//                  lf_step_kernel   : LF_c of the pile-0 ranges                   (updateIntervals on vectRangeDollarPile, :691)
//                  compact/merge    : drop empty, merge adjacent                  (MergeAndRemove, :911-953)
//                  combine          : F <- Dol' ++ F' per pattern                 (:696-700)
// The first k steps of every pattern come from a seed table (frontiers of all 4^L suffixes, L <= k) that is
// built with the very same step code run level-synchronously over all keys.
#include <cub/cub.cuh>

#include <algorithm>

#include "internal.cuh"

namespace edsbwt {

// ===================================================================================================
// small device helpers
// ===================================================================================================
__device__ __forceinline__ uint32_t upper_bound_u32(const uint32_t* __restrict__ a, uint32_t n, uint32_t v) {
    uint32_t lo = 0, hi = n;                       // first index with a[idx] > v
    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (__ldg(a + mid) <= v) lo = mid + 1; else hi = mid; }
    return lo;
}
__device__ __forceinline__ uint32_t lower_bound_u32(const uint32_t* __restrict__ a, uint32_t n, uint32_t v) {
    uint32_t lo = 0, hi = n;                       // first index with a[idx] >= v
    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (__ldg(a + mid) < v) lo = mid + 1; else hi = mid; }
    return lo;
}
__device__ __forceinline__ void warp_add(unsigned long long* dst, uint32_t v) {
    // warp-aggregated statistics counter
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(dst, (unsigned long long)v);
}

static inline unsigned grid_for(size_t n, unsigned block) { return (unsigned)((n + block - 1) / block); }
struct WidenU32 { __host__ __device__ __forceinline__ unsigned long long operator()(const uint32_t& v) const { return v; } };

// ===================================================================================================
// K1: Occ structure from the raw eBWT bytes  (replaces computeVectorOcc, EDSBWTsearch.cpp:1338-1442,
//     and the rrr rank supports of recoverInfo, :1586-1604)
// ===================================================================================================

__global__ void occ_pack_kernel(const uint8_t* __restrict__ bytes, uint32_t n, Lut lut, uint4* __restrict__ occ,
                                uint4* __restrict__ blkcnt, uint32_t ngroups, unsigned int* bad) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;          // one thread per 32-row group = one uint4 lane
    uint32_t p0 = 0, p1 = 0, mk = 0, c0 = 0, c1 = 0, c2 = 0, c3 = 0;
    if (g < ngroups) {
        uint32_t base = g * 32u;
        #pragma unroll 4
        for (uint32_t b = 0; b < 32; ++b) {
            uint32_t r = base + b;
            if (r < n) {
                uint8_t c = lut.code[bytes[r]];
                if (c == 255) { atomicExch(bad, r + 1); c = 4; }
                if (c == 4) mk |= 1u << b;
                else { p0 |= (uint32_t)(c & 1) << b; p1 |= (uint32_t)((c >> 1) & 1) << b; c0 += c == 0; c1 += c == 1; c2 += c == 2; c3 += c == 3; }
            }
        }
        occ[g] = make_uint4(0u, p0, p1, mk);
    }
    // sum the four groups of a block (consecutive lanes)
    c0 += __shfl_xor_sync(0xffffffffu, c0, 1); c0 += __shfl_xor_sync(0xffffffffu, c0, 2);
    c1 += __shfl_xor_sync(0xffffffffu, c1, 1); c1 += __shfl_xor_sync(0xffffffffu, c1, 2);
    c2 += __shfl_xor_sync(0xffffffffu, c2, 1); c2 += __shfl_xor_sync(0xffffffffu, c2, 2);
    c3 += __shfl_xor_sync(0xffffffffu, c3, 1); c3 += __shfl_xor_sync(0xffffffffu, c3, 2);
    if (g < ngroups && (g & 3) == 0) blkcnt[g >> 2] = make_uint4(c0, c1, c2, c3);
}
struct AddU4 { __device__ __forceinline__ uint4 operator()(const uint4& a, const uint4& b) const { return make_uint4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); } };
__global__ void occ_counts_kernel(uint4* __restrict__ occ, const uint4* __restrict__ prefix, uint32_t ngroups) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ngroups) return;
    uint4 p = prefix[g >> 2];
    uint32_t j = g & 3;
    occ[g].x = j == 0 ? p.x : j == 1 ? p.y : j == 2 ? p.z : p.w;
}

// bytes (alphabet characters) -> interleaved Occ blocks; tot[c] = occurrences of letter c, *bad = 1 + first row whose byte
// is outside the alphabet (0: none).  Used for the eBWT and for the run-head string L' of the move table.
void build_occ_blocks(const uint8_t* d_bytes, uint32_t n, const Lut& lut, DevBuf<uint4>& occ, DevBuf<uint8_t>& cubtmp, cudaStream_t s,
                      uint32_t tot[4], unsigned int* bad_out) {
    uint32_t nblk = n / OCC_ROWS + 1, ngroups = nblk * 4;
    occ.ensure(ngroups);
    DevBuf<uint4> blkcnt, prefix; blkcnt.ensure(nblk); prefix.ensure(nblk);
    DevBuf<unsigned int> bad; bad.ensure(1);
    CUDA_OK(cudaMemsetAsync(bad.p, 0, 4, s));
    occ_pack_kernel<<<grid_for(ngroups, 256), 256, 0, s>>>(d_bytes, n, lut, occ.p, blkcnt.p, ngroups, bad.p);
    size_t tmp = 0;
    CUDA_OK(cub::DeviceScan::ExclusiveScan(nullptr, tmp, blkcnt.p, prefix.p, AddU4(), make_uint4(0, 0, 0, 0), (int)nblk, s));
    cubtmp.ensure(tmp);
    CUDA_OK(cub::DeviceScan::ExclusiveScan(cubtmp.p, tmp, blkcnt.p, prefix.p, AddU4(), make_uint4(0, 0, 0, 0), (int)nblk, s));
    occ_counts_kernel<<<grid_for(ngroups, 256), 256, 0, s>>>(occ.p, prefix.p, ngroups);
    unsigned int hbad = 0;
    uint4 last_prefix, last_cnt;
    CUDA_OK(cudaMemcpyAsync(&hbad, bad.p, 4, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaMemcpyAsync(&last_prefix, prefix.p + (nblk - 1), 16, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaMemcpyAsync(&last_cnt, blkcnt.p + (nblk - 1), 16, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    tot[0] = last_prefix.x + last_cnt.x; tot[1] = last_prefix.y + last_cnt.y; tot[2] = last_prefix.z + last_cnt.z; tot[3] = last_prefix.w + last_cnt.w;
    if (bad_out) *bad_out = hbad;
}

// ---- wide alphabets: 96-byte blocks (internal.cuh), letter codes 0..7, end-marker code 8
__global__ void occw_pack_kernel(const uint8_t* __restrict__ bytes, uint32_t n, Lut lut, uint4* __restrict__ occ,
                                 uint4* __restrict__ cnt_lo, uint4* __restrict__ cnt_hi, uint32_t ngroups, unsigned int* bad) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;          // one thread per 32-row group
    uint32_t p0 = 0, p1 = 0, p2 = 0, mk = 0;
    uint32_t c[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (g < ngroups) {
        uint32_t base = g * 32u;
        for (uint32_t b = 0; b < 32; ++b) {
            uint32_t r = base + b;
            if (r < n) {
                uint8_t v = lut.code[bytes[r]];
                if (v == 255) { atomicExch(bad, r + 1); v = 8; }
                if (v == 8) mk |= 1u << b;
                else {
                    p0 |= (uint32_t)(v & 1) << b; p1 |= (uint32_t)((v >> 1) & 1) << b; p2 |= (uint32_t)((v >> 2) & 1) << b;
                    #pragma unroll
                    for (int t = 0; t < 8; ++t) c[t] += v == t;
                }
            }
        }
        occ[(size_t)(g >> 2) * OCCW_U4 + (g & 3)] = make_uint4(p0, p1, p2, mk);
    }
    #pragma unroll
    for (int t = 0; t < 8; ++t) { c[t] += __shfl_xor_sync(0xffffffffu, c[t], 1); c[t] += __shfl_xor_sync(0xffffffffu, c[t], 2); }
    if (g < ngroups && (g & 3) == 0) { cnt_lo[g >> 2] = make_uint4(c[0], c[1], c[2], c[3]); cnt_hi[g >> 2] = make_uint4(c[4], c[5], c[6], c[7]); }
}
__global__ void occw_counts_kernel(uint4* __restrict__ occ, const uint4* __restrict__ pre_lo, const uint4* __restrict__ pre_hi, uint32_t ngroups) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= ngroups) return;
    const uint32_t j = g & 3;
    if (j < 2) occ[(size_t)(g >> 2) * OCCW_U4 + 4 + j] = j ? pre_hi[g >> 2] : pre_lo[g >> 2];      // counts of letters 0..3 / 4..7
}
void build_occ_blocks_wide(const uint8_t* d_bytes, uint32_t n, const Lut& lut, DevBuf<uint4>& occ, DevBuf<uint8_t>& cubtmp, cudaStream_t s,
                           uint32_t tot[8], unsigned int* bad_out) {
    uint32_t nblk = n / OCC_ROWS + 1, ngroups = nblk * 4;
    occ.ensure((size_t)nblk * OCCW_U4);
    DevBuf<uint4> cnt, pre; cnt.ensure((size_t)nblk * 2); pre.ensure((size_t)nblk * 2);
    DevBuf<unsigned int> bad; bad.ensure(1);
    CUDA_OK(cudaMemsetAsync(bad.p, 0, 4, s));
    occw_pack_kernel<<<grid_for(ngroups, 256), 256, 0, s>>>(d_bytes, n, lut, occ.p, cnt.p, cnt.p + nblk, ngroups, bad.p);
    size_t tmp = 0;
    CUDA_OK(cub::DeviceScan::ExclusiveScan(nullptr, tmp, cnt.p, pre.p, AddU4(), make_uint4(0, 0, 0, 0), (int)nblk, s));
    cubtmp.ensure(tmp);
    CUDA_OK(cub::DeviceScan::ExclusiveScan(cubtmp.p, tmp, cnt.p, pre.p, AddU4(), make_uint4(0, 0, 0, 0), (int)nblk, s));
    CUDA_OK(cub::DeviceScan::ExclusiveScan(cubtmp.p, tmp, cnt.p + nblk, pre.p + nblk, AddU4(), make_uint4(0, 0, 0, 0), (int)nblk, s));
    occw_counts_kernel<<<grid_for(ngroups, 256), 256, 0, s>>>(occ.p, pre.p, pre.p + nblk, ngroups);
    unsigned int hbad = 0;
    uint4 lp[2], lc[2];
    CUDA_OK(cudaMemcpyAsync(&hbad, bad.p, 4, cudaMemcpyDeviceToHost, s));
    for (int h = 0; h < 2; ++h) {
        CUDA_OK(cudaMemcpyAsync(&lp[h], pre.p + (size_t)h * nblk + (nblk - 1), 16, cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaMemcpyAsync(&lc[h], cnt.p + (size_t)h * nblk + (nblk - 1), 16, cudaMemcpyDeviceToHost, s));
    }
    CUDA_OK(cudaStreamSynchronize(s));
    for (int h = 0; h < 2; ++h) { tot[4 * h] = lp[h].x + lc[h].x; tot[4 * h + 1] = lp[h].y + lc[h].y; tot[4 * h + 2] = lp[h].z + lc[h].z; tot[4 * h + 3] = lp[h].w + lc[h].w; }
    if (bad_out) *bad_out = hbad;
}

// ===================================================================================================
// SA table: row -> (string, offset), by walking every string backwards from its end-marker row
// (the inverse of the reference's locate rounds, findMultipleDollarsBackward :349-603)
// ===================================================================================================
__global__ void sa_walk_kernel(const __grid_constant__ IndexView ix, uint2* __restrict__ sa, uint32_t* __restrict__ slen) {
    uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= ix.n_text) return;
    uint32_t row = w, d = 0;                     // pile-0 row w is the suffix "$_w"
    sa[w] = make_uint2(w, 0);                    // end-marker rows: (string, distance 0); never read by locate
    while (lf_row_any(ix, row)) {
        ++d;
        sa[row] = make_uint2(w, d);              // d = distance to the end of the string
    }
    slen[w] = d;
}
// final table: one 16-byte record per row = everything a csv row needs, (string, offset in string, segment, index in segment)
__global__ void sa_offset_kernel(uint32_t n_text, uint32_t n, const uint2* __restrict__ sa, const uint32_t* __restrict__ slen,
                                 const uint2* __restrict__ str_info, uint4* __restrict__ sa4) {
    uint32_t row = n_text + blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n) return;
    const uint2 v = sa[row];
    const uint2 si = __ldg(str_info + v.x);
    sa4[row] = make_uint4(v.x, slen[v.x] - v.y, si.x, si.y);
}
// link_rec[r] for the r-th end-marker row of the eBWT (EOF_ID order): everything link() derives from it
// (preceding_dollars_finder :1287-1307 + the empty-string closure :1255); str_info[w] = (segment, index in segment)
__global__ void link_rec_kernel(IndexView ix, uint2* __restrict__ link_rec, uint2* __restrict__ str_info) {
    uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= ix.n_text) return;
    uint32_t w = ix.eof_flat[r];
    uint2 rec = make_uint2(1u, 0u);              // x > y: no predecessor
    if (w > ix.fsi) {
        uint32_t sg = ix.str_seg[w] - 1;         // previous segment
        rec = make_uint2(ix.seg_first[ix.seg_low[sg]], ix.seg_first[sg + 1] - 1);
    }
    link_rec[r] = rec;
    uint32_t sg = ix.str_seg[r];                 // r doubles as a string index for str_info
    str_info[r] = make_uint2(sg, r - ix.seg_first[sg]);
}
__global__ void zip_kernel(const uint32_t* __restrict__ lo, const uint32_t* __restrict__ hi, uint32_t n, uint2* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = make_uint2(lo[i], hi[i]);
}

// ===================================================================================================
// patterns
// ===================================================================================================
struct PatView {
    const uint64_t* pack;    // [npat * words] letter codes, cb bits each (2; 4 for a wide alphabet): char i of pattern q at bits cb*(i % per) of pack[q*words + i/per], per = 64/cb
    const uint32_t* len;     // 0 for patterns that cannot match (bad character / empty)
    uint32_t words, npat, k; // k = seed depth used for this batch
    uint32_t cb = 2, kr = 4; // kr = radix of a seed key: key = sum of code(char i of the suffix) * kr^i (4; the number of letters for a wide alphabet)
};
__global__ void encode_patterns_kernel(const uint8_t* __restrict__ bytes, const uint64_t* __restrict__ off, const uint32_t* __restrict__ ids,
                                       uint32_t npat, Lut lut, uint32_t words, uint64_t* __restrict__ pack, uint32_t* __restrict__ len,
                                       uint32_t nletters, uint32_t cb) {
    uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= npat) return;
    uint32_t g = ids ? ids[q] : q;               // local pattern q of this pass = pattern g of the batch
    uint64_t a = off[g], b = off[g + 1];
    uint32_t m = (uint32_t)(b - a);
    bool ok = m > 0;
    uint64_t wv = 0;
    const uint32_t per = 64 / cb;
    for (uint32_t i = 0; i < m; ++i) {
        uint8_t c = lut.code[bytes[a + i]];
        if (c >= nletters) { ok = false; c = 0; }
        wv |= (uint64_t)c << (cb * (i % per));
        if ((i % per) == per - 1 || i + 1 == m) { pack[(size_t)q * words + i / per] = wv; wv = 0; }
    }
    len[q] = ok ? m : 0;
}
__device__ __forceinline__ uint32_t pat_char(const PatView& pv, uint32_t q, uint32_t idx) {
    if (pv.cb == 2) return (uint32_t)(__ldg(pv.pack + (size_t)q * pv.words + (idx >> 5)) >> (2 * (idx & 31))) & 3u;
    return (uint32_t)(__ldg(pv.pack + (size_t)q * pv.words + (idx >> 4)) >> (4 * (idx & 15))) & 15u;
}

// ===================================================================================================
// seed table
// ===================================================================================================

// number of seed intervals of each pattern + where they start inside the level
__global__ void seed_count_kernel(PatView pv, SeedView sv, uint32_t* __restrict__ cnt, uint32_t* __restrict__ src) {
    uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= pv.npat) return;
    uint32_t m = pv.len[q];
    if (m == 0) { cnt[q] = 0; src[q] = 0; return; }
    uint32_t L = m < pv.k ? m : pv.k;
    uint32_t key = 0;
    for (uint32_t i = 0, mul = 1; i < L; ++i, mul *= pv.kr) key += pat_char(pv, q, m - L + i) * mul;
    uint32_t a = __ldg(sv.off[L] + key), b = __ldg(sv.off[L] + key + 1);
    cnt[q] = b - a; src[q] = a;
}
__global__ void seed_scatter_kernel(PatView pv, SeedView sv, const uint32_t* __restrict__ base, const uint32_t* __restrict__ src,
                                    uint32_t total, uint32_t* __restrict__ fpat, uint32_t* __restrict__ flo, uint32_t* __restrict__ fhi) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= total) return;
    uint32_t q = upper_bound_u32(base, pv.npat + 1, t) - 1;      // base = exclusive scan of cnt (npat+1 entries)
    uint32_t m = pv.len[q];
    uint32_t L = m < pv.k ? m : pv.k;
    uint32_t j = src[q] + (t - base[q]);
    uint2 v = __ldg(sv.iv[L] + j);
    fpat[t] = q; flo[t] = v.x; fhi[t] = v.y;
}
// level L -> the nchild (4; the number of letters for a wide alphabet) children of every key, grouped by child key (nchild*key + c)
__global__ void seed_expand_kernel(const uint32_t* __restrict__ pat, const uint2* __restrict__ iv,
                                   const uint32_t* __restrict__ off, uint32_t nF, uint32_t* __restrict__ opat, uint32_t* __restrict__ olo, uint32_t* __restrict__ ohi,
                                   uint32_t nchild) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nF) return;
    uint32_t key = pat[i], a = off[key], cnt = off[key + 1] - a, r = i - a, l = iv[i].x, h = iv[i].y;
    for (uint32_t c = 0; c < nchild; ++c) { uint32_t d = nchild * a + c * cnt + r; opat[d] = nchild * key + c; olo[d] = l; ohi[d] = h; }
}
__global__ void level_keys_kernel(const uint32_t* __restrict__ off, uint32_t nkeys, uint32_t nF, uint32_t* __restrict__ keyof) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nF) return;
    keyof[i] = upper_bound_u32(off, nkeys + 1, i) - 1;
}
__global__ void key_offsets_kernel(const uint32_t* __restrict__ pat, uint32_t nF, uint32_t nkeys, uint32_t* __restrict__ off) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > nkeys) return;
    off[k] = k == nkeys ? nF : lower_bound_u32(pat, nF, k);
}

// ===================================================================================================
// the LF step (hot kernel)
// ===================================================================================================
// One 4-lane group per interval: each lane issues ONE 128-bit load per Occ block (64 B block = 4 coalesced
// lanes), ranks come from popcounts + two xor-shuffles.  When both interval ends fall into the same block
// (the common case after the seed phase) the block is fetched once.
//   KEYMODE   : seed-table construction, "pattern" id is the child key, its character is key & 3
//   otherwise : the character comes from the packed pattern; finished patterns pass through unchanged
template <bool KEYMODE, bool WIDE = false>
__global__ void __launch_bounds__(256)
lf_step_kernel(const __grid_constant__ IndexView ix, PatView pv, uint32_t step, const uint32_t* __restrict__ pat, const uint32_t* __restrict__ lo,
               const uint32_t* __restrict__ hi, uint32_t nF, uint32_t* __restrict__ olo, uint32_t* __restrict__ ohi,
               uint32_t* __restrict__ dcnt, uint32_t* __restrict__ dstart, uint32_t* __restrict__ keep,
               StatsDev* __restrict__ st, uint32_t kc, int is_dol) {
    uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t i = gid >> 2; int lane4 = gid & 3;
    bool act = i < nF;
    uint32_t ii = act ? i : 0;
    uint32_t q = __ldg(pat + ii), a = __ldg(lo + ii), b = __ldg(hi + ii) + 1;      // [a, b) half-open
    int c = 0; bool pass = false; uint32_t L = 0, m = 0;
    if (KEYMODE) c = WIDE ? (int)(q % pv.kr) : (int)(q & 3);
    else {
        m = __ldg(pv.len + q);
        L = (m < pv.k ? m : pv.k) + step;
        if (L >= m) pass = true; else c = (int)pat_char(pv, q, m - L - 1);
    }
    bool doload = act && !pass;
    uint32_t blk_a = a >> 7, blk_b = b >> 7;
    uint32_t rca, rda, rcb, rdb, Cc;
    if (WIDE) {
        // the 96-byte layout: planes + mask in entry lane4, the counts of letters 2*lane4, 2*lane4+1 in the 8 bytes at 64 + 8*lane4
        uint4 va = make_uint4(0, 0, 0, 0), vb; uint2 ca = make_uint2(0, 0), cb2;
        if (doload) { const uint4* p = ix.occ + (size_t)blk_a * OCCW_U4; va = __ldg(p + lane4); ca = __ldg(reinterpret_cast<const uint2*>(p + 4) + lane4); }
        vb = va; cb2 = ca;
        if (doload && blk_b != blk_a) { const uint4* p = ix.occ + (size_t)blk_b * OCCW_U4; vb = __ldg(p + lane4); cb2 = __ldg(reinterpret_cast<const uint2*>(p + 4) + lane4); }
        rank4w_from(va, ca, a, c, lane4, rca, rda);
        rank4w_from(vb, cb2, b, c, lane4, rcb, rdb);
        Cc = ix.Clw[c];
    } else {
        uint4 va = make_uint4(0, 0, 0, 0), vb;
        if (doload) va = occ_load(ix.occ, blk_a, lane4);
        vb = va;
        if (doload && blk_b != blk_a) vb = occ_load(ix.occ, blk_b, lane4);
        rank4_from(va, a, c, lane4, rca, rda);
        rank4_from(vb, b, c, lane4, rcb, rdb);
        Cc = ix.Cl[c];
    }
    uint32_t nlo, nhi, nd = 0;
    bool kept;
    if (pass) { nlo = a; nhi = b - 1; kept = true; }
    else { nlo = Cc + rca; nhi = Cc + rcb - 1; kept = rcb > rca; nd = rdb - rda; }
    if (act && lane4 == 0) {
        olo[i] = nlo; ohi[i] = nhi; keep[i] = kept ? 1u : 0u;
        if (dcnt) { dcnt[i] = nd; dstart[i] = rda; }
    }
    if (st) {
        bool mine = act && lane4 == 0 && !pass;
        bool beyond = KEYMODE ? false : (L >= kc);
        warp_add(&st->lf_intervals, mine ? 1u : 0u);
        warp_add(&st->S, (mine && beyond) ? 1u : 0u);
        warp_add(&st->E, (mine && beyond) ? nd : 0u);
        if (!KEYMODE && !is_dol) {
            uint32_t mk = m < kc ? m : kc;           // the canonical seed entry is the frontier at suffix length min(m, kc)
            uint32_t Lv = (m < pv.k ? m : pv.k) + step;
            warp_add(&st->I, (act && lane4 == 0 && m > 0 && Lv == mk) ? 1u : 0u);
        }
    }
}
// I-counter for the frontier left after the last step
__global__ void count_I_kernel(PatView pv, uint32_t step, const uint32_t* __restrict__ pat, uint32_t nF, StatsDev* st, uint32_t kc) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    bool hit = false;
    if (i < nF) { uint32_t m = pv.len[pat[i]]; hit = m > 0 && ((m < pv.k ? m : pv.k) + step == (m < kc ? m : kc)); }
    warp_add(&st->I, hit ? 1u : 0u);
}

// ---- compaction + MergeAndRemove -----------------------------------------------------------------
__global__ void compact_kernel(const uint32_t* __restrict__ pat, const uint32_t* __restrict__ lo, const uint32_t* __restrict__ hi,
                               const uint32_t* __restrict__ keep, const uint32_t* __restrict__ pos, uint32_t n,
                               uint32_t* __restrict__ opat, uint32_t* __restrict__ olo, uint32_t* __restrict__ ohi) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !keep[i]) return;
    uint32_t d = pos[i];
    opat[d] = pat[i]; olo[d] = lo[i]; ohi[d] = hi[i];
}
// head[i] = 1 unless interval i continues interval i-1 (same pattern, still active, lo == prev.hi + 1)
template <bool KEYMODE>
__global__ void head_flags_kernel(PatView pv, uint32_t step, const uint32_t* __restrict__ pat, const uint32_t* __restrict__ lo,
                                  const uint32_t* __restrict__ hi, const uint32_t* __restrict__ n_ptr, uint32_t nmax, uint32_t* __restrict__ head) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nmax) return;
    uint32_t n = *n_ptr;
    if (i >= n) { head[i] = 0; return; }
    uint32_t h = 1;
    if (i > 0 && pat[i - 1] == pat[i] && lo[i] == hi[i - 1] + 1) {
        bool pass = false;
        if (!KEYMODE) { uint32_t m = pv.len[pat[i]]; pass = (m < pv.k ? m : pv.k) + step >= m; }
        if (!pass) h = 0;
    }
    head[i] = h;
}
__global__ void merge_kernel(const uint32_t* __restrict__ pat, const uint32_t* __restrict__ lo, const uint32_t* __restrict__ hi,
                             const uint32_t* __restrict__ head, const uint32_t* __restrict__ pos, const uint32_t* __restrict__ n_ptr, uint32_t nmax,
                             uint32_t* __restrict__ opat, uint32_t* __restrict__ olo, uint32_t* __restrict__ ohi) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t n = *n_ptr;
    if (i >= nmax || i >= n) return;
    uint32_t run = pos[i] + head[i] - 1;                 // pos = exclusive scan of head
    if (head[i]) { opat[run] = pat[i]; olo[run] = lo[i]; }
    if (i + 1 == n || head[i + 1]) ohi[run] = hi[i];
}

// ---- link -------------------------------------------------------------------------------------------
__global__ void enum_dollars_kernel(IndexView ix, const uint32_t* __restrict__ pat, const uint32_t* __restrict__ dofs,
                                    const uint32_t* __restrict__ dstart, uint32_t nF, uint32_t D, uint64_t* __restrict__ keys,
                                    unsigned int* __restrict__ nvalid) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    bool valid = false;
    if (t < D) {
        uint32_t i = upper_bound_u32(dofs, nF + 1, t) - 1;
        uint32_t w = __ldg(ix.eof_flat + dstart[i] + (t - dofs[i]));
        uint64_t key = ~0ull;
        if (w > ix.fsi) { key = ((uint64_t)pat[i] << 32) | (uint64_t)(__ldg(ix.str_seg + w) - 1); valid = true; }
        keys[t] = key;
    }
    unsigned bal = __ballot_sync(0xffffffffu, valid);
    if ((threadIdx.x & 31) == 0 && bal) atomicAdd(nvalid, (unsigned)__popc(bal));
}
// sorted (pattern, segment) keys -> heads of maximal runs of reachable segments (closure through segments
// holding the empty string is the precomputed seg_low[])
__global__ void dol_heads_kernel(IndexView ix, const uint64_t* __restrict__ keys, uint32_t Dv, uint32_t* __restrict__ head) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= Dv) return;
    uint64_t k = keys[t];
    uint32_t h = 1;
    if (t > 0) {
        uint64_t p = keys[t - 1];
        if ((uint32_t)(p >> 32) == (uint32_t)(k >> 32) && __ldg(ix.seg_low + (uint32_t)k) <= (uint32_t)p + 1) h = 0;
    }
    head[t] = h;
}
__global__ void dol_build_kernel(IndexView ix, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ head, const uint32_t* __restrict__ pos,
                                 uint32_t Dv, uint32_t* __restrict__ dpat, uint32_t* __restrict__ dlo, uint32_t* __restrict__ dhi) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= Dv) return;
    uint64_t k = keys[t];
    uint32_t s = (uint32_t)k, run = pos[t] + head[t] - 1;
    if (head[t]) { dpat[run] = (uint32_t)(k >> 32); dlo[run] = __ldg(ix.seg_first + __ldg(ix.seg_low + s)); }
    if (t + 1 == Dv || head[t + 1]) dhi[run] = __ldg(ix.seg_first + s + 1) - 1;
}
// F <- Dol' ++ F' per pattern (both inputs sorted by pattern)
__global__ void combine_kernel(const uint32_t* __restrict__ fpat, const uint32_t* __restrict__ flo, const uint32_t* __restrict__ fhi, uint32_t nf,
                               const uint32_t* __restrict__ dpat, const uint32_t* __restrict__ dlo, const uint32_t* __restrict__ dhi, uint32_t nd,
                               uint32_t* __restrict__ opat, uint32_t* __restrict__ olo, uint32_t* __restrict__ ohi) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < nf) {
        uint32_t q = fpat[t], d = t + upper_bound_u32(dpat, nd, q);
        opat[d] = q; olo[d] = flo[t]; ohi[d] = fhi[t];
    } else if (t < nf + nd) {
        uint32_t j = t - nf, q = dpat[j], d = j + lower_bound_u32(fpat, nf, q);
        opat[d] = q; olo[d] = dlo[j]; ohi[d] = dhi[j];
    }
}

// ---- results ----------------------------------------------------------------------------------------
// level-synchronous frontier (grouped by LOCAL pattern index) -> staging slots + per-pattern list / count of the batch
__global__ void ls_finalize_kernel(const uint32_t* __restrict__ pat, const uint32_t* __restrict__ lo, const uint32_t* __restrict__ hi, uint32_t nF,
                                   const uint32_t* __restrict__ ids, uint32_t sbase, uint2* __restrict__ staging, uint2* __restrict__ plist,
                                   uint32_t* __restrict__ counts) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nF) return;
    uint32_t q = pat[i], l = lo[i], h = hi[i];
    staging[sbase + i] = make_uint2(l, h);
    uint32_t g = ids ? ids[q] : q;
    atomicAdd(counts + g, h - l + 1);                    // "Found N", EDSBWTsearch.cpp:739-746
    if (i == 0 || pat[i - 1] != q) {                     // head of the pattern's run: list start + length
        uint32_t e = upper_bound_u32(pat, nF, q);
        plist[g] = make_uint2(sbase + i, e - i);
    }
}
// Locate (indices_of_dollars_in_interval, EDSBWTsearch.cpp:1310-1332): row -> (string, offset) by table lookup or by LF-walk,
// then segment / index in segment.  Output grouped by ascending pattern: pattern q owns out[occ_ofs[q] .. occ_ofs[q+1]).
template <bool WALK>
__device__ __forceinline__ uint32_t locate_row(const IndexView& ix, uint32_t row, uint32_t pat, edsbwt_occ* __restrict__ dst) {
    uint32_t w, off;
    uint2 si;
    if (WALK) {
        off = 0;
        while (lf_row_any(ix, row)) ++off;
        w = __ldg(ix.eof_flat + rank_eof_any(ix, row));
        si = __ldg(ix.str_info + w);
    } else { const uint4 v = __ldg(ix.sa4 + row); w = v.x; off = v.y; si = make_uint2(v.z, v.w); }      // ONE 16-byte gather per occurrence
    edsbwt_occ o; o.pat = pat; o.str = w; o.seg = si.x; o.idx_in_seg = si.y; o.off = off;
    *dst = o;
    return off + 1;
}
constexpr uint32_t LOCATE_LIGHT_MAX = 32;      // patterns with more occurrences go to the warp-per-pattern kernel
// one thread per PATTERN (nearly every pattern has 0-2 occurrences): counts, plist and staged intervals are read
// coalesced, no search for the owner of a row; heavy patterns are queued for locate_heavy_kernel
template <bool WALK>
__global__ void locate_light_kernel(const __grid_constant__ IndexView ix, const uint2* __restrict__ plist, const uint32_t* __restrict__ occ_ofs, const uint2* __restrict__ staging,
                                    const uint2* const* __restrict__ seed_ivp, uint32_t npat, uint32_t pat_base, edsbwt_occ* __restrict__ out,
                                    uint32_t* __restrict__ heavy, uint32_t* __restrict__ nheavy, StatsDev* __restrict__ st) {
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t steps = 0, nrows = 0;
    if (q < npat) {
        const uint32_t o0 = __ldg(occ_ofs + q), cnt = __ldg(occ_ofs + q + 1) - o0;
        if (cnt > LOCATE_LIGHT_MAX) heavy[atomicAdd(nheavy, 1u)] = q;
        else if (cnt) {
            uint2 pl = __ldg(plist + q);
            uint32_t t = o0;
            if (pl.y == PLIST_INLINE) {                    // one interval, carried by the entry itself: rows pl.x .. pl.x + cnt - 1
                for (uint32_t row = pl.x; row < pl.x + cnt; ++row) steps += locate_row<WALK>(ix, row, q + pat_base, out + t++);
            } else {
                const uint2* ivs = staging + pl.x;
                if (pl.y & PLIST_SEED) { ivs = seed_ivp[(pl.y >> 27) & 15u] + pl.x; pl.y &= PLIST_N_MASK; }
                for (uint32_t j = 0; j < pl.y; ++j) {
                    const uint2 iv = __ldg(ivs + j);
                    for (uint32_t row = iv.x; row <= iv.y; ++row) steps += locate_row<WALK>(ix, row, q + pat_base, out + t++);
                }
            }
            nrows = cnt;
        }
    }
    if (st) { warp_add(&st->L, steps); warp_add(&st->O, nrows); }
}
// one WARP per heavy pattern: the intervals are walked once, the lanes share the rows of each interval
template <bool WALK>
__global__ void locate_heavy_kernel(const __grid_constant__ IndexView ix, const uint2* __restrict__ plist, const uint32_t* __restrict__ occ_ofs, const uint2* __restrict__ staging,
                                    const uint2* const* __restrict__ seed_ivp, uint32_t pat_base, edsbwt_occ* __restrict__ out,
                                    const uint32_t* __restrict__ heavy, const uint32_t* __restrict__ nheavy, StatsDev* __restrict__ st) {
    const uint32_t lane = threadIdx.x & 31, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = (gridDim.x * blockDim.x) >> 5;
    const uint32_t nh = *nheavy;
    uint32_t steps = 0, nrows = 0;
    for (uint32_t i = warp; i < nh; i += nwarps) {
        const uint32_t q = heavy[i];
        uint2 pl = __ldg(plist + q);
        const uint2* ivs = staging + pl.x;
        uint32_t t = __ldg(occ_ofs + q);
        const bool inl = pl.y == PLIST_INLINE;             // one interval carried by the entry itself
        const uint32_t inl_hi = pl.x + (__ldg(occ_ofs + q + 1) - t) - 1;
        if (inl) pl.y = 1;
        else if (pl.y & PLIST_SEED) { ivs = seed_ivp[(pl.y >> 27) & 15u] + pl.x; pl.y &= PLIST_N_MASK; }
        // 32 intervals at a time: lane l holds interval j0 + l; an exclusive warp scan of the lengths places every row
        for (uint32_t j0 = 0; j0 < pl.y; j0 += 32) {
            uint2 iv = make_uint2(1u, 0u);
            if (j0 + lane < pl.y) iv = inl ? make_uint2(pl.x, inl_hi) : __ldg(ivs + j0 + lane);
            const uint32_t len = iv.y + 1 - iv.x;                      // 0 for the padding lanes
            uint32_t pre = len;
            for (int o = 1; o < 32; o <<= 1) { uint32_t v = __shfl_up_sync(0xffffffffu, pre, o); if (lane >= (uint32_t)o) pre += v; }
            const uint32_t total = __shfl_sync(0xffffffffu, pre, 31);
            pre -= len;                                                // exclusive
            // rows k = 0 .. total-1 of this group: owner = last lane with pre <= k (binary search over the 32 prefix values by shuffles)
            for (uint32_t k = lane; k < ((total + 31) & ~31u); k += 32) {
                uint32_t lo = 0;
                #pragma unroll
                for (int b = 16; b; b >>= 1) { const uint32_t pv = __shfl_sync(0xffffffffu, pre, (lo + b) & 31); if (lo + b < 32 && pv <= k) lo += b; }
                // skip empty (padding / zero-length) owners: pre is non-decreasing, the LAST lane with pre <= k has len > 0 when k < total
                const uint32_t ivx = __shfl_sync(0xffffffffu, iv.x, lo), p0 = __shfl_sync(0xffffffffu, pre, lo);
                if (k < total) { steps += locate_row<WALK>(ix, ivx + (k - p0), q + pat_base, out + t + k); ++nrows; }
            }
            t += total;
        }
    }
    if (st) { warp_add(&st->L, steps); warp_add(&st->O, nrows); }
}

// ===================================================================================================
// host side
// ===================================================================================================
struct Tri {
    DevBuf<uint32_t> pat, lo, hi;
    void ensure(size_t n) { pat.ensure(n); lo.ensure(n); hi.ensure(n); }
    void swap(Tri& o) { std::swap(pat.p, o.pat.p); std::swap(lo.p, o.lo.p); std::swap(hi.p, o.hi.p);
                        std::swap(pat.cap, o.pat.cap); std::swap(lo.cap, o.lo.cap); std::swap(hi.cap, o.hi.cap); }
};

struct Workspace {
    Tri cur, raw, cmp, mrg, dol, dolraw, nxt;
    DevBuf<uint32_t> keep, pos, dcnt, dstart, dofs, head, len, cnt, src, base, plen;
    DevBuf<uint64_t> keys, keys2, ppack;
    DevBuf<uint8_t> cubtmp;
    DevBuf<unsigned int> scalars;        // [0] nvalid
    DevBuf<uint2> plist, staging;        // per-pattern final lists (fused + level-synchronous results)
    DevBuf<uint32_t> fcursor, ovf_ids, ovf_ids2, occ_ofs, heavy;
    DevBuf<uint64_t> pwords, uoff;       // 2-bit pattern words of the batch (pack_patterns_kernel); offsets generated for fixed-length batches
    DevBuf<uint32_t> pinvalid;
    DevBuf<uint8_t> ubytes;              // characters of packed patterns, unpacked for the tiers that read characters
    StatsDev* d_stats = nullptr;
    uint32_t* h_pin = nullptr;           // pinned read-back area
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, evk0 = nullptr, evk1 = nullptr, evf0 = nullptr, evf1 = nullptr, evl0 = nullptr, evl1 = nullptr;
    uint64_t launches = 0;
    double lf_ms = 0;
    ~Workspace() { if (d_stats) cudaFree(d_stats); if (h_pin) cudaFreeHost(h_pin); if (ev0) cudaEventDestroy(ev0); if (ev1) cudaEventDestroy(ev1);
                   if (evk0) cudaEventDestroy(evk0); if (evk1) cudaEventDestroy(evk1); if (evf0) cudaEventDestroy(evf0); if (evf1) cudaEventDestroy(evf1);
                   if (evl0) cudaEventDestroy(evl0); if (evl1) cudaEventDestroy(evl1); }
};

Index::~Index() {
    for (SeedLevel* s : seed) delete s;
    delete ws;
    delete pipe;
    delete move;
    free_csv_scratch(csv);
    if (push) cudaStreamDestroy(push);
    if (stream) cudaStreamDestroy(stream);
}
uint64_t Index::device_bytes() const {
    uint64_t b = occ.bytes() + eof_flat.bytes() + eof_inv.bytes() + str_seg.bytes() + seg_first.bytes() + seg_low.bytes() + sa4.bytes() + link_rec.bytes() + str_info.bytes();
    for (SeedLevel* s : seed) if (s) b += s->off.bytes() + s->iv.bytes() + s->cnt.bytes() + s->first.bytes();
    if (move) b += move->bytes();
    return b;
}

static void exclusive_sum(Workspace& ws, const uint32_t* in, uint32_t* out, size_t n, cudaStream_t s) {
    size_t tmp = 0;
    CUDA_OK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, in, out, (int)n, s));
    ws.cubtmp.ensure(tmp);
    CUDA_OK(cub::DeviceScan::ExclusiveSum(ws.cubtmp.p, tmp, in, out, (int)n, s));
    ws.launches += 2;
}
static uint32_t read_u32(Workspace& ws, const uint32_t* d, cudaStream_t s) {
    CUDA_OK(cudaMemcpyAsync(ws.h_pin, d, 4, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    return ws.h_pin[0];
}
Lut make_lut(const HostIndex& h) {
    Lut l; memset(l.code, 255, sizeof l.code);
    l.code[h.alphabet[0]] = h.sigma > 5 ? 8 : 4;      // end-marker: the value the Occ pack kernel of the layout tests for
    for (uint32_t a = 1; a < h.sigma; ++a) l.code[h.alphabet[a]] = (uint8_t)(a - 1);
    return l;
}

// drop empty intervals and merge adjacent ones: raw(+keep) -> out ; returns the new count
template <bool KEYMODE>
static uint32_t compact_and_merge(Index& ix, Workspace& ws, PatView pv, uint32_t step, const uint32_t* pat, Tri& raw, uint32_t n, Tri& cmp, Tri& out) {
    cudaStream_t s = ix.stream;
    if (n == 0) return 0;
    ws.pos.ensure(n + 1); ws.head.ensure(n + 1); cmp.ensure(n); out.ensure(n);
    // keep[n] = 0 so that pos[n] = number kept
    CUDA_OK(cudaMemsetAsync(ws.keep.p + n, 0, 4, s));
    exclusive_sum(ws, ws.keep.p, ws.pos.p, n + 1, s);
    compact_kernel<<<grid_for(n, 256), 256, 0, s>>>(pat, raw.lo.p, raw.hi.p, ws.keep.p, ws.pos.p, n, cmp.pat.p, cmp.lo.p, cmp.hi.p);
    // head flags on the compacted list (count read on the device: no host round trip)
    const uint32_t* d_nkeep = ws.pos.p + n;
    uint32_t* d_nk_copy = ws.scalars.p + 1;        // pos is overwritten by the next scan: park the count
    CUDA_OK(cudaMemcpyAsync(d_nk_copy, d_nkeep, 4, cudaMemcpyDeviceToDevice, s));
    head_flags_kernel<KEYMODE><<<grid_for(n + 1, 256), 256, 0, s>>>(pv, step, cmp.pat.p, cmp.lo.p, cmp.hi.p, d_nk_copy, n + 1, ws.head.p);
    exclusive_sum(ws, ws.head.p, ws.pos.p, n + 1, s);
    merge_kernel<<<grid_for(n, 256), 256, 0, s>>>(cmp.pat.p, cmp.lo.p, cmp.hi.p, ws.head.p, ws.pos.p, d_nk_copy, n, out.pat.p, out.lo.p, out.hi.p);
    ws.launches += 4;
    return read_u32(ws, ws.pos.p + n, s);
}

// One level-synchronous backward step over the whole frontier ws.cur[0..nF) -> ws.cur ; returns new nF
template <bool KEYMODE>
static uint32_t engine_step(Index& ix, Workspace& ws, PatView pv, uint32_t step, uint32_t nF, StatsDev* st, uint32_t kc, bool time_lf) {
    cudaStream_t s = ix.stream;
    if (nF == 0) return 0;
    ws.raw.ensure(nF); ws.keep.ensure(nF + 1); ws.dcnt.ensure(nF + 1); ws.dstart.ensure(nF); ws.dofs.ensure(nF + 2);
    if (time_lf) CUDA_OK(cudaEventRecord(ws.evk0, s));
    if (ix.wide) lf_step_kernel<KEYMODE, true><<<grid_for((size_t)nF * 4, 256), 256, 0, s>>>(ix.view, pv, step, ws.cur.pat.p, ws.cur.lo.p, ws.cur.hi.p, nF,
        ws.raw.lo.p, ws.raw.hi.p, ws.dcnt.p, ws.dstart.p, ws.keep.p, st, kc, 0);
    else lf_step_kernel<KEYMODE><<<grid_for((size_t)nF * 4, 256), 256, 0, s>>>(ix.view, pv, step, ws.cur.pat.p, ws.cur.lo.p, ws.cur.hi.p, nF,
        ws.raw.lo.p, ws.raw.hi.p, ws.dcnt.p, ws.dstart.p, ws.keep.p, st, kc, 0);
    if (time_lf) { CUDA_OK(cudaEventRecord(ws.evk1, s)); }
    ws.launches += 1;
    // end-markers inside the frontier -> link
    CUDA_OK(cudaMemsetAsync(ws.dcnt.p + nF, 0, 4, s));
    exclusive_sum(ws, ws.dcnt.p, ws.dofs.p, nF + 1, s);
    uint32_t D = read_u32(ws, ws.dofs.p + nF, s);
    if (time_lf) { float ms = 0; CUDA_OK(cudaEventElapsedTime(&ms, ws.evk0, ws.evk1)); ws.lf_ms += ms; }
    uint32_t nDol = 0;
    if (D > 0) {
        ws.keys.ensure(D); ws.keys2.ensure(D);
        CUDA_OK(cudaMemsetAsync(ws.scalars.p, 0, 4, s));
        enum_dollars_kernel<<<grid_for(D, 256), 256, 0, s>>>(ix.view, ws.cur.pat.p, ws.dofs.p, ws.dstart.p, nF, D, ws.keys.p, ws.scalars.p);
        ws.launches += 1;
        uint32_t Dv = read_u32(ws, ws.scalars.p, s);
        if (Dv > 0) {
            size_t tmp = 0;
            int end_bit = 64;
            CUDA_OK(cub::DeviceRadixSort::SortKeys(nullptr, tmp, ws.keys.p, ws.keys2.p, (int)D, 0, end_bit, s));
            ws.cubtmp.ensure(tmp);
            CUDA_OK(cub::DeviceRadixSort::SortKeys(ws.cubtmp.p, tmp, ws.keys.p, ws.keys2.p, (int)D, 0, end_bit, s));
            ws.launches += 4;
            ws.head.ensure(Dv + 1); ws.pos.ensure(Dv + 1); ws.dol.ensure(Dv);
            dol_heads_kernel<<<grid_for(Dv, 256), 256, 0, s>>>(ix.view, ws.keys2.p, Dv, ws.head.p);
            CUDA_OK(cudaMemsetAsync(ws.head.p + Dv, 0, 4, s));
            exclusive_sum(ws, ws.head.p, ws.pos.p, Dv + 1, s);
            dol_build_kernel<<<grid_for(Dv, 256), 256, 0, s>>>(ix.view, ws.keys2.p, ws.head.p, ws.pos.p, Dv, ws.dol.pat.p, ws.dol.lo.p, ws.dol.hi.p);
            ws.launches += 2;
            nDol = read_u32(ws, ws.pos.p + Dv, s);
        }
    }
    // F' : drop empty, merge adjacent
    uint32_t nF2 = compact_and_merge<KEYMODE>(ix, ws, pv, step, ws.cur.pat.p, ws.raw, nF, ws.cmp, ws.mrg);
    if (nDol == 0) { ws.cur.swap(ws.mrg); return nF2; }
    // Dol' : LF of the pile-0 ranges, then the same clean-up
    ws.dolraw.ensure(nDol); ws.keep.ensure(nDol + 1); ws.dcnt.ensure(nDol + 1); ws.dstart.ensure(nDol + 1);
    if (time_lf) CUDA_OK(cudaEventRecord(ws.evk0, s));
    if (ix.wide) lf_step_kernel<KEYMODE, true><<<grid_for((size_t)nDol * 4, 256), 256, 0, s>>>(ix.view, pv, step, ws.dol.pat.p, ws.dol.lo.p, ws.dol.hi.p, nDol,
        ws.dolraw.lo.p, ws.dolraw.hi.p, ws.dcnt.p, ws.dstart.p, ws.keep.p, st, kc, 1);
    else lf_step_kernel<KEYMODE><<<grid_for((size_t)nDol * 4, 256), 256, 0, s>>>(ix.view, pv, step, ws.dol.pat.p, ws.dol.lo.p, ws.dol.hi.p, nDol,
        ws.dolraw.lo.p, ws.dolraw.hi.p, ws.dcnt.p, ws.dstart.p, ws.keep.p, st, kc, 1);
    if (time_lf) CUDA_OK(cudaEventRecord(ws.evk1, s));
    ws.launches += 1;
    Tri& dcmp = ws.cmp;         // cmp is free again (mrg holds F'')
    uint32_t nDol2 = compact_and_merge<KEYMODE>(ix, ws, pv, step, ws.dol.pat.p, ws.dolraw, nDol, dcmp, ws.nxt);
    if (time_lf) { float ms = 0; CUDA_OK(cudaEventElapsedTime(&ms, ws.evk0, ws.evk1)); ws.lf_ms += ms; }
    // ws.nxt = Dol'' ; ws.mrg = F'' ; combine into cur
    uint32_t total = nF2 + nDol2;
    if (nDol2 == 0) { ws.cur.swap(ws.mrg); return nF2; }
    ws.cur.ensure(total);
    combine_kernel<<<grid_for(total, 256), 256, 0, s>>>(ws.mrg.pat.p, ws.mrg.lo.p, ws.mrg.hi.p, nF2, ws.nxt.pat.p, ws.nxt.lo.p, ws.nxt.hi.p, nDol2,
                                                       ws.cur.pat.p, ws.cur.lo.p, ws.cur.hi.p);
    ws.launches += 1;
    return total;
}

// cnt[key] = number of rows in the frontier of key: patterns no longer than the seed depth are answered by one lookup
__global__ void seed_counts_kernel(const uint32_t* __restrict__ off, const uint2* __restrict__ iv, uint32_t nkeys, uint32_t* __restrict__ cnt) {
    uint32_t key = blockIdx.x * blockDim.x + threadIdx.x;
    if (key >= nkeys) return;
    uint32_t c = 0;
    for (uint32_t i = off[key], e = off[key + 1]; i < e; ++i) { const uint2 v = __ldg(iv + i); c += v.y - v.x + 1; }
    cnt[key] = c;
}

// first[key] = {first interval, offset of the entry, number of intervals}: the seed lookup of the hot kernel in ONE 16-byte gather
__global__ void seed_first_kernel(const uint32_t* __restrict__ off, const uint2* __restrict__ iv, uint32_t nkeys, uint4* __restrict__ first) {
    uint32_t key = blockIdx.x * blockDim.x + threadIdx.x;
    if (key >= nkeys) return;
    const uint32_t a = off[key], n0 = off[key + 1] - a;
    uint2 v = make_uint2(1u, 0u);
    if (n0) v = __ldg(iv + a);
    first[key] = make_uint4(v.x, v.y, a, n0);
}

// ---------------------------------------------------------------------------------------------------
// index build on the device
// ---------------------------------------------------------------------------------------------------
static void build_seed_table(Index& ix, uint32_t k) {
    Workspace& ws = *ix.ws;
    cudaStream_t s = ix.stream;
    ix.seed.assign(k + 1, nullptr);
    const uint32_t nchild = ix.key_radix;      // 4 children per key; one per letter for a wide alphabet
    auto keys_at = [&](uint32_t L) { uint64_t v = 1; for (uint32_t i = 0; i < L; ++i) v *= nchild; return (uint32_t)v; };
    // level 1: the piles
    {
        std::vector<uint32_t> off(nchild + 1, 0), lo, hi;
        for (uint32_t c = 0; c < nchild; ++c) {
            off[c] = (uint32_t)lo.size();
            const uint32_t first = ix.host.pile_start[c + 1 <= ix.host.sigma ? c + 1 : ix.host.sigma], next = ix.host.pile_start[c + 2 <= ix.host.sigma ? c + 2 : ix.host.sigma];
            if (c < ix.view.nletters && next > first) { lo.push_back(first); hi.push_back(next - 1); }
        }
        off[nchild] = (uint32_t)lo.size();
        SeedLevel* L1 = new SeedLevel; ix.seed[1] = L1;
        std::vector<uint2> iv(lo.size());
        for (size_t i = 0; i < lo.size(); ++i) iv[i] = make_uint2(lo[i], hi[i]);
        L1->off.ensure(nchild + 1); L1->iv.ensure(lo.size() + 1); L1->n_iv = lo.size();
        CUDA_OK(cudaMemcpyAsync(L1->off.p, off.data(), (nchild + 1) * 4, cudaMemcpyHostToDevice, s));
        if (!iv.empty()) CUDA_OK(cudaMemcpyAsync(L1->iv.p, iv.data(), iv.size() * 8, cudaMemcpyHostToDevice, s));
        CUDA_OK(cudaStreamSynchronize(s));
        ix.seed_intervals = lo.size();
    }
    PatView pv{}; pv.k = 0; pv.kr = nchild;
    for (uint32_t L = 1; L < k; ++L) {
        SeedLevel* cur = ix.seed[L];
        uint32_t nkeys = keys_at(L), nF = (uint32_t)cur->n_iv;
        size_t nE = (size_t)nF * nchild;
        if (nE > 0xfffffff0ull) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "seed level %u too large", L + 1);
        ws.cur.ensure(nE ? nE : 1); ws.nxt.ensure(nF ? nF : 1);
        // pat array of the level = key of every interval (recovered from off)
        uint32_t newF = 0;
        if (nF) {
            // key of interval i = upper_bound(off, i) - 1
            ws.src.ensure(nF);
            level_keys_kernel<<<grid_for(nF, 256), 256, 0, s>>>(cur->off.p, nkeys, nF, ws.src.p);
            seed_expand_kernel<<<grid_for(nF, 256), 256, 0, s>>>(ws.src.p, cur->iv.p, cur->off.p, nF, ws.cur.pat.p, ws.cur.lo.p, ws.cur.hi.p, nchild);
            ws.launches += 2;
            newF = engine_step<true>(ix, ws, pv, 0, (uint32_t)nE, nullptr, 0, false);
        }
        SeedLevel* nx = new SeedLevel; ix.seed[L + 1] = nx;
        uint32_t nk2 = nkeys * nchild;
        nx->off.exact((size_t)nk2 + 1); nx->iv.exact((size_t)newF + 1); nx->n_iv = newF;
        if (newF) zip_kernel<<<grid_for(newF, 256), 256, 0, s>>>(ws.cur.lo.p, ws.cur.hi.p, newF, nx->iv.p);
        key_offsets_kernel<<<grid_for((size_t)nk2 + 1, 256), 256, 0, s>>>(ws.cur.pat.p, newF, nk2, nx->off.p);
        CUDA_OK(cudaStreamSynchronize(s));
        ix.seed_intervals += newF;
    }
    // per-key row counts of the levels below the deepest one (the deepest level's lists are short: summed on the fly)
    std::vector<const uint2*> ivp(17, nullptr);
    for (uint32_t L = 1; L <= k; ++L) {
        SeedLevel* lv = ix.seed[L];
        ivp[L] = lv->iv.p;
        if (L == k && k > 8) continue;
        const uint32_t nkeys = keys_at(L);
        lv->cnt.exact(nkeys);
        seed_counts_kernel<<<grid_for(nkeys, 256), 256, 0, s>>>(lv->off.p, lv->iv.p, nkeys, lv->cnt.p);
    }
    {                                   // deepest level: the seed lookup of the hot kernel in one 16-byte gather
        SeedLevel* lv = ix.seed[k];
        const uint32_t nkeys = keys_at(k);
        lv->first.exact(nkeys);
        seed_first_kernel<<<grid_for(nkeys, 256), 256, 0, s>>>(lv->off.p, lv->iv.p, nkeys, lv->first.p);
    }
    ix.seed_ivp.ensure(17);
    CUDA_OK(cudaMemcpyAsync(ix.seed_ivp.p, ivp.data(), 17 * sizeof(const uint2*), cudaMemcpyHostToDevice, s));
    CUDA_OK(cudaStreamSynchronize(s));
    CUDA_OK(cudaGetLastError());
}
void init_index_runtime(Index& ix, const edsbwt_open_opts& o) {
    HostIndex& h = ix.host;
    CUDA_OK(cudaSetDevice(ix.device));
    CUDA_OK(cudaStreamCreateWithFlags(&ix.stream, cudaStreamNonBlocking));
    cudaStream_t s = ix.stream;
    ix.ws = new Workspace;
    Workspace& ws = *ix.ws;
    CUDA_OK(cudaMalloc(&ws.d_stats, 2 * sizeof(StatsDev)));      // [1]: 64-bit scratch (occurrence total)
    CUDA_OK(cudaMallocHost(&ws.h_pin, 256));
    CUDA_OK(cudaEventCreate(&ws.ev0)); CUDA_OK(cudaEventCreate(&ws.ev1)); CUDA_OK(cudaEventCreate(&ws.evk0)); CUDA_OK(cudaEventCreate(&ws.evk1));
    CUDA_OK(cudaEventCreate(&ws.evf0)); CUDA_OK(cudaEventCreate(&ws.evf1)); CUDA_OK(cudaEventCreate(&ws.evl0)); CUDA_OK(cudaEventCreate(&ws.evl1));
    ix.engine = (o.engine == EDSBWT_ENGINE_LEVELSYNC || o.engine == EDSBWT_ENGINE_FUSED_STEP || o.engine == EDSBWT_ENGINE_ITEMS) ? o.engine : EDSBWT_ENGINE_FUSED;
    ix.fused_cap = o.fused_cap > 0 ? (uint32_t)o.fused_cap : 13u;       // 13 slots: 12 CTAs/SM (40 registers) and >= 60 KB of L1 left (A/B: profiles/r01_ab_keep_sort_sizes.txt)
    ix.fused_cap_auto = o.fused_cap <= 0;
    if (ix.fused_cap < 4) ix.fused_cap = 4;
    if (ix.fused_cap > FUSED_BIG_CAP) ix.fused_cap = FUSED_BIG_CAP;
    ix.fused_cap2 = o.fused_cap2 > 0 ? std::min<uint32_t>((uint32_t)o.fused_cap2, FUSED_BIG_CAP) : FUSED_BIG_CAP;
    if (ix.fused_cap2 < 4) ix.fused_cap2 = 4;
    { Lut l = make_lut(h); ix.lut.ensure(256); CUDA_OK(cudaMemcpyAsync(ix.lut.p, l.code, 256, cudaMemcpyHostToDevice, s)); CUDA_OK(cudaStreamSynchronize(s)); }
    ws.scalars.ensure(16);
    if (const char* e = getenv("EDSBWT_OCC_LIMIT")) { const unsigned long long v = strtoull(e, nullptr, 10); if (v && v < ix.occ_limit) ix.occ_limit = v; }
}

void build_device_index(Index& ix, const edsbwt_open_opts& o) {
    HostIndex& h = ix.host;
    init_index_runtime(ix, o);
    cudaStream_t s = ix.stream;
    Workspace& ws = *ix.ws;
    IndexView& v = ix.view;
    v.n = h.n; v.n_text = h.n_text; v.n_seg = h.n_seg; v.nletters = h.sigma - 1;
    for (uint32_t c = 0; c < 5; ++c) v.Cl[c] = c < h.sigma - 1 ? h.pile_start[c + 1] : h.n;
    // 5..7 letters: 3-bit codes in 96-byte Occ blocks, seed keys in radix nletters
    ix.wide = h.sigma > 5; ix.key_radix = ix.wide ? h.sigma - 1 : 4;
    v.wide = ix.wide ? 1u : 0u;
    for (uint32_t c = 0; c < 9; ++c) v.Clw[c] = c < h.sigma - 1 ? h.pile_start[c + 1] : h.n;
    if (ix.wide && ix.engine == EDSBWT_ENGINE_ITEMS) ix.engine = EDSBWT_ENGINE_FUSED;      // the item kernel reads 2-bit words

    // ---- K1: Occ blocks
    {
        DevBuf<uint8_t> d_bytes; d_bytes.ensure((size_t)h.n + 64);
        CUDA_OK(cudaMemcpyAsync(d_bytes.p, h.ebwt.data(), h.n, cudaMemcpyHostToDevice, s));
        uint32_t tot[8] = {0}; unsigned int hbad = 0;
        if (ix.wide) build_occ_blocks_wide(d_bytes.p, h.n, make_lut(h), ix.occ, ws.cubtmp, s, tot, &hbad);
        else build_occ_blocks(d_bytes.p, h.n, make_lut(h), ix.occ, ws.cubtmp, s, tot, &hbad);
        if (hbad) EDS_FAIL(EDSBWT_E_FORMAT, "eBWT row %u holds a byte outside the alphabet of info.aux", hbad - 1);
        // totals must reproduce the pile sizes implied by tableOcc
        for (uint32_t c = 0; c + 1 < h.sigma; ++c) {
            uint32_t expect = 0;
            for (uint32_t p = 0; p < h.sigma; ++p) expect += h.table_occ[p * h.sigma + c + 1];
            if (tot[c] != expect) EDS_FAIL(EDSBWT_E_FORMAT, "letter %c occurs %u times in the piles, tableOcc says %u", h.alphabet[c + 1], tot[c], expect);
            // pile of letter c holds exactly the suffixes starting with c = number of occurrences of c
            if (h.pile_start[c + 2] - h.pile_start[c + 1] != expect) EDS_FAIL(EDSBWT_E_FORMAT, "pile %u has %u rows but letter %c occurs %u times", c + 1, h.pile_start[c + 2] - h.pile_start[c + 1], h.alphabet[c + 1], expect);
        }
    }
    v.occ = ix.occ.p;
    // ---- keep the Occ blocks resident in L2 (B200: 126 MB; the cfg4 Occ array is 142 MB): every search step is one random
    //      64-byte fetch from this array, everything else the search touches (seed table, patterns, results) streams through.
    //      MEASURED (profiles/r01_l2_persist_ab.txt): 4.84 ms with the window vs 4.54 ms without at cfg4 / k = 14 -- the
    //      carve-out takes more from the streaming traffic than it gives the Occ fetches, so it is OFF unless
    //      EDSBWT_L2_PERSIST=1.
    {
        const char* e = getenv("EDSBWT_L2_PERSIST");
        cudaDeviceProp prop{};
        if (e && e[0] == '1' && cudaGetDeviceProperties(&prop, ix.device) == cudaSuccess && prop.persistingL2CacheMaxSize > 0) {
            size_t occ_bytes = (size_t)(h.n / OCC_ROWS + 1) * 64;
            size_t carve = std::min<size_t>(occ_bytes, (size_t)prop.persistingL2CacheMaxSize);
            size_t win = std::min<size_t>(occ_bytes, (size_t)prop.accessPolicyMaxWindowSize);
            if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve) == cudaSuccess && win) {
                cudaStreamAttrValue attr{};
                attr.accessPolicyWindow.base_ptr = (void*)ix.occ.p;
                attr.accessPolicyWindow.num_bytes = win;
                attr.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)carve / (double)win);
                attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
                attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
                if (cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &attr) == cudaSuccess) ix.l2_persist_bytes = carve;
            }
            cudaGetLastError();
        }
    }
    // ---- end-marker array and segment structure (rank/select on .bitvector, closure through empty strings)
    ix.eof_flat.ensure(h.n_text + 1); ix.str_seg.ensure(h.n_text + 1); ix.seg_first.ensure((size_t)h.n_seg + 3); ix.seg_low.ensure((size_t)h.n_seg + 2);
    {
        std::vector<uint32_t> str_seg(h.n_text), seg_first(h.n_seg + 2, 0), seg_low(h.n_seg + 1, 0);
        std::vector<uint8_t> has_empty(h.n_seg + 1, 0);
        uint32_t seg = 0;
        for (uint32_t w = 0; w < h.n_text; ++w) {
            if (h.seg_first_bit[w]) seg_first[++seg] = w;
            str_seg[w] = seg;
            if (h.ebwt[w] == h.alphabet[0]) has_empty[seg] = 1;        // pile-0 row w holds '#' <=> string w is empty
        }
        seg_first[h.n_seg + 1] = h.n_text;
        for (uint32_t sg = 1; sg <= h.n_seg; ++sg) seg_low[sg] = (sg == 1 || !has_empty[sg]) ? sg : seg_low[sg - 1];
        v.fsi = h.n_seg >= 2 ? seg_first[2] - 1 : h.n_text - 1;      // first_symbol_index (EDSBWTsearch.cpp:168)
        CUDA_OK(cudaMemcpyAsync(ix.eof_flat.p, h.eof_flat.data(), (size_t)h.n_text * 4, cudaMemcpyHostToDevice, s));
        CUDA_OK(cudaMemcpyAsync(ix.str_seg.p, str_seg.data(), (size_t)h.n_text * 4, cudaMemcpyHostToDevice, s));
        CUDA_OK(cudaMemcpyAsync(ix.seg_first.p, seg_first.data(), seg_first.size() * 4, cudaMemcpyHostToDevice, s));
        CUDA_OK(cudaMemcpyAsync(ix.seg_low.p, seg_low.data(), seg_low.size() * 4, cudaMemcpyHostToDevice, s));
        CUDA_OK(cudaStreamSynchronize(s));
    }
    v.eof_flat = ix.eof_flat.p; v.str_seg = ix.str_seg.p; v.seg_first = ix.seg_first.p; v.seg_low = ix.seg_low.p;
    // ---- link records + string info (one gather each in the hot kernels)
    ix.link_rec.ensure((size_t)h.n_text + 1); ix.str_info.ensure((size_t)h.n_text + 1);
    link_rec_kernel<<<grid_for(h.n_text, 256), 256, 0, s>>>(v, ix.link_rec.p, ix.str_info.p);
    v.link_rec = ix.link_rec.p; v.str_info = ix.str_info.p;
    // ---- SA table
    ix.locate_mode = o.locate_mode == EDSBWT_LOCATE_WALK ? EDSBWT_LOCATE_WALK : EDSBWT_LOCATE_SA;
    if (ix.locate_mode == EDSBWT_LOCATE_SA) {
        DevBuf<uint2> sa; sa.ensure((size_t)h.n + 1);                 // (string, distance to the end): build scratch
        ix.sa4.exact((size_t)h.n + 1);
        DevBuf<uint32_t> slen; slen.ensure(h.n_text + 1);
        CUDA_OK(cudaMemsetAsync(sa.p, 0xff, (size_t)h.n * 8, s));
        CUDA_OK(cudaMemsetAsync(ix.sa4.p, 0xff, (size_t)h.n_text * 16, s));     // end-marker rows are never located
        sa_walk_kernel<<<grid_for(h.n_text, 128), 128, 0, s>>>(v, sa.p, slen.p);
        if (h.n > h.n_text) sa_offset_kernel<<<grid_for(h.n - h.n_text, 256), 256, 0, s>>>(h.n_text, h.n, sa.p, slen.p, ix.str_info.p, ix.sa4.p);
        CUDA_OK(cudaStreamSynchronize(s));
        v.sa4 = ix.sa4.p;
    }
    CUDA_OK(cudaStreamSynchronize(s));
    // ---- seed table
    int k = o.seed_k;
    // depth = floor(log_radix n), at most 2^28 keys (4^14; 5^12 for {A,C,G,N,T}) and at most the 16 characters of a wide pattern window
    auto keys_of = [&](int kk) { uint64_t v = 1; for (int i = 0; i < kk; ++i) v *= ix.key_radix; return v; };
    int kmax = 1; while (kmax < 14 && keys_of(kmax + 1) <= (1ull << 28)) ++kmax;
    if (k < 0) { k = 1; while (k < kmax && keys_of(k + 1) <= (uint64_t)h.n) ++k; }
    if (k < 1) k = 1;
    if (k > kmax) k = kmax;
    ix.seed_k = (uint32_t)k;
    build_seed_table(ix, ix.seed_k);
    CUDA_OK(cudaGetLastError());
    if (o.build_move && !ix.wide) build_move_table(ix);
    h.ebwt.clear(); h.ebwt.shrink_to_fit();
    h.seg_first_bit.clear(); h.seg_first_bit.shrink_to_fit();
    h.eof_flat.clear(); h.eof_flat.shrink_to_fit();
}

// ---------------------------------------------------------------------------------------------------
// batched search
// ---------------------------------------------------------------------------------------------------
// Level-synchronous search of a subset (ids == nullptr: the whole batch).  Leaves the final frontier, grouped by LOCAL
// pattern index, in ws.cur and returns its length.
// alive[g] = longest suffix of pattern g with a non-empty frontier (edsbwt_search_trace): seed part ...
__global__ void ls_alive_init_kernel(PatView pv, SeedView sv, const uint32_t* __restrict__ cnt, const uint32_t* __restrict__ ids, uint32_t* __restrict__ alive) {
    uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= pv.npat) return;
    uint32_t m = pv.len[q], la = 0;
    if (m) {
        uint32_t L0 = m < pv.k ? m : pv.k;
        if (cnt[q]) la = L0;
        else for (uint32_t L = L0 - 1; L >= 1; --L) {
            uint32_t key = 0;
            for (uint32_t i = 0, mul = 1; i < L; ++i, mul *= pv.kr) key += pat_char(pv, q, m - L + i) * mul;
            if (__ldg(sv.off[L] + key + 1) > __ldg(sv.off[L] + key)) { la = L; break; }
        }
    }
    alive[ids ? ids[q] : q] = la;
}
// ... and one mark per backward step for every pattern still present in the frontier
__global__ void ls_alive_mark_kernel(PatView pv, uint32_t step, const uint32_t* __restrict__ pat, uint32_t nF, const uint32_t* __restrict__ ids, uint32_t* __restrict__ alive) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nF) return;
    uint32_t q = pat[i];
    if (i && pat[i - 1] == q) return;
    uint32_t m = pv.len[q], L = (m < pv.k ? m : pv.k) + step + 1;
    alive[ids ? ids[q] : q] = L < m ? L : m;
}

// longest pattern of a subset (only needed by the level-synchronous tier; computed on demand so that the host never
// walks the offsets of a batch)
__global__ void max_len_kernel(const uint64_t* __restrict__ off, const uint32_t* __restrict__ ids, uint32_t nsub, unsigned int* __restrict__ out) {
    uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t m = 0;
    if (q < nsub) { uint32_t g = ids ? ids[q] : q; uint64_t l = off[g + 1] - off[g]; m = l > 0xffffffffull ? 0u : (uint32_t)l; }
    for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}

static uint32_t levelsync_search(Index& ix, const uint8_t* d_bytes, const uint64_t* d_off, const uint32_t* ids, uint32_t nsub, uint32_t max_len,
                                 StatsDev* st, uint32_t& steps, uint64_t& peak, uint32_t* d_alive, uint32_t k_eff) {
    Workspace& ws = *ix.ws;
    cudaStream_t s = ix.stream;
    const uint32_t cb = ix.wide ? 4u : 2u, per = 64 / cb;
    uint32_t words = (max_len + per - 1) / per; if (!words) words = 1;
    ws.ppack.ensure((size_t)nsub * words); ws.plen.ensure(nsub);
    Lut lut = make_lut(ix.host);
    encode_patterns_kernel<<<grid_for(nsub, 128), 128, 0, s>>>(d_bytes, d_off, ids, nsub, lut, words, ws.ppack.p, ws.plen.p, ix.view.nletters, cb);
    PatView pv{ws.ppack.p, ws.plen.p, words, nsub, k_eff, cb, ix.key_radix};
    SeedView sv{};
    for (uint32_t L = 1; L <= ix.seed_k; ++L) { sv.off[L] = ix.seed[L]->off.p; sv.iv[L] = ix.seed[L]->iv.p; sv.cnt[L] = ix.seed[L]->cnt.p; }
            sv.first = ix.seed[ix.seed_k]->first.p; sv.first_level = ix.seed_k;
    ws.cnt.ensure(nsub + 1); ws.src.ensure(nsub); ws.base.ensure(nsub + 1);
    seed_count_kernel<<<grid_for(nsub, 256), 256, 0, s>>>(pv, sv, ws.cnt.p, ws.src.p);
    CUDA_OK(cudaMemsetAsync(ws.cnt.p + nsub, 0, 4, s));
    exclusive_sum(ws, ws.cnt.p, ws.base.p, (size_t)nsub + 1, s);
    ws.launches += 2;
    uint32_t nF = read_u32(ws, ws.base.p + nsub, s);
    if (d_alive) { ls_alive_init_kernel<<<grid_for(nsub, 256), 256, 0, s>>>(pv, sv, ws.cnt.p, ids, d_alive); ws.launches += 1; }
    ws.cur.ensure(nF ? nF : 1);
    if (nF) { seed_scatter_kernel<<<grid_for(nF, 256), 256, 0, s>>>(pv, sv, ws.base.p, ws.src.p, nF, ws.cur.pat.p, ws.cur.lo.p, ws.cur.hi.p); ws.launches += 1; }
    if (nF > peak) peak = nF;
    uint32_t total_steps = max_len > k_eff ? max_len - k_eff : 0;
    for (uint32_t t = 0; t < total_steps && nF; ++t) {
        nF = engine_step<false>(ix, ws, pv, t, nF, st, ix.canonical_k, true);
        if (d_alive && nF) { ls_alive_mark_kernel<<<grid_for(nF, 256), 256, 0, s>>>(pv, t, ws.cur.pat.p, nF, ids, d_alive); ws.launches += 1; }
        ++steps;
        if (nF > peak) peak = nF;
    }
    if (st && nF) { count_I_kernel<<<grid_for(nF, 256), 256, 0, s>>>(pv, steps, ws.cur.pat.p, nF, st, ix.canonical_k); ws.launches += 1; }
    return nF;
}

// Clears the few control words of a search on the SMs.  cudaMemsetAsync runs on a copy engine: while the chunk pipeline keeps the
// copy engines busy with 17 MB / 43 MB transfers, each of these 32-byte memsets waited 40-70 us behind them and held up the
// search kernel queued after it (profiles/r02_e2e_timeline.txt: 0.2 ms per chunk).
__global__ void zero_words_kernel(uint32_t* __restrict__ a, uint32_t na, uint32_t* __restrict__ b, uint32_t nb) {
    const uint32_t i = threadIdx.x;
    if (i < na) a[i] = 0u;
    if (i < nb) b[i] = 0u;
}
// Read-back of the control words through the SMs: the cursors (and, after the scans, the occurrence total) are stored straight
// into the pinned host words of the workspace (cudaMallocHost memory is device-accessible under UVA).  As 32-byte
// cudaMemcpyAsync calls they ran on the device-to-host copy engine, IN ORDER behind the previous chunk's 35-60 MB tuple copy:
// the host saw the cursors up to 0.3 ms after the search kernel had ended, and the locate kernels and the next chunk's search
// started that much later (profiles/r02_e2e_timeline.txt).  total != nullptr: also occ_ofs[npat] = low word of the total.
__global__ void publish_control_kernel(const uint32_t* __restrict__ cursors, const unsigned long long* __restrict__ total,
                                       uint32_t* __restrict__ ofs_end, volatile uint32_t* __restrict__ host_words) {
    const uint32_t i = threadIdx.x;
    if (i < 8) host_words[i] = cursors[i];
    if (total && i == 8) {
        const unsigned long long t = *total;
        *ofs_end = (uint32_t)t;
        host_words[16] = (uint32_t)t; host_words[17] = (uint32_t)(t >> 32);
    }
    __threadfence_system();
}
__global__ void fixed_offsets_kernel(uint64_t* __restrict__ off, uint32_t n, uint64_t first, uint64_t len) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i <= n) off[i] = first + (uint64_t)i * len;
}

void run_search(Index& ix, const uint8_t* d_bytes, const uint64_t* d_off, uint32_t npat, uint64_t total_bytes, uint32_t max_len,
                bool locate, int lf_variant, SearchOut& out, uint32_t pat_base, const PatHint* hint) {
    if (lf_variant != EDSBWT_LF_RANK && lf_variant != EDSBWT_LF_MOVE) EDS_FAIL(EDSBWT_E_ARG, "unknown LF variant %d", lf_variant);
    const bool move = lf_variant == EDSBWT_LF_MOVE;
    if (ix.wide && move) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "the move-table variant packs 2-bit letter codes: an alphabet of %u letters is searched with EDSBWT_LF_RANK only", ix.host.sigma - 1);
    if (ix.wide && hint && hint->d_words) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "packed 2-bit patterns cannot address an alphabet of %u letters: pass bytes", ix.host.sigma - 1);
    if (move && out.want_alive) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "edsbwt_search_trace is an EDSBWTsearch (RECOVERBW=0) feature; MOVE_EDSBWTSearch always locates");
    const uint32_t fixed_len = hint ? hint->fixed_len : 0u, stride = (fixed_len + 31) / 32;
    const uint64_t* caller_words = hint ? hint->d_words : nullptr;
    if (npat && !d_off && !fixed_len) EDS_FAIL(EDSBWT_E_ARG, "a batch needs pattern offsets or a fixed pattern length");
    if (npat && !d_bytes && !caller_words) EDS_FAIL(EDSBWT_E_ARG, "a batch needs pattern bytes or packed words");
    if (caller_words && !fixed_len) EDS_FAIL(EDSBWT_E_ARG, "packed patterns are fixed-length batches");
    if (fixed_len && max_len == 0) max_len = fixed_len;
    CUDA_OK(cudaSetDevice(ix.device));
    std::lock_guard<std::mutex> lock(ix.mu);
    if (move) build_move_table(ix);            // no-op once built
    Workspace& ws = *ix.ws;
    cudaStream_t s = ix.stream;
    ws.launches = 0; ws.lf_ms = 0;
    const bool stats = ix.collect_counters != 0;
    // counters are defined relative to the canonical seed depth: a counting run starts from the seed level
    // min(k, canonical_k) so that every step beyond the canonical depth is really executed and counted
    const uint32_t k_eff = stats ? std::min(ix.seed_k, ix.canonical_k) : ix.seed_k;
    StatsDev* st = stats ? ws.d_stats : nullptr;
    static_assert(sizeof(StatsDev) / 4 <= 32, "zero_words_kernel clears at most 32 words per array");
    if (!npat) CUDA_OK(cudaMemsetAsync(ws.d_stats, 0, sizeof(StatsDev), s));
    CUDA_OK(cudaEventRecord(ws.ev0, s));
    out.npat = npat; out.nocc = 0;
    out.counts.ensure(npat ? npat : 1);
    uint32_t* d_alive = nullptr;
    if (out.want_alive) { out.alive.ensure(npat ? npat : 1); d_alive = out.alive.p; }
    uint32_t steps = 0, novf = 0, novf1 = 0; uint64_t peak = 0;
    bool fused_ran = false, locate_ran = false;
    // The step-loop kernel, the move kernel and the level-synchronous engine read CHARACTERS + offsets; the item kernel reads
    // 2-bit words.  Whatever form the caller did not supply is produced on the device, and only when a tier needs it.
    const uint8_t* cbytes = d_bytes; const uint64_t* coff = d_off;
    auto need_chars = [&](const uint32_t* sub_ids, uint32_t nsub) {
        if (!cbytes) {                                       // packed input: back to characters (only the listed patterns)
            ws.ubytes.ensure((size_t)npat * fixed_len + 64);
            uint64_t letters = ix.host.alphabet[0];          // bits 0..7: the end-marker (written for patterns flagged invalid), then the letters
            for (uint32_t a = 1; a < ix.host.sigma && a <= 4; ++a) letters |= (uint64_t)ix.host.alphabet[a] << (8 * a);
            launch_unpack_patterns(caller_words, fixed_len, stride, sub_ids, nsub, letters, hint->d_invalid, ws.ubytes.p, s);
            cbytes = ws.ubytes.p; ws.launches += 1;
            coff = nullptr;
        }
        if (!coff) {
            ws.uoff.ensure((size_t)npat + 1);
            fixed_offsets_kernel<<<grid_for((size_t)npat + 1, 256), 256, 0, s>>>(ws.uoff.p, npat, d_bytes ? hint->fixed_base : 0ull, fixed_len);
            coff = ws.uoff.p; ws.launches += 1;
        }
    };
    // total in 64 bits (the 32-bit offsets wrap silently when a batch of short patterns has more than 2^32 occurrences), the
    // exclusive scan of the counts, entry npat = the total (written on the device), and the total on its way to h_pin[16..17]
    bool scanned = false;
    auto enqueue_occ_offsets = [&]() {
        ws.occ_ofs.ensure((size_t)npat + 2);
        size_t tmp = 0, tmp2 = 0;
        cub::TransformInputIterator<unsigned long long, WidenU32, const uint32_t*> wide(out.counts.p, WidenU32());
        unsigned long long* d_total = reinterpret_cast<unsigned long long*>(ws.d_stats + 1);
        CUDA_OK(cub::DeviceReduce::Sum(nullptr, tmp2, wide, d_total, (int)npat, s));
        CUDA_OK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, out.counts.p, ws.occ_ofs.p, (int)npat, s));
        ws.cubtmp.ensure(std::max(tmp, tmp2));
        CUDA_OK(cub::DeviceReduce::Sum(ws.cubtmp.p, tmp2, wide, d_total, (int)npat, s));
        CUDA_OK(cub::DeviceScan::ExclusiveSum(ws.cubtmp.p, tmp, out.counts.p, ws.occ_ofs.p, (int)npat, s));
        ws.launches += 4;
        // occ_ofs[npat] = low word of the total (the total itself when it fits: checked on the host), cursors + total to the host
        publish_control_kernel<<<1, 32, 0, s>>>(ws.fcursor.p, d_total, ws.occ_ofs.p + npat, ws.h_pin);
        ws.launches += 1;
    };
    if (npat) {
        ws.plist.ensure(npat); ws.fcursor.ensure(8);
        // counters + cursors ([0] staging slots, [1] / [2] overflow of pass 1 / 2, [3] heavy patterns of the locate, [4] item-kernel work, [5] pack error)
        zero_words_kernel<<<1, 32, 0, s>>>(reinterpret_cast<uint32_t*>(ws.d_stats), (uint32_t)(sizeof(StatsDev) / 4), ws.fcursor.p, 8u);
        ws.launches += 1;
        uint32_t nstage = 0;
        const uint32_t* ids = nullptr;
        if (ix.engine != EDSBWT_ENGINE_LEVELSYNC || move) {
            // first pass: the item kernel for batches that are (or are asked to be) 2-bit words, the step-loop kernel for bytes
            const bool items = !move && !ix.wide && (ix.engine == EDSBWT_ENGINE_ITEMS || (ix.engine == EDSBWT_ENGINE_FUSED && caller_words));
            size_t want = (size_t)npat * 2 + (1u << 16);
            if (want > 0xfffffff0ull) want = 0xfffffff0ull;
            ws.staging.ensure(want); ws.ovf_ids.ensure(npat);
            SeedView sv{};
            for (uint32_t L = 1; L <= ix.seed_k; ++L) { sv.off[L] = ix.seed[L]->off.p; sv.iv[L] = ix.seed[L]->iv.p; sv.cnt[L] = ix.seed[L]->cnt.p; }
            sv.first = ix.seed[ix.seed_k]->first.p; sv.first_level = ix.seed_k;
            FusedOut fo{ws.plist.p, out.counts.p, ws.staging.p, (uint32_t)std::min<size_t>(ws.staging.cap, 0xfffffff0ull), ws.fcursor.p, ws.fcursor.p + 1, ws.ovf_ids.p, d_alive};
            // MOVE: 16-byte frontier entries (b, e, b', e'), so half the slots fit the same shared memory
            const uint32_t cap1 = move ? std::min<uint32_t>(ix.fused_cap, 12u) : ix.fused_cap;     // MOVE: 24 KB per CTA, 8 CTAs/SM
            const uint32_t cap2 = move ? std::min<uint32_t>(ix.fused_cap2, FUSED_BIG_CAP / 2) : ix.fused_cap2;
            if (items) {
                // ---- the hot kernel (items.cu): patterns as 2-bit words, persistent warps, lanes refilled
                PackedPats pp{};
                if (caller_words) { pp.words = caller_words; pp.invalid = hint->d_invalid; pp.poff = nullptr; }
                else {
                    const uint64_t nwords = fixed_len ? (uint64_t)npat * stride : (total_bytes >> 5) + npat + 1;
                    if (nwords > 0xfffffff0ull) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "more than 2^32 pattern words in one batch; split the batch");
                    ws.pwords.ensure(nwords + 1); ws.pinvalid.ensure(((size_t)npat + 31) / 32 + 1);
                    launch_pack_patterns(d_bytes, fixed_len ? nullptr : d_off, fixed_len, hint ? hint->fixed_base : 0ull, stride, npat, ix.lut.p,
                                         ws.pwords.p, nwords, ws.pinvalid.p, ws.fcursor.p + 5, s);
                    ws.launches += 1;
                    pp.words = ws.pwords.p; pp.invalid = ws.pinvalid.p; pp.poff = fixed_len ? nullptr : d_off;
                }
                pp.fixed_len = fixed_len; pp.stride = stride;
                CUDA_OK(cudaEventRecord(ws.evf0, s));
                launch_items_search(ix.view, sv, k_eff, pp, npat, cap1, fo, ws.fcursor.p + 4, st, ix.canonical_k, s);
                CUDA_OK(cudaEventRecord(ws.evf1, s));
            } else {
                need_chars(nullptr, npat);
                CUDA_OK(cudaEventRecord(ws.evf0, s));
                if (move) { MoveOut mo{fo.plist, fo.counts, fo.staging, fo.staging_cap, fo.cursor, fo.ovf_count, fo.ovf_ids};
                            launch_move_search(ix.view, ix.move->view, sv, k_eff, cbytes, coff, nullptr, npat, ix.lut.p, cap1, mo, s); }
                else launch_fused_search(ix.view, sv, k_eff, cbytes, coff, nullptr, npat, ix.lut.p, cap1, fo, st, ix.canonical_k, s);      // EDSBWT_ENGINE_FUSED_STEP
                CUDA_OK(cudaEventRecord(ws.evf1, s));
            }
            ws.launches += 1; fused_ran = true;
            // The counts are final unless a pattern overflowed (none at cfg4): the occurrence total and the offsets of the locate
            // are queued right behind the kernel, and ONE host round trip brings back the cursors and the total together.
            if (locate) { enqueue_occ_offsets(); scanned = true; }
            else { publish_control_kernel<<<1, 32, 0, s>>>(ws.fcursor.p, nullptr, nullptr, ws.h_pin); ws.launches += 1; }
            CUDA_OK(cudaStreamSynchronize(s));
            nstage = ws.h_pin[0]; novf = ws.h_pin[1];
            if (novf) scanned = false;
            if (ws.h_pin[5]) EDS_FAIL(EDSBWT_E_ARG, "total_bytes (%llu) is smaller than the span of the pattern offsets", (unsigned long long)total_bytes);
            novf1 = novf;
            // variant-dense indexes (cfg3: a site every 6 bp) carry longer frontiers: a second pass costs a full pattern latency,
            // so batches after one that overflowed noticeably run with 20 slots
            if (!move && !stats && ix.fused_cap_auto && ix.fused_cap < 20 && (uint64_t)novf * 1000 > npat) ix.fused_cap = 20;   // (a counting run starts from a shallower seed level: not representative)
            ids = ws.ovf_ids.p;
            if (novf && cap1 < cap2) {
                // ---- second pass: the step-loop kernel with the largest frontier that fits one CTA per SM, only over the
                //      patterns that did not fit; what still does not fit goes to the level-synchronous engine
                need_chars(ws.ovf_ids.p, novf);
                ws.ovf_ids2.ensure(novf);
                size_t want2 = (size_t)nstage + (size_t)novf * cap2;
                if (want2 > 0xfffffff0ull) want2 = 0xfffffff0ull;
                ws.staging.ensure(want2, s, true, nstage);
                FusedOut fo2{ws.plist.p, out.counts.p, ws.staging.p, (uint32_t)std::min<size_t>(ws.staging.cap, 0xfffffff0ull), ws.fcursor.p, ws.fcursor.p + 2, ws.ovf_ids2.p, d_alive};
                if (move) { MoveOut mo{fo2.plist, fo2.counts, fo2.staging, fo2.staging_cap, fo2.cursor, fo2.ovf_count, fo2.ovf_ids};
                            launch_move_search(ix.view, ix.move->view, sv, k_eff, cbytes, coff, ws.ovf_ids.p, novf, ix.lut.p, cap2, mo, s); }
                else launch_fused_search(ix.view, sv, k_eff, cbytes, coff, ws.ovf_ids.p, novf, ix.lut.p, cap2, fo2, st, ix.canonical_k, s);
                ws.launches += 1;
                CUDA_OK(cudaMemcpyAsync(ws.h_pin, ws.fcursor.p, 12, cudaMemcpyDeviceToHost, s));
                CUDA_OK(cudaStreamSynchronize(s));
                nstage = ws.h_pin[0]; novf = ws.h_pin[2];
                ids = ws.ovf_ids2.p;
            }
        } else {
            CUDA_OK(cudaMemsetAsync(out.counts.p, 0, (size_t)npat * 4, s));
            CUDA_OK(cudaMemsetAsync(ws.plist.p, 0, (size_t)npat * 8, s));
            novf = npat;
        }
        if (novf) {
            // ---- level-synchronous engine: the whole batch (ENGINE_LEVELSYNC) or the patterns whose frontier outgrew the fused kernel
            need_chars(ids, novf);
            if (max_len == 0) {
                CUDA_OK(cudaMemsetAsync(ws.scalars.p + 2, 0, 4, s));
                max_len_kernel<<<grid_for(novf, 256), 256, 0, s>>>(coff, ids, novf, ws.scalars.p + 2);
                max_len = read_u32(ws, ws.scalars.p + 2, s);
                ws.launches += 1;
            }
            uint32_t nF = levelsync_search(ix, cbytes, coff, ids, novf, max_len, st, steps, peak, d_alive, k_eff);
            if (nF) {
                if ((uint64_t)nstage + nF > 0xfffffff0ull) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "more than 2^32 result intervals in one batch; split the batch");
                ws.staging.ensure((size_t)nstage + nF, s, true, nstage);
                ls_finalize_kernel<<<grid_for(nF, 256), 256, 0, s>>>(ws.cur.pat.p, ws.cur.lo.p, ws.cur.hi.p, nF, ids, nstage, ws.staging.p, ws.plist.p, out.counts.p);
                ws.launches += 1;
            }
        }
        if (locate) {
            if (!scanned) { enqueue_occ_offsets(); CUDA_OK(cudaStreamSynchronize(s)); }
            const uint64_t nocc = *reinterpret_cast<const unsigned long long*>(ws.h_pin + 16);
            if (nocc > ix.occ_limit) throw OccOverflow{};
            out.nocc = nocc;
            if (nocc) {
                out.occ.ensure(nocc);
                CUDA_OK(cudaEventRecord(ws.evl0, s));
                if (move) launch_move_locate(ix.view, ix.move->view, ws.plist.p, ws.occ_ofs.p, ws.staging.p, npat, nocc, pat_base, out.occ.p, st, s);
                else {
                    // light patterns: one thread each; heavy ones (> 32 occurrences) are queued on the device and taken by warps
                    ws.heavy.ensure(npat); uint32_t* nheavy = ws.fcursor.p + 3;      // cleared with the cursors at the start of the search
                    const unsigned hgrid = 148 * 8;
                    if (ix.locate_mode == EDSBWT_LOCATE_WALK) {
                        locate_light_kernel<true><<<grid_for(npat, 256), 256, 0, s>>>(ix.view, ws.plist.p, ws.occ_ofs.p, ws.staging.p, ix.seed_ivp.p, npat, pat_base, out.occ.p, ws.heavy.p, nheavy, st);
                        locate_heavy_kernel<true><<<hgrid, 128, 0, s>>>(ix.view, ws.plist.p, ws.occ_ofs.p, ws.staging.p, ix.seed_ivp.p, pat_base, out.occ.p, ws.heavy.p, nheavy, st);
                    } else {
                        locate_light_kernel<false><<<grid_for(npat, 256), 256, 0, s>>>(ix.view, ws.plist.p, ws.occ_ofs.p, ws.staging.p, ix.seed_ivp.p, npat, pat_base, out.occ.p, ws.heavy.p, nheavy, st);
                        locate_heavy_kernel<false><<<hgrid, 128, 0, s>>>(ix.view, ws.plist.p, ws.occ_ofs.p, ws.staging.p, ix.seed_ivp.p, pat_base, out.occ.p, ws.heavy.p, nheavy, st);
                    }
                    ws.launches += 1;
                }
                CUDA_OK(cudaEventRecord(ws.evl1, s));
                ws.launches += 1; locate_ran = true;
                // csv row order of the reference; MOVE's order is the frontier order the kernel emits
                if (ix.row_order == EDSBWT_ROWS_REFERENCE && !move) ws.launches += order_rows_reference(ix, out, ws.occ_ofs.p, npat, s);
            }
        }
    }
    CUDA_OK(cudaEventRecord(ws.ev1, s));
    // The chunk pipeline of the host-buffer calls (api.cu) does not wait for the locate kernels here: its copy-out stream takes
    // the dependency, and the next chunk's kernels queue up behind this one's without a host round trip in between.  The
    // per-kernel times of edsbwt_last_stats are then not available (0); counters and launch counts are.
    const bool deferred = hint && hint->after && !stats;
    if (deferred) CUDA_OK(cudaStreamWaitEvent(hint->after, ws.ev1, 0));
    else CUDA_OK(cudaStreamSynchronize(s));
    CUDA_OK(cudaGetLastError());
    StatsDev hs{};
    if (stats) CUDA_OK(cudaMemcpy(&hs, ws.d_stats, sizeof hs, cudaMemcpyDeviceToHost));
    float ms = 0; if (!deferred) CUDA_OK(cudaEventElapsedTime(&ms, ws.ev0, ws.ev1));
    edsbwt_stats& ls = ix.last;
    ls = edsbwt_stats{};
    ls.I = hs.I; ls.S = hs.S; ls.E = hs.E; ls.L = hs.L; ls.O = hs.O; ls.lf_intervals = hs.lf_intervals; ls.occ_blocks = hs.occ_blocks;
    ls.steps = steps; ls.kernel_launches = ws.launches; ls.lf_kernel_ms = ws.lf_ms; ls.total_ms = ms; ls.frontier_peak = peak;
    ls.overflow_patterns = (ix.engine != EDSBWT_ENGINE_LEVELSYNC || move) ? novf1 : 0; ls.levelsync_patterns = novf;
    if (fused_ran) { float fm = 0; CUDA_OK(cudaEventElapsedTime(&fm, ws.evf0, ws.evf1)); ls.search_kernel_ms = fm; }      // (completed before the cursors were read)
    if (locate_ran && !deferred) { float lm = 0; CUDA_OK(cudaEventElapsedTime(&lm, ws.evl0, ws.evl1)); ls.locate_kernel_ms = lm; }
}

}  // namespace edsbwt
