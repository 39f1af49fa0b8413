// The hot kernel: one thread per pattern runs the WHOLE backward search after the seed lookup.
//
//   reference per-pattern loop     EDSBWT::backwardSearch   EDSBWTsearch.cpp:619-772
//   LF of every frontier interval  updateIntervals / updateSingleInterval / rankManySymbols / MergeAndRemove   :841-1209
//   segment jump                   link / dollars_in_interval / preceding_dollars_finder   :1212-1307
//
// Per thread: the frontier (sorted lists Dol' ++ F' of SURVEY App. B.2) lives in shared memory, `cap` (lo,hi) slots per
// thread, column-major so that lanes touching the same slot index do not conflict.  Each interval-step is ONE 64-byte Occ
// block fetched with two 256-bit loads (two when the interval spans blocks): letter rank, '#' rank and '#' count all come
// out of those 64 bytes.  A '#' inside the frontier costs one 8-byte gather of link_rec[] (the pile-0 range of the
// preceding segment(s), closure through empty strings included, precomputed at load); the ranges are kept as a sorted,
// merged set at the top of the thread's slot column and LF-mapped in a second pass, exactly like the reference's
// vectRangeDollarPile.  Patterns whose lists do not fit are handed to the level-synchronous engine (engine.cu), which has
// no capacity limit -- still on the GPU.
#include "internal.cuh"

namespace edsbwt {

namespace {

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace

// WIDE: the 3-bit layout of alphabets of 5..7 letters (96-byte blocks, 4-bit pattern windows of 16 characters, seed keys in
// radix nletters); every WIDE branch below is a compile-time one, the 2-bit instantiation is the code it was before.
template <bool STATS, int MINB, bool WIDE = false>
__global__ void __launch_bounds__(FUSED_THREADS, MINB)
fused_search_kernel(const __grid_constant__ IndexView ix, const __grid_constant__ SeedView sv, uint32_t k, const uint8_t* __restrict__ bytes, const uint64_t* __restrict__ poff,
                    const uint32_t* __restrict__ ids, uint32_t npat, const uint8_t* __restrict__ lut, uint32_t cap, FusedOut out,
                    StatsDev* __restrict__ st, uint32_t kc) {
    extern __shared__ uint2 s_list[];                 // [cap][FUSED_THREADS]
    __shared__ uint8_t s_lut[256];
    // kernel-parameter arrays indexed at run time (seed level, letter) would be copied to LOCAL memory by the compiler
    // (ncu: 114 M local sectors per launch): keep them in shared memory instead
    __shared__ const uint32_t* s_off[17];
    __shared__ const uint2* s_iv[17];
    __shared__ const uint32_t* s_cnt[17];
    __shared__ uint32_t s_Cl[WIDE ? 9 : 5];
    constexpr uint32_t PER = WIDE ? 16u : 32u;          // characters per 64-bit window
    for (int i = threadIdx.x; i < 256; i += FUSED_THREADS) s_lut[i] = lut[i];
    if (threadIdx.x < 17) { s_off[threadIdx.x] = sv.off[threadIdx.x]; s_iv[threadIdx.x] = sv.iv[threadIdx.x]; s_cnt[threadIdx.x] = sv.cnt[threadIdx.x]; }
    if (WIDE) { if (threadIdx.x < 9) s_Cl[threadIdx.x] = ix.Clw[threadIdx.x]; }
    else if (threadIdx.x < 5) s_Cl[threadIdx.x] = ix.Cl[threadIdx.x];
    __syncthreads();
#define LIST(j) s_list[(j) * FUSED_THREADS + threadIdx.x]
#define DOL(i) s_list[(cap - 1 - (i)) * FUSED_THREADS + threadIdx.x]

    const uint32_t tq = blockIdx.x * FUSED_THREADS + threadIdx.x;
    const bool live = tq < npat;
    const uint32_t q = live ? (ids ? ids[tq] : tq) : 0u;      // second pass (large cap): only the patterns listed in ids
    uint32_t nF = 0, m = 0, L0 = 0, key = 0, win_lo = 0;
    uint64_t base = 0, win = 0;                       // win: 2-bit codes of pattern characters [win_lo, win_lo + 32)
    bool ovf = false, direct = false;                 // direct: the answer is the seed entry itself (pattern no longer than the seed depth)
    unsigned long long sI = 0, sS = 0, sE = 0, sLF = 0, sBlk = 0;
    if (live) {
        base = poff[q];
        uint64_t len64 = poff[q + 1] - base;
        m = len64 > 0xffffffffull ? 0u : (uint32_t)len64;
        L0 = m < k ? m : k;
        bool ok = m > 0;
        // the whole pattern is checked against the alphabet first (a rejected pattern yields 0 and no trace position);
        // the window keeps the last 32 characters, which hold the seed key
        for (uint32_t hi = m; hi > 0 && ok; ) {
            const uint32_t lo = hi > PER ? hi - PER : 0;
            uint64_t cw;
            ok = WIDE ? fetch_codes_w(bytes + base + lo, hi - lo, s_lut, cw) : fetch_codes(bytes + base + lo, hi - lo, s_lut, cw);
            if (hi == m) { win = cw; win_lo = lo; }
            hi = lo;
        }
        if (ok) {
            if (WIDE) {                                   // radix-nletters key of the last L0 characters (L0 <= 14 < 16: inside the window)
                key = 0;
                for (uint32_t i = L0; i > 0; --i) key = key * ix.nletters + ((uint32_t)(win >> (4 * (m - L0 - win_lo + i - 1))) & 7u);
            } else
            key = (uint32_t)(win >> (2 * (m - L0 - win_lo))) & (uint32_t)((1ull << (2 * L0)) - 1);
            uint32_t a, n0; uint2 iv0 = make_uint2(1u, 0u);
            if (sv.first && L0 == sv.first_level) { const uint4 e = __ldg(sv.first + key); iv0 = make_uint2(e.x, e.y); a = e.z; n0 = e.w; }
            else { a = __ldg(s_off[L0] + key); n0 = __ldg(s_off[L0] + key + 1) - a; if (n0) iv0 = __ldg(s_iv[L0] + a); }
            if (n0 && m == L0 && n0 <= PLIST_N_MASK) {
                // pattern no longer than the seed depth: nothing is left to search.  A short key's frontier can hold thousands
                // of intervals, so it is neither copied nor staged: the count comes from the per-key table (or a sum at the deepest level) and the plist entry
                // points into the seed table.  The lane's own (unused) slot column carries both to the epilogue.
                direct = true;
                uint32_t dc = 0;
                if (s_cnt[L0]) dc = __ldg(s_cnt[L0] + key);
                else for (uint32_t j = 0; j < n0; ++j) { const uint2 v = __ldg(s_iv[L0] + a + j); dc += v.y - v.x + 1; }
                LIST(0) = make_uint2(a, PLIST_SEED | (L0 << 27) | n0);
                LIST(1) = make_uint2(dc, n0);
            }
            else if (n0 > cap) ovf = true;
            else { if (n0) LIST(0) = iv0; for (uint32_t j = 1; j < n0; ++j) LIST(j) = __ldg(s_iv[L0] + a + j); nF = n0; }
        } else m = 0;
    }
    const uint32_t steps = m - L0;
    const uint32_t mk = m < kc ? m : kc;
    uint32_t s = 0;
    for (; s < steps && nF && !ovf; ++s) {
        const uint32_t ci = m - L0 - s - 1;
        if (ci < win_lo) {                                                // the next 32 characters to the left
            const uint32_t lo = win_lo > PER ? win_lo - PER : 0;
            if (WIDE) fetch_codes_w(bytes + base + lo, win_lo - lo, s_lut, win); else fetch_codes(bytes + base + lo, win_lo - lo, s_lut, win);
            win_lo = lo;
        }
        const int c = WIDE ? (int)((win >> (4 * (ci - win_lo))) & 7u) : (int)((win >> (2 * (ci - win_lo))) & 3u);
        const uint32_t x0 = (c & 1) ? 0u : 0xffffffffu, x1 = (c & 2) ? 0u : 0xffffffffu;
        const uint32_t Cc = s_Cl[c];
        const bool beyond = STATS && (L0 + s >= kc);
        if (STATS && L0 + s == mk) sI += nF;
        uint32_t nout = 0, ndol = 0;
        // ---- pass 1: F' = LF_c(F) in place, and the set of pile-0 ranges reached through the '#' rows of F
        for (uint32_t j = 0; j < nF && !ovf; ++j) {
            const uint2 iv = LIST(j);
            uint32_t rca, rcb, rda, rdb, nblk;
            if (WIDE) lf_range_w(ix.occ, iv.x, iv.y + 1, c, rca, rcb, rda, rdb, nblk);
            else lf_range(ix.occ, iv.x, iv.y + 1, c, x0, x1, rca, rcb, rda, rdb, nblk);
            if (STATS) { sLF += 1; sBlk += nblk; if (beyond) { sS += 1; sE += rdb - rda; } }
            if (rcb > rca) {
                const uint32_t nlo = Cc + rca, nhi = Cc + rcb - 1;
                if (nout && LIST(nout - 1).y + 1 == nlo) LIST(nout - 1).y = nhi;
                else { LIST(nout) = make_uint2(nlo, nhi); ++nout; }
            }
            for (uint32_t e = rda; e < rdb; ++e) {
                const uint2 rec = __ldg(ix.link_rec + e);
                if (rec.x > rec.y) continue;                              // string of the first segment: no predecessor
                uint32_t l = rec.x, h = rec.y, i = 0;
                while (i < ndol && DOL(i).y + 1 < l) ++i;
                if (i < ndol && DOL(i).x <= h + 1) {                      // overlaps / touches D[i]: merge, then absorb followers
                    const uint2 d = DOL(i);
                    l = l < d.x ? l : d.x; h = h > d.y ? h : d.y;
                    uint32_t t = i + 1;
                    while (t < ndol && DOL(t).x <= h + 1) { const uint32_t y = DOL(t).y; h = h > y ? h : y; ++t; }
                    DOL(i) = make_uint2(l, h);
                    if (t > i + 1) { for (uint32_t u = t; u < ndol; ++u) DOL(i + 1 + (u - t)) = DOL(u); ndol -= t - (i + 1); }
                } else {                                                  // insert before i
                    if (nF + ndol + 1 >= cap) { ovf = true; break; }
                    for (uint32_t u = ndol; u > i; --u) DOL(u) = DOL(u - 1);
                    DOL(i) = make_uint2(l, h);
                    ++ndol;
                }
            }
        }
        if (ovf) break;
        // ---- pass 2: Dol' = LF_c(Dol), pile-0 rows (in place in the top region), then F <- Dol' ++ F'
        if (ndol) {
            uint32_t ndp = 0;
            for (uint32_t i = 0; i < ndol; ++i) {
                const uint2 iv = DOL(i);
                uint32_t rca, rcb, rda, rdb, nblk;
                if (WIDE) lf_range_w(ix.occ, iv.x, iv.y + 1, c, rca, rcb, rda, rdb, nblk);
            else lf_range(ix.occ, iv.x, iv.y + 1, c, x0, x1, rca, rcb, rda, rdb, nblk);
                if (STATS) { sLF += 1; sBlk += nblk; if (beyond) { sS += 1; sE += rdb - rda; } }
                if (rcb > rca) {
                    const uint32_t nlo = Cc + rca, nhi = Cc + rcb - 1;
                    if (ndp && DOL(ndp - 1).y + 1 == nlo) DOL(ndp - 1).y = nhi;
                    else { DOL(ndp) = make_uint2(nlo, nhi); ++ndp; }
                }
            }
            if (ndp) {
                if (nout + 2 * ndp > cap) { ovf = true; break; }
                for (uint32_t j = nout; j > 0; --j) LIST(j - 1 + ndp) = LIST(j - 1);
                for (uint32_t i = 0; i < ndp; ++i) LIST(i) = DOL(i);
                nout += ndp;
            }
        }
        nF = nout;
    }
    if (STATS && !ovf && live && m && L0 + s == mk) sI += direct ? LIST(1).y : nF;            // frontier left at suffix length min(m, kc)
    if (out.alive && live && !ovf) {
        // where the frontier died ("Pattern not found, interrupted in position", EDSBWTsearch.cpp:714-717)
        uint32_t la = 0;
        if (m == 0) la = 0;
        else if (nF || direct) la = m;
        else if (s > 0) la = L0 + s - 1;
        else for (uint32_t L = L0 - 1; L >= 1; --L) {                     // empty seed entry: longest shorter suffix that is not
            uint32_t kl = key >> (2 * (L0 - L));
            if (WIDE) { kl = key; for (uint32_t t = L; t < L0; ++t) kl /= ix.nletters; }
            if (__ldg(s_off[L] + kl + 1) > __ldg(s_off[L] + kl)) { la = L; break; }
        }
        out.alive[q] = la;
    }
    // ---- results: per-pattern count + the final interval list, staged contiguously per warp
    uint32_t cnt = 0;
    if (ovf) nF = 0;
    for (uint32_t j = 0; j < nF; ++j) { const uint2 iv = LIST(j); cnt += iv.y - iv.x + 1; }
    if (direct) cnt = LIST(1).x;
    const int lane = threadIdx.x & 31;
    uint32_t pre = nF;                                                  // inclusive warp scan
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, pre, o); if (lane >= o) pre += t; }
    const uint32_t total = __shfl_sync(0xffffffffu, pre, 31);
    uint32_t wbase = 0;
    if (lane == 31 && total) wbase = atomicAdd(out.cursor, total);
    wbase = __shfl_sync(0xffffffffu, wbase, 31);
    const bool nospace = total && (uint64_t)wbase + total > out.staging_cap;
    if (nospace && nF) { ovf = true; nF = 0; cnt = 0; }
    if (live) {
        const uint32_t my = wbase + pre - nF;
        for (uint32_t j = 0; j < nF; ++j) out.staging[my + j] = LIST(j);
        out.plist[q] = direct ? LIST(0) : make_uint2(my, nF);
        out.counts[q] = cnt;
        if (ovf) out.ovf_ids[atomicAdd(out.ovf_count, 1u)] = q;
    }
    if (STATS) {
        if (ovf) sI = sS = sE = sLF = sBlk = 0;          // the level-synchronous engine redoes (and recounts) this pattern
        const unsigned long long v[5] = {sI, sS, sE, sLF, sBlk};
        unsigned long long* dst[5] = {&st->I, &st->S, &st->E, &st->lf_intervals, &st->occ_blocks};
        #pragma unroll
        for (int i = 0; i < 5; ++i) {
            uint32_t x = warp_sum((uint32_t)v[i]);
            if (lane == 0 && x) atomicAdd(dst[i], (unsigned long long)x);
        }
    }
#undef LIST
#undef DOL
}

namespace {
template <bool STATS, int MINB, bool WIDE = false>
void launch_instance(unsigned grid, size_t smem, cudaStream_t s, const IndexView& ix, const SeedView& sv, uint32_t k, const uint8_t* d_bytes, const uint64_t* d_off,
                     const uint32_t* d_ids, uint32_t npat, const uint8_t* d_lut, uint32_t cap, const FusedOut& out, StatsDev* st, uint32_t kc) {
    CUDA_OK(cudaFuncSetAttribute(fused_search_kernel<STATS, MINB, WIDE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));      // per device: set at every launch
    fused_search_kernel<STATS, MINB, WIDE><<<grid, FUSED_THREADS, smem, s>>>(ix, sv, k, d_bytes, d_off, d_ids, npat, d_lut, cap, out, st, kc);
}
}  // namespace

void launch_fused_search(const IndexView& ix, const SeedView& sv, uint32_t k, const uint8_t* d_bytes, const uint64_t* d_off,
                         const uint32_t* d_ids, uint32_t npat, const uint8_t* d_lut, uint32_t cap, const FusedOut& out, StatsDev* st, uint32_t kc, cudaStream_t s) {
    if (npat == 0) return;
    const size_t smem = (size_t)cap * FUSED_THREADS * sizeof(uint2);
    const unsigned grid = (npat + FUSED_THREADS - 1) / FUSED_THREADS;
#define INSTANCE(STATS, MINB) launch_instance<STATS, MINB>(grid, smem, s, ix, sv, k, d_bytes, d_off, d_ids, npat, d_lut, cap, out, st, kc)
    if (ix.wide) {                  // 24 registers of Occ block per interval: 8 CTAs/SM (64 registers)
        if (st) launch_instance<true, 8, true>(grid, smem, s, ix, sv, k, d_bytes, d_off, d_ids, npat, d_lut, cap, out, st, kc);
        else launch_instance<false, 8, true>(grid, smem, s, ix, sv, k, d_bytes, d_off, d_ids, npat, d_lut, cap, out, st, kc);
    }
    else if (st) INSTANCE(true, 10);
    else if (cap <= 14) INSTANCE(false, 12);      // 12 CTAs/SM: 40 registers (40 B of spill in a cold path), at most 14 KB of frontier each; 3.27 -> 3.09 ms
    else INSTANCE(false, 10);
#undef INSTANCE
    CUDA_OK(cudaGetLastError());
}

}  // namespace edsbwt
