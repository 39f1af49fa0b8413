// MOVE_EDSBWTSearch variant: LF through an a-balanced move table instead of blocked rank.
//
//   reference                                        here
//   build_MLF.cpp:53-164 (Move-r move_data_structure_l_ from the (i, LF(i)) pairs of X_runs.txt,
//     L_ from X_runs.aux; da_to_everything.cpp:62-109, 435-463)            build_move_table()   (K6, on the GPU, from the Occ blocks)
//   MOVE_EDSBWT::retrieve_MLF / findInputInterval    MOVE_EDSBWTSearch.cpp:178-218     row_run0[] (run_of_row over pile 0)
//   backward_search_step / updateSingleInterval      :376-510                          move_lf() inside move_search_kernel
//   link / dollars_in_interval / preceding_dollars_finder   :512-625                   '#' rank over run heads + link_rec gather
//   locate by M_LF.move walk                         :326-369                          move_locate_kernel
//
// The reference's <base>_MLF.aux is Move-r's private serialisation (submodule absent from the reference checkout,
// SURVEY 8(c)): it cannot be read, so the table is rebuilt from the pile files at load.  Its interval boundaries may
// differ from Move-r's (any a-balanced refinement of the runs is a valid move table); search results do not depend on
// them and are pinned by equality with the EDSBWTsearch occurrence sets (a10 vs a13 emit the same tuples).
#include <cub/cub.cuh>

#include <algorithm>

#include "internal.cuh"

namespace edsbwt {

namespace {

inline unsigned grid_for(size_t n, unsigned block) { return (unsigned)((n + block - 1) / block); }

// ---------------------------------------------------------------------------------------------------
// table construction
// ---------------------------------------------------------------------------------------------------
// run heads: row 0, every row whose symbol differs from the previous row's, and every '#' row
// (da_to_everything.cpp:440-442: "every end-marker is its own run")
struct IsRunHead {
    const uint4* occ;
    __device__ __forceinline__ bool operator()(const uint32_t& row) const {
        if (row == 0) return true;
        uint32_t s = occ_symbol(occ, row);
        return s == 0 || s != occ_symbol(occ, row - 1);
    }
};

// LF of an arbitrary row (da_to_everything.cpp:78-103): letters through Occ, the k-th '#' row -> row k of pile 0
__device__ __forceinline__ uint32_t lf_row(const IndexView& ix, uint32_t row, uint32_t& sym) {
    sym = occ_symbol(ix.occ, row);
    if (sym == 0) return occ_rank_eof(ix.occ, row);
    return ix.Cl[sym - 1] + occ_rank1(ix.occ, row, (int)sym - 1);
}

// coarse[b] = number of starts < 128*b (lower bound), b = 0 .. nb-1
__global__ void coarse_kernel(const uint32_t* __restrict__ P, uint32_t r, uint32_t nb, uint32_t* __restrict__ coarse) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb) return;
    uint64_t v = (uint64_t)b * OCC_ROWS;
    uint32_t lo = 0, hi = r;
    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if ((uint64_t)__ldg(P + mid) < v) lo = mid + 1; else hi = mid; }
    coarse[b] = lo;
}
// number of starts < v (v <= n), using the coarse table: at most 8 probes
__device__ __forceinline__ uint32_t starts_below(const uint32_t* __restrict__ P, const uint32_t* __restrict__ coarse, uint32_t v) {
    uint32_t b = v >> 7;
    uint32_t lo = __ldg(coarse + b), hi = __ldg(coarse + b + 1);
    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (__ldg(P + mid) < v) lo = mid + 1; else hi = mid; }
    return lo;
}

// one balancing round: interval i = [P[i], P[i+1]) maps onto [q, q+d); if that output interval holds cnt >= 2a input
// starts it is cut so that every piece holds a of them (the last one a + cnt % a < 2a)
__global__ void heavy_count_kernel(IndexView ix, const uint32_t* __restrict__ P, const uint32_t* __restrict__ coarse, uint32_t r, uint32_t a,
                                   uint32_t* __restrict__ nsplit) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= r) return;
    uint32_t p = P[i], d = (i + 1 < r ? P[i + 1] : ix.n) - p, sym;
    uint32_t ns = 0;
    if (d >= 2 * a) {                                    // an interval of d rows cannot hold more than d starts
        uint32_t q = lf_row(ix, p, sym);
        uint32_t lo = starts_below(P, coarse, q), hi = starts_below(P, coarse, q + d);
        uint32_t cnt = hi - lo;
        if (cnt >= 2 * a) ns = cnt / a - 1;
    }
    nsplit[i] = ns;
}
__global__ void heavy_emit_kernel(IndexView ix, const uint32_t* __restrict__ P, const uint32_t* __restrict__ coarse, uint32_t r, uint32_t a,
                                  const uint32_t* __restrict__ nsplit, const uint32_t* __restrict__ ofs, uint32_t* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= r) return;
    uint32_t ns = nsplit[i];
    if (!ns) return;
    uint32_t p = P[i], sym;
    uint32_t q = lf_row(ix, p, sym);
    uint32_t lo = starts_below(P, coarse, q);
    for (uint32_t t = 1; t <= ns; ++t) out[ofs[i] + t - 1] = p + (P[lo + t * a] - q);
}

__global__ void table_kernel(IndexView ix, const uint32_t* __restrict__ P, const uint32_t* __restrict__ coarse, uint32_t r,
                             uint4* __restrict__ T, uint8_t* __restrict__ Lbytes, Lut lutinv) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > r) return;
    if (i == r) { T[r] = make_uint4(ix.n, ix.n, r, 0xffu); return; }
    uint32_t p = P[i], sym;
    uint32_t q = lf_row(ix, p, sym);
    uint32_t idx = starts_below(P, coarse, q + 1) - 1;          // last start <= q
    T[i] = make_uint4(p, q, idx, sym);
    Lbytes[i] = lutinv.code[sym];                                // back to the alphabet byte for the Occ packer
}
__global__ void sel_kernel(const uint4* __restrict__ T, const uint4* __restrict__ occL, uint32_t r, MoveView mv, uint32_t* __restrict__ sel) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= r) return;
    uint32_t sym = T[i].w;
    if (sym == 0) return;
    sel[mv.selbase[sym - 1] + occ_rank1(occL, i, (int)sym - 1)] = i;
}
__global__ void row_run0_kernel(const uint32_t* __restrict__ P, const uint32_t* __restrict__ coarse, uint32_t n_text, uint32_t* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_text) out[i] = starts_below(P, coarse, i + 1) - 1;
}

}  // namespace

void build_move_table(Index& ix) {
    if (ix.move) return;
    if (ix.wide) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "the move table packs 2-bit letter codes: not built for an alphabet of %u letters", ix.host.sigma - 1);
    CUDA_OK(cudaSetDevice(ix.device));
    // a table written earlier by build_MLF / edsbwt_move_table_save for these very index files (the reference's
    // retrieve_MLF, MOVE_EDSBWTSearch.cpp:186-192): stale or foreign files are ignored and the table is built
    if (!ix.base.empty() && move_cache_load(ix, ix.base + "_MLF.edsbwt")) return;
    cudaStream_t s = ix.stream;
    const IndexView& v = ix.view;
    const uint32_t n = v.n, a = ix.move_a < 2 ? 2 : ix.move_a;
    MoveTable* mt = new MoveTable; mt->a = a;
    struct Guard { MoveTable* p; ~Guard() { delete p; } } guard{mt};
    DevBuf<uint8_t> tmp;
    DevBuf<uint32_t> P, P2, nsplit, ofs, extra, d_num;
    d_num.ensure(4);
    // ---- run heads
    P.ensure((size_t)n + 1);
    {
        cub::CountingInputIterator<uint32_t> rows(0);
        size_t tb = 0;
        CUDA_OK(cub::DeviceSelect::If(nullptr, tb, rows, P.p, d_num.p, (int)n, IsRunHead{v.occ}, s));
        tmp.ensure(tb);
        CUDA_OK(cub::DeviceSelect::If(tmp.p, tb, rows, P.p, d_num.p, (int)n, IsRunHead{v.occ}, s));
    }
    uint32_t r = 0;
    CUDA_OK(cudaMemcpyAsync(&r, d_num.p, 4, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    mt->r_runs = r;
    const uint32_t nb = n / OCC_ROWS + 3;
    mt->coarse.ensure(nb);
    // ---- balancing rounds
    for (uint32_t round = 0;; ++round) {
        if (round > 64) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "move table balancing did not converge");
        coarse_kernel<<<grid_for(nb, 256), 256, 0, s>>>(P.p, r, nb, mt->coarse.p);
        nsplit.ensure((size_t)r + 1); ofs.ensure((size_t)r + 1);
        CUDA_OK(cudaMemsetAsync(nsplit.p + r, 0, 4, s));
        heavy_count_kernel<<<grid_for(r, 256), 256, 0, s>>>(v, P.p, mt->coarse.p, r, a, nsplit.p);
        size_t tb = 0;
        CUDA_OK(cub::DeviceScan::ExclusiveSum(nullptr, tb, nsplit.p, ofs.p, (int)r + 1, s));
        tmp.ensure(tb);
        CUDA_OK(cub::DeviceScan::ExclusiveSum(tmp.p, tb, nsplit.p, ofs.p, (int)r + 1, s));
        uint32_t nnew = 0;
        CUDA_OK(cudaMemcpyAsync(&nnew, ofs.p + r, 4, cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));
        mt->rounds = round;
        if (nnew == 0) break;
        if ((uint64_t)r + nnew > n) EDS_FAIL(EDSBWT_E_FORMAT, "move table: more intervals than rows");
        // new starts are appended and the whole column re-sorted (they are few; one radix sort per round)
        heavy_emit_kernel<<<grid_for(r, 256), 256, 0, s>>>(v, P.p, mt->coarse.p, r, a, nsplit.p, ofs.p, P.p + r);
        P2.ensure((size_t)r + nnew + 1);
        tb = 0;
        CUDA_OK(cub::DeviceRadixSort::SortKeys(nullptr, tb, P.p, P2.p, (int)(r + nnew), 0, 32, s));
        tmp.ensure(tb);
        CUDA_OK(cub::DeviceRadixSort::SortKeys(tmp.p, tb, P.p, P2.p, (int)(r + nnew), 0, 32, s));
        std::swap(P.p, P2.p); std::swap(P.cap, P2.cap);
        P.ensure((size_t)n + 1, s, true, (size_t)r + nnew);       // room for the next round's appends
        r += nnew;
    }
    mt->r = r;
    // ---- table entries, L' structure, select arrays, pile-0 row -> interval
    mt->T.ensure((size_t)r + 2);
    DevBuf<uint8_t> Lbytes; Lbytes.ensure((size_t)r + 64);
    Lut lutinv; memset(lutinv.code, 0, sizeof lutinv.code);
    lutinv.code[0] = ix.host.alphabet[0];
    for (uint32_t c = 1; c < ix.host.sigma; ++c) lutinv.code[c] = ix.host.alphabet[c];
    table_kernel<<<grid_for((size_t)r + 1, 256), 256, 0, s>>>(v, P.p, mt->coarse.p, r, mt->T.p, Lbytes.p, lutinv);
    uint32_t tot[4]; unsigned int bad = 0;
    build_occ_blocks(Lbytes.p, r, make_lut(ix.host), mt->occL, tmp, s, tot, &bad);
    if (bad) EDS_FAIL(EDSBWT_E_FORMAT, "move table: run symbol outside the alphabet");
    MoveView& mv = mt->view;
    mv.T = mt->T.p; mv.r = r; mv.occL = mt->occL.p; mv.coarse = mt->coarse.p;
    mv.selbase[0] = 0;
    for (int c = 0; c < 4; ++c) mv.selbase[c + 1] = mv.selbase[c] + tot[c];
    mt->sel.ensure((size_t)mv.selbase[4] + 1);
    sel_kernel<<<grid_for(r, 256), 256, 0, s>>>(mt->T.p, mt->occL.p, r, mv, mt->sel.p);
    mv.sel = mt->sel.p;
    mt->row_run0.ensure((size_t)v.n_text + 1);
    row_run0_kernel<<<grid_for(v.n_text, 256), 256, 0, s>>>(P.p, mt->coarse.p, v.n_text, mt->row_run0.p);
    mv.row_run0 = mt->row_run0.p;
    CUDA_OK(cudaStreamSynchronize(s));
    CUDA_OK(cudaGetLastError());
    guard.p = nullptr;
    ix.move = mt;
}

// ---------------------------------------------------------------------------------------------------
// search
// ---------------------------------------------------------------------------------------------------
namespace {

constexpr int MOVE_THREADS = 128;

// interval containing `row` (binary search over the p column, bracketed by the coarse table)
__device__ __forceinline__ uint32_t run_of_row(const MoveView& mv, uint32_t row) {
    uint32_t b = (row + 1) >> 7;
    uint32_t lo = __ldg(mv.coarse + b), hi = __ldg(mv.coarse + b + 1);      // number of starts < row + 1 lies in [lo, hi]
    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (__ldg(&mv.T[mid].x) < row + 1) lo = mid + 1; else hi = mid; }
    return lo - 1;
}
// M_LF.move (Move-r): (row, interval) -> (LF(row), interval containing it); < 2a iterations of the loop (a-balancedness)
__device__ __forceinline__ void move_step(const MoveView& mv, uint32_t& row, uint32_t& run) {
    const uint4 e = __ldg(mv.T + run);
    row = e.y + (row - e.x);
    run = e.z;
    while (row >= __ldg(&mv.T[run + 1].x)) ++run;
}
// the two rank pairs of an interval, (rank_c, rank_#) of L'[0, posA) and L'[0, posB), posA <= posB: the run heads b' and
// e' + 1 nearly always share one 64-byte block of the L' structure, which is then fetched once (two 256-bit loads)
__device__ __forceinline__ void rank2_pair(const uint4* __restrict__ occL, uint32_t posA, uint32_t posB, int c,
                                           uint32_t& kb, uint32_t& l, uint32_t& ke1, uint32_t& u) {
    const uint32_t x0 = (c & 1) ? 0u : 0xffffffffu, x1 = (c & 2) ? 0u : 0xffffffffu;
    const uint32_t blkA = posA >> 7, blkB = posB >> 7;
    Blk B;
    load_blk(occL, blkA, B);
    uint32_t ka, kk;
    rank_in_blk(B, posA & 127u, c, x0, x1, kb, ka);
    l = blkA * OCC_ROWS - letters_before(B) + ka;
    if (blkB != blkA) load_blk(occL, blkB, B);
    rank_in_blk(B, posB & 127u, c, x0, x1, ke1, kk);
    u = blkB * OCC_ROWS - letters_before(B) + kk;
}

// One backward step of one interval by move-LF (updateSingleInterval, MOVE_EDSBWTSearch.cpp:424-510).
// in: (b, e, b_, e_) and the ranks kb = rank_c(L', b_), ke1 = rank_c(L', e_ + 1).  false: no row of the interval holds c.
__device__ __forceinline__ bool move_lf(const MoveView& mv, int c, uint32_t kb, uint32_t ke1, uint32_t& b, uint32_t& e, uint32_t& b_, uint32_t& e_) {
    if (ke1 <= kb) return false;                                   // no run of c among runs b_ .. e_
    const uint32_t sy = (uint32_t)c + 1;
    if (__ldg(&mv.T[b_].w) != sy) { b_ = __ldg(mv.sel + mv.selbase[c] + kb); b = __ldg(&mv.T[b_].x); }             // :436-442
    if (__ldg(&mv.T[e_].w) != sy) { e_ = __ldg(mv.sel + mv.selbase[c] + ke1 - 1); e = __ldg(&mv.T[e_ + 1].x) - 1; }   // :444-447
    if (b > e) return false;
    if (b_ == e_) {                                                // :455-482: one move, e follows b
        const uint32_t diff = e - b;
        move_step(mv, b, b_);
        e = b + diff; e_ = b_;
        while (e >= __ldg(&mv.T[e_ + 1].x)) ++e_;
    } else { move_step(mv, b, b_); move_step(mv, e, e_); }         // :484-494
    return true;
}

}  // namespace

template <int MINB>
__global__ void __launch_bounds__(MOVE_THREADS, MINB)
move_search_kernel(const __grid_constant__ IndexView ix, const __grid_constant__ MoveView mv, const __grid_constant__ SeedView sv, uint32_t k, const uint8_t* __restrict__ bytes, const uint64_t* __restrict__ poff,
                   const uint32_t* __restrict__ ids, uint32_t npat, const uint8_t* __restrict__ lut, uint32_t cap, MoveOut out) {
    extern __shared__ uint4 m_list[];                 // [cap][MOVE_THREADS] of (b, e, b', e')
    __shared__ uint8_t s_lut[256];
    for (int i = threadIdx.x; i < 256; i += MOVE_THREADS) s_lut[i] = lut[i];
    __syncthreads();
#define LIST(j) m_list[(j) * MOVE_THREADS + threadIdx.x]
#define DOL(i) m_list[(cap - 1 - (i)) * MOVE_THREADS + threadIdx.x]
    const uint32_t tq = blockIdx.x * MOVE_THREADS + threadIdx.x;
    const bool live = tq < npat;
    const uint32_t q = live ? (ids ? ids[tq] : tq) : 0u;
    uint32_t nF = 0, m = 0, L0 = 0, win_lo = 0;
    uint64_t base = 0, win = 0;                       // win: 2-bit codes of pattern characters [win_lo, win_lo + 32)
    bool ovf = false;
    if (live) {
        base = poff[q];
        uint64_t len64 = poff[q + 1] - base;
        m = len64 > 0xffffffffull ? 0u : (uint32_t)len64;
        L0 = m < k ? m : k;
        bool ok = m > 0;
        for (uint32_t hi = m; hi > 0 && ok; ) {      // alphabet check of the whole pattern; the window keeps its last 32 characters
            const uint32_t lo = hi > 32 ? hi - 32 : 0;
            uint64_t cw;
            ok = fetch_codes(bytes + base + lo, hi - lo, s_lut, cw);
            if (hi == m) { win = cw; win_lo = lo; }
            hi = lo;
        }
        if (ok) {
            const uint32_t key = (uint32_t)(win >> (2 * (m - L0 - win_lo))) & (uint32_t)((1ull << (2 * L0)) - 1);
            // the first min(m,k) steps come from the seed table (rows); the companion input intervals are looked up
            uint32_t a = __ldg(sv.off[L0] + key), n0 = __ldg(sv.off[L0] + key + 1) - a;
            if (n0 > cap) ovf = true;
            else { for (uint32_t j = 0; j < n0; ++j) { uint2 iv = __ldg(sv.iv[L0] + a + j); LIST(j) = make_uint4(iv.x, iv.y, run_of_row(mv, iv.x), run_of_row(mv, iv.y)); } nF = n0; }
        } else m = 0;
    }
    const uint32_t steps = m - L0;
    for (uint32_t s = 0; s < steps && nF && !ovf; ++s) {
        const uint32_t ci = m - L0 - s - 1;
        if (ci < win_lo) {
            const uint32_t lo = win_lo > 32 ? win_lo - 32 : 0;
            fetch_codes(bytes + base + lo, win_lo - lo, s_lut, win);
            win_lo = lo;
        }
        const int c = (int)((win >> (2 * (ci - win_lo))) & 3u);
        uint32_t nout = 0, ndol = 0;
        // ---- pass 1: link (the '#' runs among runs b'..e' of every interval) + F' = LF_c(F) in place
        for (uint32_t j = 0; j < nF && !ovf; ++j) {
            uint4 iv = LIST(j);
            uint32_t kb, l, ke1, u;
            rank2_pair(mv.occL, iv.z, iv.w + 1, c, kb, l, ke1, u);  // dollars_in_interval: _RS_L_.rank('#', b') .. rank('#', e'+1), :609-610
            if (move_lf(mv, c, kb, ke1, iv.x, iv.y, iv.z, iv.w)) {
                if (nout && LIST(nout - 1).y + 1 == iv.x) { uint4 t = LIST(nout - 1); t.y = iv.y; t.w = iv.w; LIST(nout - 1) = t; }
                else { LIST(nout) = iv; ++nout; }
            }
            for (uint32_t e = l; e < u; ++e) {
                const uint2 rec = __ldg(ix.link_rec + e);
                if (rec.x > rec.y) continue;
                uint32_t lo = rec.x, hi = rec.y, i = 0;
                while (i < ndol && DOL(i).y + 1 < lo) ++i;
                if (i < ndol && DOL(i).x <= hi + 1) {
                    const uint4 d = DOL(i);
                    lo = lo < d.x ? lo : d.x; hi = hi > d.y ? hi : d.y;
                    uint32_t t = i + 1;
                    while (t < ndol && DOL(t).x <= hi + 1) { const uint32_t y = DOL(t).y; hi = hi > y ? hi : y; ++t; }
                    DOL(i) = make_uint4(lo, hi, 0, 0);
                    if (t > i + 1) { for (uint32_t w = t; w < ndol; ++w) DOL(i + 1 + (w - t)) = DOL(w); ndol -= t - (i + 1); }
                } else {
                    if (nF + ndol + 1 >= cap) { ovf = true; break; }
                    for (uint32_t w = ndol; w > i; --w) DOL(w) = DOL(w - 1);
                    DOL(i) = make_uint4(lo, hi, 0, 0);
                    ++ndol;
                }
            }
        }
        if (ovf) break;
        // ---- pass 2: Dol' = LF_c(Dol) (pile-0 ranges with their input intervals, preceding_dollars_finder :598-601), F <- Dol' ++ F'
        if (ndol) {
            uint32_t ndp = 0;
            for (uint32_t i = 0; i < ndol; ++i) {
                uint4 iv = DOL(i);
                iv.z = __ldg(mv.row_run0 + iv.x); iv.w = __ldg(mv.row_run0 + iv.y);
                uint32_t kb, l, ke1, u;
                rank2_pair(mv.occL, iv.z, iv.w + 1, c, kb, l, ke1, u);
                if (move_lf(mv, c, kb, ke1, iv.x, iv.y, iv.z, iv.w)) {
                    if (ndp && DOL(ndp - 1).y + 1 == iv.x) { uint4 t = DOL(ndp - 1); t.y = iv.y; t.w = iv.w; DOL(ndp - 1) = t; }
                    else { DOL(ndp) = iv; ++ndp; }
                }
            }
            if (ndp) {
                if (nout + 2 * ndp > cap) { ovf = true; break; }
                // the reference merges across the whole concatenated list (:309-324), i.e. also across the seam
                uint32_t shift = ndp;
                if (nout && DOL(ndp - 1).y + 1 == LIST(0).x) { uint4 t = DOL(ndp - 1); t.y = LIST(0).y; t.w = LIST(0).w; DOL(ndp - 1) = t;
                                                               for (uint32_t j = 1; j < nout; ++j) LIST(j - 1) = LIST(j); --nout; }
                for (uint32_t j = nout; j > 0; --j) LIST(j - 1 + shift) = LIST(j - 1);
                for (uint32_t i = 0; i < ndp; ++i) LIST(i) = DOL(i);
                nout += ndp;
            }
        }
        nF = nout;
    }
    // ---- results (same staging protocol as fused_search_kernel)
    uint32_t cnt = 0;
    if (ovf) nF = 0;
    for (uint32_t j = 0; j < nF; ++j) { const uint4 iv = LIST(j); cnt += iv.y - iv.x + 1; }
    const int lane = threadIdx.x & 31;
    uint32_t pre = nF;
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, pre, o); if (lane >= o) pre += t; }
    const uint32_t total = __shfl_sync(0xffffffffu, pre, 31);
    uint32_t wbase = 0;
    if (lane == 31 && total) wbase = atomicAdd(out.cursor, total);
    wbase = __shfl_sync(0xffffffffu, wbase, 31);
    const bool nospace = total && (uint64_t)wbase + total > out.staging_cap;
    if (nospace && nF) { ovf = true; nF = 0; cnt = 0; }
    if (live) {
        const uint32_t my = wbase + pre - nF;
        for (uint32_t j = 0; j < nF; ++j) { const uint4 iv = LIST(j); out.staging[my + j] = make_uint2(iv.x, iv.y); }
        out.plist[q] = make_uint2(my, nF);
        out.counts[q] = cnt;
        if (ovf) out.ovf_ids[atomicAdd(out.ovf_count, 1u)] = q;
    }
#undef LIST
#undef DOL
}

// one thread per occurrence row: walk M_LF.move until the run symbol is '#', counting the steps = offset in the string
// (MOVE_EDSBWTSearch.cpp:340-368); then k = rank('#', run) over the run heads, w = EOF_ID_Copy[k], segment / index in segment
__global__ void move_locate_kernel(const __grid_constant__ IndexView ix, const __grid_constant__ MoveView mv, const uint2* __restrict__ plist, const uint32_t* __restrict__ occ_ofs,
                                   const uint2* __restrict__ staging, uint32_t npat, uint64_t nocc, uint32_t pat_base,
                                   edsbwt_occ* __restrict__ out, StatsDev* __restrict__ st) {
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t steps = 0;
    if (t < nocc) {
        uint32_t lo = 0, hi = npat + 1;                              // first index with occ_ofs[idx] > t
        while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (__ldg(occ_ofs + mid) <= (uint32_t)t) lo = mid + 1; else hi = mid; }
        uint32_t q = lo - 1;
        uint32_t r = (uint32_t)t - __ldg(occ_ofs + q);
        uint2 pl = __ldg(plist + q);
        uint32_t row = 0;
        for (uint32_t j = 0; j < pl.y; ++j) {
            uint2 iv = __ldg(staging + pl.x + j);
            uint32_t len = iv.y - iv.x + 1;
            if (r < len) { row = iv.x + r; break; }
            r -= len;
        }
        uint32_t run = run_of_row(mv, row), off = 0;
        while (__ldg(&mv.T[run].w) != 0) { move_step(mv, row, run); ++off; }
        uint32_t w = __ldg(ix.eof_flat + occ_rank_eof(mv.occL, run));
        uint2 si = __ldg(ix.str_info + w);
        edsbwt_occ o; o.pat = q + pat_base; o.str = w; o.seg = si.x; o.idx_in_seg = si.y; o.off = off;
        out[t] = o;
        steps = off + 1;
    }
    if (st) {
        uint32_t v = steps, one = steps ? 1u : 0u;
        for (int o = 16; o; o >>= 1) { v += __shfl_xor_sync(0xffffffffu, v, o); one += __shfl_xor_sync(0xffffffffu, one, o); }
        if ((threadIdx.x & 31) == 0 && v) { atomicAdd(&st->L, (unsigned long long)v); atomicAdd(&st->O, (unsigned long long)one); }
    }
}

void launch_move_search(const IndexView& ix, const MoveView& mv, const SeedView& sv, uint32_t k, const uint8_t* d_bytes, const uint64_t* d_off,
                        const uint32_t* d_ids, uint32_t npat, const uint8_t* d_lut, uint32_t cap, const MoveOut& out, cudaStream_t s) {
    if (npat == 0) return;
    const size_t smem = (size_t)cap * MOVE_THREADS * sizeof(uint4);
    if (cap <= 14) {                   // first pass: 64 registers, 8 CTAs per SM (cfg3: 2.07 -> 1.56 ms against 4 CTAs at 72 registers)
        CUDA_OK(cudaFuncSetAttribute(move_search_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        move_search_kernel<8><<<(npat + MOVE_THREADS - 1) / MOVE_THREADS, MOVE_THREADS, smem, s>>>(ix, mv, sv, k, d_bytes, d_off, d_ids, npat, d_lut, cap, out);
    } else {
        CUDA_OK(cudaFuncSetAttribute(move_search_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        move_search_kernel<4><<<(npat + MOVE_THREADS - 1) / MOVE_THREADS, MOVE_THREADS, smem, s>>>(ix, mv, sv, k, d_bytes, d_off, d_ids, npat, d_lut, cap, out);
    }
    CUDA_OK(cudaGetLastError());
}
void launch_move_locate(const IndexView& ix, const MoveView& mv, const uint2* plist, const uint32_t* occ_ofs, const uint2* staging,
                        uint32_t npat, uint64_t nocc, uint32_t pat_base, edsbwt_occ* out, StatsDev* st, cudaStream_t s) {
    if (nocc == 0) return;
    move_locate_kernel<<<grid_for(nocc, 256), 256, 0, s>>>(ix, mv, plist, occ_ofs, staging, npat, nocc, pat_base, out, st);
    CUDA_OK(cudaGetLastError());
}

}  // namespace edsbwt
