// The hot kernel (round 2): persistent warps, one ITEM per lane per sweep, lanes refilled with new patterns.
//
//   reference per-pattern loop     EDSBWT::backwardSearch   EDSBWTsearch.cpp:619-772   (the loop :660-722 is what a lane runs)
//   LF of every frontier interval  updateIntervals / updateSingleInterval / rankManySymbols / MergeAndRemove   :841-1209
//   segment jump                   link / dollars_in_interval / preceding_dollars_finder   :1212-1307
//
// The step-loop kernel of round 1 (fused.cu) gave every pattern a thread and walked its frontier in a per-thread loop:
// ncu counted 14.9 of 32 lanes active per instruction (patterns that died early, frontiers of different length, one lane
// in twenty waiting alone for a link record).  Here a warp is a pool of 32 independent state machines.  In every sweep
// each lane (1) ISSUES the one memory fetch its pattern needs next -- a 64-byte Occ block (two 256-bit loads), a seed
// entry, a link record, a further seed interval or the next 32 pattern characters -- and only then (2) CONSUMES it and
// moves its state on.  All 32 fetches of a sweep are in flight together whatever they are, so a link record or a seed
// lookup costs its lane one sweep instead of stalling the warp, and a lane whose pattern is finished takes the next
// pattern from the warp's pool (32 patterns fetched coalesced, refilled from a global cursor: the warp-cooperative part)
// in the very next sweep.  Semantics are those of the step loop, statement for statement: same frontier lists
// (Dol' ++ F'), same merge rule, same overflow conditions, same byte-model counters.
//
// Patterns arrive as 2-bit codes (PackedPats, internal.cuh): produced by pack_patterns_kernel below from the caller's
// bytes, or by the caller itself (edsbwt_search_into_packed).
#include "internal.cuh"

namespace edsbwt {

// ---------------------------------------------------------------------------------------------------
// bytes -> 2-bit words.  Pattern q of m characters occupies ceil(m/32) words; word w holds characters
// [max(0, m - 32(w+1)), m - 32w) with the LAST of them in bits 62..63 (a short head is padded with zero bits below):
// the backward search consumes word 0 from the top down, the seed key is its top min(m, k) characters.
// ---------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
pack_patterns_kernel(const uint8_t* __restrict__ bytes, const uint64_t* __restrict__ poff, uint32_t fixed_len, uint64_t fixed_base, uint32_t stride,
                     uint32_t npat, const uint8_t* __restrict__ lut, uint64_t* __restrict__ words, uint64_t nwords, uint32_t* __restrict__ invalid,
                     uint32_t* __restrict__ err) {
    __shared__ uint8_t s_lut[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) s_lut[i] = lut[i];
    __syncthreads();
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    bool bad = false;
    if (q < npat) {
        uint64_t base, wb; uint32_t m;
        if (poff) {
            base = __ldg(poff + q);
            const uint64_t len64 = __ldg(poff + q + 1) - base;
            m = len64 > 0xffffffffull ? 0u : (uint32_t)len64;
            wb = ((base - __ldg(poff)) >> 5) + q;
        } else { base = fixed_base + (uint64_t)q * fixed_len; m = fixed_len; wb = (uint64_t)q * stride; }
        if (wb + (m + 31) / 32 > nwords) { *err = 1u; bad = true; m = 0; }      // the caller's total_bytes does not cover the offsets: nothing is written, the call fails
        for (uint32_t hi = m, w = 0; hi > 0; ++w) {
            const uint32_t lo = hi > 32 ? hi - 32 : 0, n = hi - lo;
            uint64_t cw;
            if (!fetch_codes(bytes + base + lo, n, s_lut, cw)) bad = true;
            words[wb + w] = cw << (2 * (32 - n));
            hi = lo;
        }
    }
    const uint32_t bal = __ballot_sync(0xffffffffu, bad);       // q >> 5 is the same for the whole warp
    if ((threadIdx.x & 31) == 0 && q < npat) invalid[q >> 5] = bal;
}

void launch_pack_patterns(const uint8_t* d_bytes, const uint64_t* d_off, uint32_t fixed_len, uint64_t fixed_base, uint32_t stride, uint32_t npat,
                          const uint8_t* d_lut, uint64_t* words, uint64_t nwords, uint32_t* invalid, uint32_t* err, cudaStream_t s) {
    if (npat == 0) return;
    pack_patterns_kernel<<<(npat + 127) / 128, 128, 0, s>>>(d_bytes, d_off, fixed_len, fixed_base, stride, npat, d_lut, words, nwords, invalid, err);
    CUDA_OK(cudaGetLastError());
}

// words -> characters of the patterns listed in ids (nullptr: all), pattern q at bytes[q * fixed_len ..): the rare patterns
// that outgrow the item kernel's frontier go to the step-loop kernel / the level-synchronous engine, which read characters.
// Only needed when the caller handed in packed patterns, i.e. no bytes exist on the device.  A pattern flagged invalid
// gets the end-marker byte, which no tier accepts.  letters: the (up to) four alphabet bytes, letter c in bits 8c..8c+7.
__global__ void unpack_patterns_kernel(const uint64_t* __restrict__ words, uint32_t fixed_len, uint32_t stride, const uint32_t* __restrict__ ids,
                                       uint32_t nsub, uint32_t letters, uint32_t bad_byte, const uint32_t* __restrict__ invalid, uint8_t* __restrict__ bytes) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nsub) return;
    const uint32_t q = ids ? ids[t] : t;
    const bool inv = invalid && ((__ldg(invalid + (q >> 5)) >> (q & 31u)) & 1u);
    for (uint32_t i = 0; i < fixed_len; ++i) {
        const uint32_t back = fixed_len - 1 - i;             // characters from the end: word back / 32, position 31 - back % 32
        const uint64_t w = __ldg(words + (uint64_t)q * stride + (back >> 5));
        const uint32_t c = (uint32_t)(w >> (2 * (31 - (back & 31u)))) & 3u;
        bytes[(uint64_t)q * fixed_len + i] = inv ? (uint8_t)bad_byte : (uint8_t)(letters >> (8 * c));
    }
}
void launch_unpack_patterns(const uint64_t* words, uint32_t fixed_len, uint32_t stride, const uint32_t* ids, uint32_t nsub, uint64_t letters,
                            const uint32_t* invalid, uint8_t* bytes, cudaStream_t s) {
    if (nsub == 0) return;
    unpack_patterns_kernel<<<(nsub + 127) / 128, 128, 0, s>>>(words, fixed_len, stride, ids, nsub, (uint32_t)(letters >> 8), (uint32_t)(letters & 255u), invalid, bytes);
    CUDA_OK(cudaGetLastError());
}

namespace {

enum : uint32_t {
    K_IDLE = 0,      // no pattern: takes one from the warp's pool at the top of the next sweep
    K_FIN,           // pattern finished: its results are written, then the lane takes the next pattern (same place)
    K_OCCF,          // fetch: Occ block of F[j]           (pass 1: F' = LF_c(F), '#' rows -> link)
    K_OCCD,          // fetch: Occ block of Dol[j]         (pass 2: Dol' = LF_c(Dol))
    K_LINK,          // fetch (8 B): link record of the next '#' row inside F[j]
    K_SEEDIV,        // fetch (8 B): a further interval of the seed entry
    K_WORD,          // fetch (8 B): the next 32 pattern characters
    K_ENDP2          // transition, resolved inside the sweep that raises it: F <- Dol' ++ F'
};

__device__ __forceinline__ void ld64(const void* p, uint32_t* r) {
    asm volatile("ld.global.nc.v2.b32 {%0,%1}, [%2];" : "=r"(r[0]), "=r"(r[1]) : "l"(p));
}

constexpr int ITEM_WARPS = FUSED_THREADS / 32;

}  // namespace

template <bool STATS, int MINB>
__global__ void __launch_bounds__(FUSED_THREADS, MINB)
search_items_kernel(const __grid_constant__ IndexView ix, const __grid_constant__ SeedView sv, uint32_t k, const __grid_constant__ PackedPats pp,
                    uint32_t npat, uint32_t cap, FusedOut out, uint32_t* __restrict__ work, StatsDev* __restrict__ st, uint32_t kc) {
    extern __shared__ uint2 s_list[];                 // [cap][FUSED_THREADS]: column of the lane = its frontier (F from the bottom, Dol from the top)
    __shared__ const uint32_t* s_off[17];             // runtime-indexed kernel parameters would live in local memory
    __shared__ const uint2* s_iv[17];
    __shared__ const uint32_t* s_cnt[17];
    __shared__ uint32_t s_Cl[5];
    // the warp's pool: 32 consecutive patterns, fetched coalesced and with their seed entries already looked up by all 32
    // lanes at once; a lane that needs a pattern reads entry pool_pos + (its rank among the needy lanes)
    __shared__ uint64_t s_pw[ITEM_WARPS][32];         // word 0 of the pattern
    __shared__ uint4 s_ps[ITEM_WARPS][32];            // seed entry {first interval lo, hi, offset of the entry, number of intervals}
    __shared__ uint2 s_pm[ITEM_WARPS][32];            // {length (0: nothing to search), first word}
    if (threadIdx.x < 17) { s_off[threadIdx.x] = sv.off[threadIdx.x]; s_iv[threadIdx.x] = sv.iv[threadIdx.x]; s_cnt[threadIdx.x] = sv.cnt[threadIdx.x]; }
    if (threadIdx.x < 5) s_Cl[threadIdx.x] = ix.Cl[threadIdx.x];
    __syncthreads();
#define SLOT(i) s_list[(i) * FUSED_THREADS + threadIdx.x]
#define LIST(j) SLOT(j)
#define DOL(i) SLOT(cap - 1 - (i))
    constexpr uint32_t FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;

    uint32_t pool_base = 0, pool_pos = 0, pool_n = 0;
    bool exhausted = false;
    auto grab_pool = [&]() {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(work, 32u);
        base = __shfl_sync(FULL, base, 0);
        pool_pos = 0; pool_base = base;
        if (base >= npat) { exhausted = true; pool_n = 0; return; }
        pool_n = npat - base < 32u ? npat - base : 32u;
        if (lane < pool_n) {
            const uint32_t pq = base + lane;
            uint32_t mm, pwb;
            if (pp.poff) {
                const uint64_t a = __ldg(pp.poff + pq), len64 = __ldg(pp.poff + pq + 1) - a;
                mm = len64 > 0xffffffffull ? 0u : (uint32_t)len64;
                pwb = (uint32_t)(((a - __ldg(pp.poff)) >> 5) + pq);
            } else { mm = pp.fixed_len; pwb = pq * pp.stride; }
            if (pp.invalid && ((__ldg(pp.invalid + (pq >> 5)) >> (pq & 31u)) & 1u)) mm = 0;
            uint64_t w = 0; uint4 se = make_uint4(1u, 0u, 0u, 0u);
            if (mm) {
                w = __ldg(pp.words + pwb);
                // seed entry of the last min(m, k) characters: the frontier after that many backward steps
                const uint32_t L0 = mm < k ? mm : k, key = (uint32_t)(w >> (64 - 2 * L0));
                if (sv.first && L0 == sv.first_level) se = __ldg(sv.first + key);
                else {
                    se.z = __ldg(s_off[L0] + key); se.w = __ldg(s_off[L0] + key + 1) - se.z;
                    if (se.w) { const uint2 v = __ldg(s_iv[L0] + se.z); se.x = v.x; se.y = v.y; }
                }
            }
            s_pw[wid][lane] = w; s_ps[wid][lane] = se; s_pm[wid][lane] = make_uint2(mm, pwb);
        }
        __syncwarp();
    };

    // ---- lane state
    uint32_t kind = K_IDLE, q = 0, m = 0, spos = 0, wb = 0;     // spos = length of the pattern suffix matched so far
    uint64_t win = 0;                                             // word spos >> 5 of the pattern
    uint32_t nF = 0, j = 0, nout = 0, ndol = 0;
    uint32_t aux = 0;                                             // pass 1: link records left of F[j]; pass 2: length of Dol'
    uint32_t cur = 0;                                             // K_LINK: link record; K_SEEDIV: seed interval
    bool ovf = false;
    unsigned long long sI = 0, sS = 0, sE = 0, sLF = 0, sBlk = 0;     // counting runs: totals of the lane ...
    uint32_t tI = 0, tS = 0, tE = 0, tLF = 0, tBlk = 0;               // ... and of its current pattern (dropped when it overflows)

    // the frontier of the suffix of length spos is complete: what the lane fetches next
    auto next_step = [&]() {
        if (STATS && spos == (m < kc ? m : kc)) tI += nF;
        if (spos == m || nF == 0) kind = K_FIN;
        else if ((spos & 31u) == 0) kind = K_WORD;
        else { kind = K_OCCF; j = 0; nout = 0; ndol = 0; }
    };

    grab_pool();
    for (;;) {
        // ---------------- finished lanes write their results; they and the idle lanes take the next patterns of the pool
        const uint32_t needy = __ballot_sync(FULL, kind <= K_FIN);
        if (needy) {
            if (kind == K_FIN) {
                // count + the final interval list.  One interval (nearly every found pattern) travels inside the plist entry
                // itself -- its end follows from the count -- longer lists are staged contiguously
                uint32_t cnt = 0; uint2 pl = make_uint2(0u, 0u);
                if (ovf) nF = 0;
                if (nF == 1) { const uint2 v = LIST(0); cnt = v.y - v.x + 1; pl = make_uint2(v.x, PLIST_INLINE); }
                else if (nF > 1) {
                    for (uint32_t t = 0; t < nF; ++t) { const uint2 v = LIST(t); cnt += v.y - v.x + 1; }
                    const uint32_t my = atomicAdd(out.cursor, nF);
                    if ((uint64_t)my + nF > out.staging_cap) { ovf = true; cnt = 0; }
                    else { for (uint32_t t = 0; t < nF; ++t) out.staging[my + t] = LIST(t); pl = make_uint2(my, nF); }
                }
                out.plist[q] = pl; out.counts[q] = cnt;
                if (ovf) out.ovf_ids[atomicAdd(out.ovf_count, 1u)] = q;
                else {
                    if (out.alive) {
                        // where the frontier died ("Pattern not found, interrupted in position", EDSBWTsearch.cpp:714-717)
                        const uint32_t L0 = m < k ? m : k;
                        uint32_t la = 0;
                        if (nF) la = m;
                        else if (spos > L0) la = spos - 1;
                        else {                                    // empty seed entry: longest shorter suffix that is not (win is still word 0)
                            const uint32_t key = (uint32_t)(win >> (64 - 2 * L0));
                            for (uint32_t L = L0 - 1; L >= 1; --L) {
                                const uint32_t kl = key >> (2 * (L0 - L));
                                if (__ldg(s_off[L] + kl + 1) > __ldg(s_off[L] + kl)) { la = L; break; }
                            }
                        }
                        out.alive[q] = la;
                    }
                    if (STATS) { sI += tI; sS += tS; sE += tE; sLF += tLF; sBlk += tBlk; }
                }
                kind = K_IDLE;
            }
            const uint32_t avail = pool_n - pool_pos;
            if (avail) {
                const uint32_t r = __popc(needy & ((1u << lane) - 1u));
                if (((needy >> lane) & 1u) && r < avail) {        // (every needy lane is idle by now)
                    const uint32_t e = pool_pos + r;
                    const uint2 pm = s_pm[wid][e];
                    q = pool_base + e; m = pm.x; wb = pm.y; ovf = false;
                    if (STATS) tI = tS = tE = tLF = tBlk = 0;
                    if (m == 0) {                                 // empty pattern or a byte outside the alphabet: 0 occurrences
                        out.counts[q] = 0; out.plist[q] = make_uint2(0u, 0u);
                        if (out.alive) out.alive[q] = 0;
                    } else {
                        win = s_pw[wid][e];
                        const uint4 se = s_ps[wid][e];
                        const uint32_t L0 = m < k ? m : k, n0 = se.w;
                        spos = L0;
                        if (n0 && m == L0 && n0 <= PLIST_N_MASK) {
                            // pattern no longer than the seed depth: the table IS the answer (count from the per-key table, the
                            // result list is a reference into the seed intervals; a short key's frontier can hold thousands)
                            const uint32_t key = (uint32_t)(win >> (64 - 2 * L0));
                            uint32_t dc = 0;
                            if (s_cnt[L0]) dc = __ldg(s_cnt[L0] + key);
                            else for (uint32_t t = 0; t < n0; ++t) { const uint2 v = __ldg(s_iv[L0] + se.z + t); dc += v.y - v.x + 1; }
                            out.counts[q] = dc; out.plist[q] = make_uint2(se.z, PLIST_SEED | (L0 << 27) | n0);
                            if (out.alive) out.alive[q] = m;
                            if (STATS) { if (m <= kc) sI += n0; }
                        } else if (n0 > cap) { ovf = true; kind = K_FIN; }
                        else {
                            nF = n0;
                            if (n0) LIST(0) = make_uint2(se.x, se.y);
                            if (n0 > 1) { kind = K_SEEDIV; j = 1; cur = se.z + 1; }
                            else next_step();
                        }
                    }
                }
                const uint32_t taken = __popc(needy);
                pool_pos += taken < avail ? taken : avail;
                if (pool_pos >= pool_n && !exhausted) { __syncwarp(); grab_pool(); }      // its loads land while the warp sweeps on
            } else if (exhausted) {
                if (__all_sync(FULL, kind == K_IDLE)) break;
            } else grab_pool();
        }

        // ---------------- issue: one fetch per lane, all in flight together
        // The 8-byte fetches go out FIRST, ahead of the Occ fetches.  The loads are volatile asm and keep their order, so placed
        // here the compiler cannot sink them into the branches that consume them; placed after the Occ loads it did (same
        // condition, no dependence), the warp then ran "Occ lanes: load, wait, consume" and only afterwards "link lanes: load,
        // wait, consume", and the latencies added up: three long-scoreboard stall sites instead of one in ncu, kernel 12 %
        // slower (profiles/r02_items_ab.txt).
        // (Registers of their own for the 8 bytes: sharing B, the Occ load behind them waited for them to land -- a
        // write-after-write hazard on the scoreboard, 10 % of all stall samples.)
        uint32_t B[16], S[2];
        if (kind == K_LINK) ld64(ix.link_rec + cur, S);
        else if (kind == K_SEEDIV) ld64(s_iv[m < k ? m : k] + cur, S);
        else if (kind == K_WORD) ld64(pp.words + wb + (spos >> 5), S);
        const bool occ_item = kind == K_OCCF || kind == K_OCCD;
        const uint32_t slot = kind == K_OCCD ? cap - 1 - j : j;  // pass 2 walks the Dol set at the top of the column
        if (occ_item) {
            const uint4* p = ix.occ + (size_t)(SLOT(slot).x >> 7) * 4;
            ld256(p, B); ld256(p + 2, B + 8);
        }

        // ---------------- consume
        if (occ_item) {
            const uint2 iv = SLOT(slot);
            const int c = (int)((win >> (2 * (31 - (spos & 31u)))) & 3u);
            const uint32_t x0 = (c & 1) ? 0u : 0xffffffffu, x1 = (c & 2) ? 0u : 0xffffffffu;
            uint32_t rca, rcb, rda, rdb, nblk;
            lf_range_loaded(*reinterpret_cast<Blk*>(B), ix.occ, iv.x, iv.y + 1, c, x0, x1, rca, rcb, rda, rdb, nblk);
            if (STATS) { tLF += 1; tBlk += nblk; if (spos >= kc) { tS += 1; tE += rdb - rda; } }
            // one code path for both passes: the output list grows from the bottom (F', length nout) or from the top (Dol', length aux)
            const bool pass2 = kind == K_OCCD;
            uint32_t no = pass2 ? aux : nout;
            if (rcb > rca) {
                const uint32_t Cc = (c & 2) ? ((c & 1) ? ix.Cl[3] : ix.Cl[2]) : ((c & 1) ? ix.Cl[1] : ix.Cl[0]);
                const uint32_t nlo = Cc + rca, nhi = Cc + rcb - 1;
                const uint32_t prev = pass2 ? cap - no : no - 1;      // slot of the last output (valid when no > 0)
                if (no && SLOT(prev).y + 1 == nlo) SLOT(prev).y = nhi;
                else { SLOT(pass2 ? cap - 1 - no : no) = make_uint2(nlo, nhi); ++no; }
            }
            ++j;
            if (pass2) { aux = no; if (j == ndol) kind = K_ENDP2; }
            else {
                nout = no;
                if (rdb > rda) { kind = K_LINK; cur = rda; aux = rdb - rda; }
                else if (j == nF) {
                    if (ndol) { kind = K_OCCD; j = 0; aux = 0; }
                    else { nF = nout; ++spos; next_step(); }
                }
            }
        } else if (kind == K_LINK) {
            // pile-0 range [l, h] of the preceding segment(s) -> sorted, merged Dol set at the top of the column
            uint32_t l = S[0], h = S[1];
            if (l <= h) {                                         // l > h: string of the first segment, no predecessor
                uint32_t i = 0;
                while (i < ndol && DOL(i).y + 1 < l) ++i;
                if (i < ndol && DOL(i).x <= h + 1) {              // overlaps / touches D[i]: merge, then absorb followers
                    const uint2 d = DOL(i);
                    l = l < d.x ? l : d.x; h = h > d.y ? h : d.y;
                    uint32_t t = i + 1;
                    while (t < ndol && DOL(t).x <= h + 1) { const uint32_t y = DOL(t).y; h = h > y ? h : y; ++t; }
                    DOL(i) = make_uint2(l, h);
                    if (t > i + 1) { for (uint32_t u = t; u < ndol; ++u) DOL(i + 1 + (u - t)) = DOL(u); ndol -= t - (i + 1); }
                } else if (nF + ndol + 1 >= cap) { ovf = true; kind = K_FIN; }
                else {                                            // insert before i
                    for (uint32_t u = ndol; u > i; --u) DOL(u) = DOL(u - 1);
                    DOL(i) = make_uint2(l, h);
                    ++ndol;
                }
            }
            if (!ovf) {
                ++cur;
                if (--aux == 0) {                                 // (j already counts F[j] as done)
                    if (j < nF) kind = K_OCCF;
                    else if (ndol) { kind = K_OCCD; j = 0; aux = 0; }
                    else { nF = nout; ++spos; next_step(); }
                }
            }
        } else if (kind == K_SEEDIV) {
            LIST(j) = make_uint2(S[0], S[1]);
            ++cur;
            if (++j == nF) next_step();
        } else if (kind == K_WORD) {
            win = ((uint64_t)S[1] << 32) | S[0];
            kind = K_OCCF; j = 0; nout = 0; ndol = 0;
        }
        if (kind == K_ENDP2) {                                    // F <- Dol' ++ F'   (aux = length of Dol')
            if (aux) {
                if (nout + 2 * aux > cap) { ovf = true; kind = K_FIN; }
                else {
                    for (uint32_t t = nout; t > 0; --t) LIST(t - 1 + aux) = LIST(t - 1);
                    for (uint32_t i = 0; i < aux; ++i) LIST(i) = DOL(i);
                    nout += aux;
                }
            }
            if (!ovf) { nF = nout; ++spos; next_step(); }
        }
    }
    if (STATS) {
        const unsigned long long v[5] = {sI, sS, sE, sLF, sBlk};
        unsigned long long* dst[5] = {&st->I, &st->S, &st->E, &st->lf_intervals, &st->occ_blocks};
        #pragma unroll
        for (int i = 0; i < 5; ++i) {
            unsigned long long x = v[i];
            for (int o = 16; o; o >>= 1) x += __shfl_xor_sync(FULL, x, o);
            if (lane == 0 && x) atomicAdd(dst[i], x);
        }
    }
#undef SLOT
#undef LIST
#undef DOL
}

namespace {
template <bool STATS, int MINB>
void launch_items_instance(size_t smem, cudaStream_t s, const IndexView& ix, const SeedView& sv, uint32_t k, const PackedPats& pp, uint32_t npat,
                           uint32_t cap, const FusedOut& out, uint32_t* work, StatsDev* st, uint32_t kc) {
    auto kern = search_items_kernel<STATS, MINB>;
    CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));      // per device: set at every launch
    int dev = 0, sms = 148, per_sm = 1;
    CUDA_OK(cudaGetDevice(&dev));
    CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, FUSED_THREADS, smem));
    if (per_sm < 1) per_sm = 1;
    // persistent: one wave of CTAs, every warp pulls pools of 32 patterns until the batch is empty
    const unsigned want = (npat + FUSED_THREADS - 1) / FUSED_THREADS, wave = (unsigned)(sms * per_sm);
    const unsigned grid = want < wave ? want : wave;
    kern<<<grid, FUSED_THREADS, smem, s>>>(ix, sv, k, pp, npat, cap, out, work, st, kc);
}
}  // namespace

// work: device counter, zeroed by the caller (the next unclaimed pattern)
void launch_items_search(const IndexView& ix, const SeedView& sv, uint32_t k, const PackedPats& pp, uint32_t npat, uint32_t cap, const FusedOut& out,
                         uint32_t* work, StatsDev* st, uint32_t kc, cudaStream_t s) {
    if (npat == 0) return;
    if (npat > 0xff000000u) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "more than 2^32 - 2^24 patterns in one search call; split the batch");
    const size_t smem = (size_t)cap * FUSED_THREADS * sizeof(uint2);
    // 10 CTAs / SM at 48 registers (A/B against 8, 9 and 12: profiles/r02_items_ab.txt); long frontiers and counting runs: 8
    if (st) launch_items_instance<true, 8>(smem, s, ix, sv, k, pp, npat, cap, out, work, st, kc);
    else if (cap > 16) launch_items_instance<false, 8>(smem, s, ix, sv, k, pp, npat, cap, out, work, st, kc);
    else launch_items_instance<false, 10>(smem, s, ix, sv, k, pp, npat, cap, out, work, st, kc);
    CUDA_OK(cudaGetLastError());
}

}  // namespace edsbwt
