// GPU index builder: string collection -> the reference's on-disk index files.
//
// Replaces the offline pipeline eds_to_fasta (eds_to_fasta.cpp:55-167) -> gsufsort --da --bwt ->
// da_to_everything (da_to_everything.cpp:113-290, 371-486) with one device-side suffix sort of the
// collection.  Only the OUTPUT CONVENTION of gsufsort matters: suffixes of all strings
// sorted with the end-markers ordered by string index; the BWT character of a whole-string row is the
// end-marker; an empty string contributes exactly one row (pile 0, '#').
//
// Algorithm (device):
//   T      = S_0 $ S_1 $ ... with letters coded 1..sigma-1 and $ = 0                      (text_kernel)
//   round 0: key(p) = the first CPK characters of suffix p packed MSB-first, everything after the
//            end-marker zeroed; stable LSD radix sort of (key, p).  Two suffixes with equal keys that
//            contain an end-marker are ordered by position == by string index (stability), which is
//            exactly the gsufsort convention, so they are resolved immediately.
//   rounds h = CPK, 2CPK, ...: prefix doubling restricted to the still-ambiguous groups
//            (key = rank[p] : rank[p+h]), until every group is a singleton.
//   BWT / pile / EOF_ID extraction from the suffix array                                   (bwt_kernel)
// Host: tableOcc, sdsl rrr_vector / bit_vector serialisation and file writing (host_io.cpp).
#include <cub/cub.cuh>

#include <algorithm>

#include "internal.cuh"

namespace edsbwt {

namespace {

inline unsigned grid_of(size_t n, unsigned block) { return (unsigned)((n + block - 1) / block); }

struct CodeLut { uint8_t code[256]; };

// T[p] for every letter: position i of the separator-free text belongs to string w = (#str_off[1..] <= i)
__global__ void text_letters_kernel(const uint8_t* __restrict__ letters, uint64_t nlet, const uint64_t* __restrict__ str_off, uint32_t nstr,
                                    CodeLut lut, uint8_t* __restrict__ T) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nlet) return;
    uint32_t lo = 0, hi = nstr;                         // first w with str_off[w+1] > i
    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (str_off[mid + 1] <= i) lo = mid + 1; else hi = mid; }
    T[i + lo] = lut.code[letters[i]];
}
__global__ void text_terms_kernel(const uint64_t* __restrict__ str_off, uint32_t nstr, uint8_t* __restrict__ T, uint32_t* __restrict__ tpos) {
    uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= nstr) return;
    uint64_t p = str_off[w + 1] + w;
    T[p] = 0; tpos[w] = (uint32_t)p;
}
// round-0 keys
__global__ void key0_kernel(const uint8_t* __restrict__ T, uint32_t n, uint32_t bits, uint32_t cpk, uint64_t* __restrict__ key, uint32_t* __restrict__ pos) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    uint64_t k = 0; bool live = true;
    for (uint32_t j = 0; j < cpk; ++j) {
        uint32_t c = 0;
        if (live) { c = T[p + j]; if (c == 0) live = false; }       // T is padded with cpk zero bytes
        k = (k << bits) | c;
    }
    key[p] = k << (64 - bits * cpk);
    pos[p] = p;
}
__device__ __forceinline__ bool key_terminated(uint64_t k, uint32_t bits, uint32_t cpk) {
    uint64_t m = (1ull << bits) - 1;
    k >>= (64 - bits * cpk);
    for (uint32_t j = 0; j < cpk; ++j) { if ((k & m) == 0) return true; k >>= bits; }
    return false;
}
// head flags after round 0 (sorted order)
__global__ void heads0_kernel(const uint64_t* __restrict__ key, uint32_t n, uint32_t bits, uint32_t cpk, uint32_t* __restrict__ head) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t k = key[i];
    head[i] = (i == 0 || k != key[i - 1] || key_terminated(k, bits, cpk)) ? i + 1 : 0;    // i+1 so that a max-scan yields the head slot
}
// rank[pos[i]] = slot of the head of i's group ; amb[i] = 1 when the group has more than one member
__global__ void ranks_kernel(const uint32_t* __restrict__ headslot, const uint32_t* __restrict__ head, const uint32_t* __restrict__ pos, uint32_t n,
                             uint32_t* __restrict__ rank, uint32_t* __restrict__ amb) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    rank[pos[i]] = headslot[i] - 1;
    bool single = head[i] != 0 && (i + 1 == n || head[i + 1] != 0);
    amb[i] = single ? 0u : 1u;
}
struct MaxOp { __device__ __forceinline__ uint32_t operator()(uint32_t a, uint32_t b) const { return a > b ? a : b; } };

// compacted ambiguous set: slot[j], p[j] ; doubling key
__global__ void gather_amb_kernel(const uint32_t* __restrict__ amb, const uint32_t* __restrict__ apos, const uint32_t* __restrict__ pos, uint32_t n,
                                  uint32_t* __restrict__ slot, uint32_t* __restrict__ p_out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !amb[i]) return;
    uint32_t j = apos[i];
    slot[j] = i; p_out[j] = pos[i];
}
__global__ void key2_kernel(const uint32_t* __restrict__ p, uint32_t m, const uint32_t* __restrict__ rank, uint32_t n, uint32_t h, uint64_t* __restrict__ key) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= m) return;
    uint32_t q = p[j];
    uint64_t r2 = (uint64_t)q + h < n ? rank[q + h] : 0;
    key[j] = ((uint64_t)rank[q] << 32) | r2;
}
// after sorting the ambiguous set: new heads (value = global slot + 1 of the head, 0 otherwise)
__global__ void heads2_kernel(const uint64_t* __restrict__ key, const uint32_t* __restrict__ slot, uint32_t m, uint32_t* __restrict__ head) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= m) return;
    head[j] = (j == 0 || key[j] != key[j - 1]) ? slot[j] + 1 : 0;
}
__global__ void update2_kernel(const uint32_t* __restrict__ headslot, const uint32_t* __restrict__ head, const uint32_t* __restrict__ slot,
                               const uint32_t* __restrict__ p_sorted, uint32_t m, uint32_t* __restrict__ rank, uint32_t* __restrict__ pos,
                               uint32_t* __restrict__ amb) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= m) return;
    uint32_t q = p_sorted[j];
    pos[slot[j]] = q;
    rank[q] = headslot[j] - 1;
    bool single = head[j] != 0 && (j + 1 == m || head[j + 1] != 0);
    amb[j] = single ? 0u : 1u;
}
__global__ void compact2_kernel(const uint32_t* __restrict__ amb, const uint32_t* __restrict__ apos, const uint32_t* __restrict__ slot,
                                const uint32_t* __restrict__ p_sorted, uint32_t m, uint32_t* __restrict__ slot_out, uint32_t* __restrict__ p_out) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= m || !amb[j]) return;
    uint32_t d = apos[j];
    slot_out[d] = slot[j]; p_out[d] = p_sorted[j];
}
// suffix array -> eBWT bytes, end-marker flags
__global__ void bwt_kernel(const uint8_t* __restrict__ T, const uint32_t* __restrict__ sa, uint32_t n, const uint8_t* __restrict__ alphabet,
                           uint8_t* __restrict__ ebwt, uint32_t* __restrict__ is_eof) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t p = sa[i];
    uint8_t c = p == 0 ? 0 : T[p - 1];          // 0 = the previous string's end-marker -> p starts a string
    ebwt[i] = alphabet[c];
    is_eof[i] = c == 0 ? 1u : 0u;
}
// EOF_ID (document array restricted to the '#' rows, in row order): string whose start is sa[i]
__global__ void eof_kernel(const uint32_t* __restrict__ sa, const uint32_t* __restrict__ is_eof, const uint32_t* __restrict__ epos, uint32_t n,
                           const uint32_t* __restrict__ tpos, uint32_t nstr, uint32_t* __restrict__ eof_id) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !is_eof[i]) return;
    uint32_t p = sa[i];
    uint32_t lo = 0, hi = nstr;                  // first w with tpos[w] >= p
    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (tpos[mid] < p) lo = mid + 1; else hi = mid; }
    eof_id[epos[i]] = lo;
}

struct Tmp {
    DevBuf<uint8_t> buf;
    void* get(size_t bytes) { buf.ensure(bytes + 256); return buf.p; }
};
void excl_sum(Tmp& tmp, const uint32_t* in, uint32_t* out, size_t n, cudaStream_t s) {
    size_t b = 0;
    CUDA_OK(cub::DeviceScan::ExclusiveSum(nullptr, b, in, out, (int64_t)n, s));
    CUDA_OK(cub::DeviceScan::ExclusiveSum(tmp.get(b), b, in, out, (int64_t)n, s));
}
void incl_max(Tmp& tmp, const uint32_t* in, uint32_t* out, size_t n, cudaStream_t s) {
    size_t b = 0;
    CUDA_OK(cub::DeviceScan::InclusiveScan(nullptr, b, in, out, MaxOp(), (int64_t)n, s));
    CUDA_OK(cub::DeviceScan::InclusiveScan(tmp.get(b), b, in, out, MaxOp(), (int64_t)n, s));
}
void sort_pairs(Tmp& tmp, cub::DoubleBuffer<uint64_t>& k, cub::DoubleBuffer<uint32_t>& v, size_t n, int begin_bit, int end_bit, cudaStream_t s) {
    size_t b = 0;
    CUDA_OK(cub::DeviceRadixSort::SortPairs(nullptr, b, k, v, (int64_t)n, begin_bit, end_bit, s));
    CUDA_OK(cub::DeviceRadixSort::SortPairs(tmp.get(b), b, k, v, (int64_t)n, begin_bit, end_bit, s));
}
uint32_t fetch_u32(const uint32_t* d, cudaStream_t s) {
    uint32_t v = 0;
    CUDA_OK(cudaMemcpyAsync(&v, d, 4, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    return v;
}

}  // namespace

// Device suffix sort of the collection.  Outputs (host): ebwt bytes [n], eof_id [nstr], alphabet, pile starts.
static void suffix_sort_collection(const uint8_t* letters, const uint64_t* str_off, uint32_t nstr, int device,
                                   std::vector<uint8_t>& ebwt, std::vector<uint32_t>& eof_id, uint8_t alphabet[256], uint32_t& sigma,
                                   std::vector<uint64_t>& pile_start) {
    const uint64_t nlet = str_off[nstr];
    const uint64_t n64 = nlet + nstr;
    if (n64 >= 0xfffffff0ull) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "collection of %llu rows exceeds the reference's 32-bit index types (Parameters.h:65-104)", (unsigned long long)n64);
    const uint32_t n = (uint32_t)n64;
    // alphabet (ascending byte values, '#' first: da_to_everything.cpp:149-163)
    uint64_t hist[256] = {0};
    for (uint64_t i = 0; i < nlet; ++i) hist[letters[i]]++;
    CodeLut lut; memset(lut.code, 0, sizeof lut.code);
    memset(alphabet, 0, 256);
    alphabet[0] = '#'; sigma = 1;
    pile_start.assign(1, 0);
    uint64_t acc = nstr;
    for (int c = 0; c < 256; ++c) if (hist[c]) {
        if (c <= '#') EDS_FAIL(EDSBWT_E_FORMAT, "letter 0x%02x does not sort above the end-marker '#'", c);
        lut.code[c] = (uint8_t)sigma; alphabet[sigma++] = (uint8_t)c;
        pile_start.push_back(acc); acc += hist[c];
    }
    pile_start.push_back(acc);
    uint32_t bits = 1; while ((1u << bits) < sigma) ++bits;
    const uint32_t cpk = 64 / bits;

    CUDA_OK(cudaSetDevice(device));
    cudaStream_t s; CUDA_OK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamDestroy(s); } } guard{s};
    Tmp tmp;
    DevBuf<uint8_t> d_T, d_alpha; DevBuf<uint32_t> d_tpos;
    d_T.ensure((size_t)n + cpk + 8); d_tpos.ensure((size_t)nstr + 1); d_alpha.ensure(256);
    CUDA_OK(cudaMemsetAsync(d_T.p, 0, (size_t)n + cpk + 8, s));
    CUDA_OK(cudaMemcpyAsync(d_alpha.p, alphabet, 256, cudaMemcpyHostToDevice, s));
    {
        DevBuf<uint8_t> d_let; DevBuf<uint64_t> d_off;
        d_let.ensure(nlet + 1); d_off.ensure((size_t)nstr + 1);
        if (nlet) CUDA_OK(cudaMemcpyAsync(d_let.p, letters, nlet, cudaMemcpyHostToDevice, s));
        CUDA_OK(cudaMemcpyAsync(d_off.p, str_off, ((size_t)nstr + 1) * 8, cudaMemcpyHostToDevice, s));
        if (nlet) text_letters_kernel<<<grid_of(nlet, 256), 256, 0, s>>>(d_let.p, nlet, d_off.p, nstr, lut, d_T.p);
        text_terms_kernel<<<grid_of(nstr, 256), 256, 0, s>>>(d_off.p, nstr, d_T.p, d_tpos.p);
        CUDA_OK(cudaStreamSynchronize(s));
    }
    // ---- round 0
    DevBuf<uint64_t> kA, kB; DevBuf<uint32_t> vA, vB, head, headslot, rank, amb, apos;
    kA.ensure(n); kB.ensure(n); vA.ensure(n); vB.ensure(n); head.ensure((size_t)n + 1); headslot.ensure((size_t)n + 1);
    rank.ensure((size_t)n + 1); amb.ensure((size_t)n + 1); apos.ensure((size_t)n + 1);
    key0_kernel<<<grid_of(n, 256), 256, 0, s>>>(d_T.p, n, bits, cpk, kA.p, vA.p);
    cub::DoubleBuffer<uint64_t> kb(kA.p, kB.p); cub::DoubleBuffer<uint32_t> vb(vA.p, vB.p);
    sort_pairs(tmp, kb, vb, n, (int)(64 - bits * cpk), 64, s);
    uint32_t* sa = vb.Current();                 // sorted positions; refined in place by the doubling rounds
    uint32_t* spare32 = vb.Alternate();
    heads0_kernel<<<grid_of(n, 256), 256, 0, s>>>(kb.Current(), n, bits, cpk, head.p);
    incl_max(tmp, head.p, headslot.p, n, s);
    ranks_kernel<<<grid_of(n, 256), 256, 0, s>>>(headslot.p, head.p, sa, n, rank.p, amb.p);
    CUDA_OK(cudaMemsetAsync(amb.p + n, 0, 4, s));
    excl_sum(tmp, amb.p, apos.p, (size_t)n + 1, s);
    uint32_t m = fetch_u32(apos.p + n, s);
    // ---- doubling on the ambiguous groups
    if (m) {
        DevBuf<uint32_t> slotA, slotB, pA, pB;
        slotA.ensure(m); slotB.ensure(m); pA.ensure(m); pB.ensure(m);
        gather_amb_kernel<<<grid_of(n, 256), 256, 0, s>>>(amb.p, apos.p, sa, n, slotA.p, pA.p);
        uint64_t* k2a = kb.Current(); uint64_t* k2b = kb.Alternate();       // reuse the round-0 key buffers (m <= n)
        uint32_t* slot = slotA.p; uint32_t* slot_alt = slotB.p;
        uint32_t* pcur = pA.p; uint32_t* palt = pB.p;
        for (uint64_t h = cpk; m; h *= 2) {
            if (h > (uint64_t)n * 2) EDS_FAIL(EDSBWT_E_FORMAT, "suffix sort did not converge");
            key2_kernel<<<grid_of(m, 256), 256, 0, s>>>(pcur, m, rank.p, n, (uint32_t)std::min<uint64_t>(h, n), k2a);
            cub::DoubleBuffer<uint64_t> k2(k2a, k2b); cub::DoubleBuffer<uint32_t> v2(pcur, palt);
            sort_pairs(tmp, k2, v2, m, 0, 64, s);
            k2a = k2.Current(); k2b = k2.Alternate(); pcur = v2.Current(); palt = v2.Alternate();
            heads2_kernel<<<grid_of(m, 256), 256, 0, s>>>(k2a, slot, m, head.p);
            incl_max(tmp, head.p, headslot.p, m, s);
            update2_kernel<<<grid_of(m, 256), 256, 0, s>>>(headslot.p, head.p, slot, pcur, m, rank.p, sa, amb.p);
            CUDA_OK(cudaMemsetAsync(amb.p + m, 0, 4, s));
            excl_sum(tmp, amb.p, apos.p, (size_t)m + 1, s);
            uint32_t m2 = fetch_u32(apos.p + m, s);
            if (m2) {
                compact2_kernel<<<grid_of(m, 256), 256, 0, s>>>(amb.p, apos.p, slot, pcur, m, slot_alt, palt);
                std::swap(slot, slot_alt); std::swap(pcur, palt);
            }
            m = m2;
        }
    }
    (void)spare32;
    // ---- eBWT + EOF_ID
    DevBuf<uint8_t> d_ebwt; d_ebwt.ensure((size_t)n + 1);
    DevBuf<uint32_t> d_eof; d_eof.ensure((size_t)nstr + 1);
    uint32_t* is_eof = amb.p; uint32_t* epos = apos.p;
    bwt_kernel<<<grid_of(n, 256), 256, 0, s>>>(d_T.p, sa, n, d_alpha.p, d_ebwt.p, is_eof);
    CUDA_OK(cudaMemsetAsync(is_eof + n, 0, 4, s));
    excl_sum(tmp, is_eof, epos, (size_t)n + 1, s);
    uint32_t neof = fetch_u32(epos + n, s);
    if (neof != nstr) EDS_FAIL(EDSBWT_E_FORMAT, "internal: %u whole-string rows for %u strings", neof, nstr);
    eof_kernel<<<grid_of(n, 256), 256, 0, s>>>(sa, is_eof, epos, n, d_tpos.p, nstr, d_eof.p);
    ebwt.resize(n); eof_id.resize(nstr);
    CUDA_OK(cudaMemcpyAsync(ebwt.data(), d_ebwt.p, n, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaMemcpyAsync(eof_id.data(), d_eof.p, (size_t)nstr * 4, cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    CUDA_OK(cudaGetLastError());
}

void build_index_files(const uint8_t* letters, const uint64_t* str_off, uint32_t nstr, const uint8_t* seg_first,
                       const std::string& out_base, int device) {
    if (str_off[0] != 0) EDS_FAIL(EDSBWT_E_ARG, "str_off[0] must be 0");
    for (uint32_t w = 0; w < nstr; ++w) if (str_off[w + 1] < str_off[w]) EDS_FAIL(EDSBWT_E_ARG, "str_off is not ascending at %u", w);
    if (!seg_first[0]) EDS_FAIL(EDSBWT_E_ARG, "string 0 must open a segment");
    std::vector<uint8_t> ebwt; std::vector<uint32_t> eof_id; std::vector<uint64_t> pile_start;
    uint8_t alphabet[256]; uint32_t sigma = 0;
    suffix_sort_collection(letters, str_off, nstr, device, ebwt, eof_id, alphabet, sigma, pile_start);
    const uint32_t n = (uint32_t)ebwt.size();
    // tableOcc + files (App. A.1-A.4)
    uint8_t code[256]; memset(code, 0, sizeof code);
    for (uint32_t a = 0; a < sigma; ++a) code[alphabet[a]] = (uint8_t)a;
    std::vector<uint32_t> table((size_t)sigma * sigma, 0);
    std::vector<uint8_t> bm;
    for (uint32_t p = 0; p < sigma; ++p) {
        const uint64_t a = pile_start[p], b = pile_start[p + 1];
        const uint8_t* pile = ebwt.data() + a;
        uint64_t cnt[256] = {0};
        bm.resize((size_t)(b - a));
        for (uint64_t i = 0; i < b - a; ++i) { cnt[pile[i]]++; bm[(size_t)i] = pile[i] == '#'; }
        for (uint32_t c = 0; c < sigma; ++c) table[(size_t)p * sigma + c] = (uint32_t)cnt[alphabet[c]];
        write_file(out_base + "_bwt_" + std::to_string(p) + ".aux", pile, (size_t)(b - a));
        std::vector<uint8_t> rrr = rrr_encode(bm.data(), bm.size());
        write_file(out_base + "_bv_" + std::to_string(p) + ".aux", rrr.data(), rrr.size());
    }
    std::vector<uint8_t> info;
    auto put32 = [&](uint32_t v) { size_t o = info.size(); info.resize(o + 4); memcpy(&info[o], &v, 4); };
    put32(n); put32(nstr); info.push_back((uint8_t)sigma);
    info.insert(info.end(), alphabet, alphabet + sigma);
    { size_t o = info.size(); info.resize(o + (size_t)nstr * 4); if (nstr) memcpy(&info[o], eof_id.data(), (size_t)nstr * 4); }
    { size_t o = info.size(); info.resize(o + table.size() * 4); memcpy(&info[o], table.data(), table.size() * 4); }
    write_file(out_base + "_info.aux", info.data(), info.size());
    std::vector<uint8_t> bv = bitvector_encode(seg_first, nstr);
    write_file(out_base + ".bitvector", bv.data(), bv.size());
}

}  // namespace edsbwt
