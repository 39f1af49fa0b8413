// Result emission: the reference's csv row ORDER and device-side decimal formatting.
//
// Row order of EDSBWTsearch (RECOVERBW=1): findMultipleDollarsBackward runs rounds r = 0, 1, ... (EDSBWTsearch.cpp:349-603);
// in round r the intervals are sorted by (pile, start) (quickSort, Sorting.cpp:8-18) and every '#' of an interval is
// printed left to right (indices_of_dollars_in_interval, :1310-1332).  The intervals of a round are disjoint, so the rows
// of pattern p come out ordered by (offset r, position of the '#' row in the eBWT) = (off, rank of the string in the
// flattened EOF_ID).  order_rows_reference() sorts the tuples of each pattern by that key (cub::DeviceSegmentedSort over
// the per-pattern segments), which makes an unsorted `diff` against the reference's csv pass.
// MOVE_EDSBWTSearch prints rows in frontier order, ascending row inside an interval (MOVE_EDSBWTSearch.cpp:326-369): that
// is the order the locate kernel already emits, nothing to do.
//
// Formatting: one thread per row, digits written into a shared-memory tile, tiles copied out coalesced; the text comes
// back over PCIe as bytes ready for fwrite (replaces one `searchOutput << ... << endl` flush per row, :1327).
#include "internal.cuh"

#include <cub/cub.cuh>

namespace edsbwt {

struct CsvScratch {
    DevBuf<uint64_t> k0, k1;
    DevBuf<uint32_t> v0, v1, len, ofs;
    DevBuf<edsbwt_occ> tmp;
    DevBuf<uint8_t> cub, text;
    char* h_text = nullptr; size_t h_cap = 0;
    uint32_t* h_pin = nullptr;
    ~CsvScratch() { if (h_text) cudaFreeHost(h_text); if (h_pin) cudaFreeHost(h_pin); }
    void host_ensure(size_t n) {
        if (n <= h_cap) return;
        if (h_text) cudaFreeHost(h_text);
        h_text = nullptr; h_cap = 0;
        CUDA_OK(cudaHostAlloc((void**)&h_text, n + n / 4 + 4096, cudaHostAllocDefault));
        h_cap = n + n / 4 + 4096;
    }
};
void free_csv_scratch(CsvScratch* c) { delete c; }
static CsvScratch& scratch(Index& ix) {
    if (!ix.csv) { ix.csv = new CsvScratch; CUDA_OK(cudaHostAlloc((void**)&ix.csv->h_pin, 64, cudaHostAllocDefault)); }
    return *ix.csv;
}
static inline unsigned grid_for(size_t n, unsigned block) { return (unsigned)((n + block - 1) / block); }

// ---------------------------------------------------------------------------------------------------
// reference row order
// ---------------------------------------------------------------------------------------------------
__global__ void eof_inv_kernel(const uint32_t* __restrict__ eof_flat, uint32_t n_text, uint32_t* __restrict__ inv) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n_text) inv[__ldg(eof_flat + k)] = k;
}
__global__ void order_keys_kernel(const edsbwt_occ* __restrict__ occ, uint64_t nocc, const uint32_t* __restrict__ eof_inv,
                                  uint64_t* __restrict__ key, uint32_t* __restrict__ val) {
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nocc) return;
    key[t] = ((uint64_t)occ[t].off << 32) | __ldg(eof_inv + occ[t].str);
    val[t] = (uint32_t)t;
}
__global__ void permute_rows_kernel(const edsbwt_occ* __restrict__ in, const uint32_t* __restrict__ perm, uint64_t nocc, edsbwt_occ* __restrict__ out) {
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < nocc) out[t] = in[__ldg(perm + t)];
}

uint64_t order_rows_reference(Index& ix, SearchOut& out, const uint32_t* d_occ_ofs, uint32_t npat, cudaStream_t s) {
    const uint64_t nocc = out.nocc;
    if (nocc < 2) return 0;
    if (nocc > 0x7fffffffull) EDS_FAIL(EDSBWT_E_UNSUPPORTED, "more than 2^31 rows to order in one batch; split the batch (edsbwt_set_chunk)");
    CsvScratch& c = scratch(ix);
    uint64_t launches = 0;
    if (!ix.eof_inv.p) {
        ix.eof_inv.ensure(ix.view.n_text);
        eof_inv_kernel<<<grid_for(ix.view.n_text, 256), 256, 0, s>>>(ix.view.eof_flat, ix.view.n_text, ix.eof_inv.p);
        ++launches;
    }
    c.k0.ensure(nocc); c.k1.ensure(nocc); c.v0.ensure(nocc); c.v1.ensure(nocc); c.tmp.ensure(nocc);
    order_keys_kernel<<<grid_for(nocc, 256), 256, 0, s>>>(out.occ.p, nocc, ix.eof_inv.p, c.k0.p, c.v0.p);
    size_t tmp = 0;
    CUDA_OK(cub::DeviceSegmentedSort::SortPairs(nullptr, tmp, c.k0.p, c.k1.p, c.v0.p, c.v1.p, (int)nocc, (int)npat, d_occ_ofs, d_occ_ofs + 1, s));
    c.cub.ensure(tmp);
    CUDA_OK(cub::DeviceSegmentedSort::SortPairs(c.cub.p, tmp, c.k0.p, c.k1.p, c.v0.p, c.v1.p, (int)nocc, (int)npat, d_occ_ofs, d_occ_ofs + 1, s));
    permute_rows_kernel<<<grid_for(nocc, 256), 256, 0, s>>>(out.occ.p, c.v1.p, nocc, c.tmp.p);
    CUDA_OK(cudaGetLastError());
    // the sorted copy becomes the result buffer
    std::swap(out.occ.p, c.tmp.p); std::swap(out.occ.cap, c.tmp.cap);
    return launches + 5;
}

// ---------------------------------------------------------------------------------------------------
// decimal formatting
// ---------------------------------------------------------------------------------------------------
constexpr int CSV_ROWS = 256;           // rows per CTA
constexpr int CSV_MAXROW = 55;          // 5 x 10 digits + 4 tabs + newline

__device__ __forceinline__ uint32_t ndigits(uint32_t v) {
    return v < 10u ? 1u : v < 100u ? 2u : v < 1000u ? 3u : v < 10000u ? 4u : v < 100000u ? 5u : v < 1000000u ? 6u
         : v < 10000000u ? 7u : v < 100000000u ? 8u : v < 1000000000u ? 9u : 10u;
}
__global__ void csv_len_kernel(const edsbwt_occ* __restrict__ occ, uint32_t n, uint32_t pat_add, uint32_t* __restrict__ len) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    edsbwt_occ o = occ[t];
    len[t] = ndigits(o.pat + pat_add) + ndigits(o.str) + ndigits(o.seg) + ndigits(o.idx_in_seg) + ndigits(o.off) + 5u;
}
__device__ __forceinline__ uint8_t* put_dec(uint8_t* end, uint32_t v) {      // writes backwards, returns the new end
    do { *--end = (uint8_t)('0' + v % 10u); v /= 10u; } while (v);
    return end;
}
__global__ void __launch_bounds__(CSV_ROWS)
csv_format_kernel(const edsbwt_occ* __restrict__ occ, uint32_t n, uint32_t pat_add, const uint32_t* __restrict__ ofs, uint8_t* __restrict__ text) {
    __shared__ uint8_t tile[CSV_ROWS * CSV_MAXROW];
    const uint32_t first = blockIdx.x * CSV_ROWS, t = first + threadIdx.x;
    const uint32_t last = min(first + CSV_ROWS, n);
    const uint32_t base = __ldg(ofs + first), stop = __ldg(ofs + last);      // ofs has n + 1 entries
    if (t < n) {
        edsbwt_occ o = occ[t];
        uint8_t* e = tile + (__ldg(ofs + t + 1) - base);
        *--e = '\n'; e = put_dec(e, o.off);
        *--e = '\t'; e = put_dec(e, o.idx_in_seg);
        *--e = '\t'; e = put_dec(e, o.seg);
        *--e = '\t'; e = put_dec(e, o.str);
        *--e = '\t'; put_dec(e, o.pat + pat_add);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < stop - base; i += CSV_ROWS) text[(size_t)base + i] = tile[i];
}

// Formats rows [0, nocc) of a DEVICE tuple array and appends the text to f.  pat_add = 1 + pat_base not yet in the tuples.
// Rows are cut into slices whose text stays below 2^31 bytes (32-bit offsets inside a slice).
uint64_t write_csv_device(Index& ix, const edsbwt_occ* d_occ, uint64_t nocc, uint32_t pat_add, FILE* f, cudaStream_t s, uint64_t* launches) {
    CsvScratch& c = scratch(ix);
    uint64_t written = 0;
    const uint64_t slice = 1u << 25;                       // 32 Mi rows x 55 B < 2^31
    for (uint64_t b = 0; b < nocc; b += slice) {
        const uint32_t n = (uint32_t)std::min<uint64_t>(slice, nocc - b);
        c.len.ensure((size_t)n + 1); c.ofs.ensure((size_t)n + 1);
        CUDA_OK(cudaMemsetAsync(c.len.p + n, 0, 4, s));
        csv_len_kernel<<<grid_for(n, 256), 256, 0, s>>>(d_occ + b, n, pat_add, c.len.p);
        size_t tmp = 0;
        CUDA_OK(cub::DeviceScan::ExclusiveSum(nullptr, tmp, c.len.p, c.ofs.p, (int)(n + 1), s));
        c.cub.ensure(tmp);
        CUDA_OK(cub::DeviceScan::ExclusiveSum(c.cub.p, tmp, c.len.p, c.ofs.p, (int)(n + 1), s));
        CUDA_OK(cudaMemcpyAsync(c.h_pin, c.ofs.p + n, 4, cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));
        const uint32_t nbytes = c.h_pin[0];
        c.text.ensure(nbytes); c.host_ensure(nbytes);
        csv_format_kernel<<<grid_for(n, CSV_ROWS), CSV_ROWS, 0, s>>>(d_occ + b, n, pat_add, c.ofs.p, c.text.p);
        CUDA_OK(cudaGetLastError());
        CUDA_OK(cudaMemcpyAsync(c.h_text, c.text.p, nbytes, cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));
        if (fwrite(c.h_text, 1, nbytes, f) != nbytes) EDS_FAIL(EDSBWT_E_IO, "short write to the csv file");
        written += nbytes;
        if (launches) *launches += 4;
    }
    return written;
}

}  // namespace edsbwt
