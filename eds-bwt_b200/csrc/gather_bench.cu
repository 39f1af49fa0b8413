// K7: random-gather microbenchmark = the random-access roofline denominator of the search kernels.
//
// Round 1's version (4-lane groups, 128-bit loads, 4 loads in flight) reported less than the search kernel itself sustains, so it
// was not a ceiling.  This one uses the search kernels' own access -- ONE thread fetches a whole 64-byte block with two
// ld.global.nc.L1::no_allocate.v8.b32 (internal.cuh ld256) -- and sweeps what decides the throughput of random access:
//   ilp        independent blocks in flight per thread (1, 2, 4, 8)
//   ctas/SM    resident CTAs of 256 threads (0 = as many as fit)
//   dependent  1: the next address of a chain is computed from the loaded data, like a backward search (the memory-level
//              parallelism is then threads x ilp chains); 0: addresses are known up front
//   mix        besides the blocks, a fraction of the accesses are 16-byte gathers over a second, larger buffer (the seed /
//              link-record / locate gathers of the search: 165 M block fetches over the 142 MB Occ array + ~30 M small gathers
//              over gigabytes at cfg4)
// The ceiling quoted by bench.py is the best independent configuration over a buffer of the size of the Occ array, and the
// best mix; the dependent rows say how much of it a chain-following kernel can reach at a given occupancy.
#include "internal.cuh"

namespace edsbwt {

namespace {

__device__ __forceinline__ uint64_t mix64(uint64_t h) {
    h *= 0x9E3779B97F4A7C15ull; h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
    return h;
}
__device__ __forceinline__ void ld128nc(const void* p, uint32_t* r) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "l"(p));
}

template <int ILP, bool DEP, int GRAN>      // GRAN: 64 (one Occ block) or 32 (one sector)
__global__ void __launch_bounds__(256)
gather_kernel(const uint4* __restrict__ buf, uint64_t nblocks, uint32_t per_thread, uint32_t seed, const uint4* __restrict__ side, uint64_t nside,
              uint32_t side_per_1024, uint32_t* __restrict__ sink) {
    const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t h[ILP];
    #pragma unroll
    for (int i = 0; i < ILP; ++i) h[i] = mix64(gid * ILP + i + ((uint64_t)seed << 40));
    uint32_t acc = 0;
    for (uint32_t it = 0; it < per_thread; ++it) {
        uint32_t r[ILP][16], sd[ILP][4];
        bool want_side[ILP];
        #pragma unroll
        for (int i = 0; i < ILP; ++i) {
            const uint64_t blk = ((h[i] >> 32) * nblocks) >> 32;            // nblocks < 2^32: multiply-shift instead of a 64-bit modulo
            const uint4* p = buf + blk * (GRAN / 16);
            ld256(p, r[i]);
            if (GRAN == 64) ld256(p + 2, r[i] + 8);
            want_side[i] = side_per_1024 && ((uint32_t)h[i] & 1023u) < side_per_1024;
            if (want_side[i]) ld128nc(side + ((((h[i] >> 11) & 0xffffffffull) * nside) >> 32), sd[i]);
        }
        #pragma unroll
        for (int i = 0; i < ILP; ++i) {
            uint32_t x = 0;
            #pragma unroll
            for (int w = 0; w < (GRAN == 64 ? 16 : 8); ++w) x ^= r[i][w];
            if (want_side[i]) x ^= sd[i][0] ^ sd[i][1] ^ sd[i][2] ^ sd[i][3];
            acc ^= x;
            h[i] = DEP ? mix64(h[i] ^ x) : mix64(h[i] + 0x632BE59BD9B4E019ull);
        }
    }
    if (acc == 0x12345678u) sink[0] = acc;      // keep the loads alive
}

template <int ILP, bool DEP, int GRAN>
float run_one(const GatherCfg& c, const uint4* buf, uint64_t nblocks, const uint4* side, uint64_t nside, uint32_t* sink, int sms, int it, uint64_t& done,
              cudaEvent_t e0, cudaEvent_t e1) {
    auto kern = gather_kernel<ILP, DEP, GRAN>;
    int per_sm = 1;
    CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, 256, 0));
    if (c.ctas_per_sm > 0 && c.ctas_per_sm < per_sm) per_sm = c.ctas_per_sm;
    const unsigned grid = (unsigned)(sms * per_sm);
    const uint64_t threads = (uint64_t)grid * 256;
    uint32_t per_thread = (uint32_t)std::max<uint64_t>(1, c.n_access / (threads * ILP));
    done = threads * ILP * per_thread;
    const uint32_t side_per_1024 = side ? (uint32_t)(c.mix_frac * 1024.0 + 0.5) : 0u;
    CUDA_OK(cudaEventRecord(e0));
    kern<<<grid, 256>>>(buf, nblocks, per_thread, (uint32_t)it * 7919u + 1u, side, nside, side_per_1024, sink);
    CUDA_OK(cudaEventRecord(e1));
    CUDA_OK(cudaEventSynchronize(e1));
    float ms = 0; CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
    return ms;
}

}  // namespace

// -> algorithmic GB/s (64 or 32 B per block + 16 B per side gather); *req_per_s = memory requests per second
double gather_bench_cfg(int device, const GatherCfg& c, double* req_per_s) {
    CUDA_OK(cudaSetDevice(device));
    if (c.granule != 32 && c.granule != 64) EDS_FAIL(EDSBWT_E_ARG, "gather_bench: granule must be 32 or 64");
    if (c.ilp != 1 && c.ilp != 2 && c.ilp != 4 && c.ilp != 8) EDS_FAIL(EDSBWT_E_ARG, "gather_bench: ilp must be 1, 2, 4 or 8");
    const uint64_t nblocks = c.buffer_bytes / c.granule;
    if (nblocks == 0 || nblocks > 0xffffffffull || c.n_access == 0) EDS_FAIL(EDSBWT_E_ARG, "gather_bench: bad buffer size");
    const uint64_t nside = c.mix_bytes / 16;
    if (nside > 0xffffffffull) EDS_FAIL(EDSBWT_E_ARG, "gather_bench: side buffer too large");
    DevBuf<uint4> buf, side; buf.exact(nblocks * (c.granule / 16));
    if (nside) side.exact(nside);
    DevBuf<uint32_t> sink; sink.ensure(4);
    CUDA_OK(cudaMemset(buf.p, 1, nblocks * c.granule));
    if (nside) CUDA_OK(cudaMemset(side.p, 2, nside * 16));
    cudaEvent_t e0, e1; CUDA_OK(cudaEventCreate(&e0)); CUDA_OK(cudaEventCreate(&e1));
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    float best = 1e30f; uint64_t done = 0;
    for (int it = 0; it < c.iters + 2; ++it) {
        float ms;
#define RUN(ILP) (c.dependent ? (c.granule == 64 ? run_one<ILP, true, 64>(c, buf.p, nblocks, side.p, nside, sink.p, sms, it, done, e0, e1) \
                                                 : run_one<ILP, true, 32>(c, buf.p, nblocks, side.p, nside, sink.p, sms, it, done, e0, e1)) \
                              : (c.granule == 64 ? run_one<ILP, false, 64>(c, buf.p, nblocks, side.p, nside, sink.p, sms, it, done, e0, e1) \
                                                 : run_one<ILP, false, 32>(c, buf.p, nblocks, side.p, nside, sink.p, sms, it, done, e0, e1)))
        if (c.ilp == 1) ms = RUN(1); else if (c.ilp == 2) ms = RUN(2); else if (c.ilp == 4) ms = RUN(4); else ms = RUN(8);
#undef RUN
        if (it >= 2 && ms < best) best = ms;
    }
    CUDA_OK(cudaGetLastError());
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    const double nside_acc = nside ? (double)done * (double)(uint32_t)(c.mix_frac * 1024.0 + 0.5) / 1024.0 : 0.0;
    if (req_per_s) *req_per_s = ((double)done + nside_acc) / (best * 1e-3);
    return ((double)done * c.granule + nside_acc * 16.0) / (best * 1e-3) / 1e9;
}

// round-1 signature (bench.py, profiles/gather_sweep.py): independent accesses, 4 in flight per thread, full occupancy
double gather_bench(int device, uint64_t buffer_bytes, uint64_t n_access, int iters, int granule) {
    if (granule == 128) EDS_FAIL(EDSBWT_E_ARG, "gather_bench: granule must be 32 or 64");
    GatherCfg c{}; c.buffer_bytes = buffer_bytes; c.n_access = n_access; c.iters = iters; c.granule = granule; c.ilp = 4;
    return gather_bench_cfg(device, c, nullptr);
}

}  // namespace edsbwt
