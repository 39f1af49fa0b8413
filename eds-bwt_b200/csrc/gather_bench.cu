// K7: 64-byte random-gather microbenchmark = the roofline denominator for the rank kernels.
// Same access shape as lf_step_kernel: a 4-lane group fetches one 64-byte block with one 128-bit load per
// lane; block addresses are pseudo-random over the whole buffer.
#include "internal.cuh"

namespace edsbwt {

template <int LANES>        // LANES x 16 bytes per access: 2 = 32 B (one sector), 4 = 64 B (one Occ block), 8 = 128 B (one line)
__global__ void __launch_bounds__(256) gather_kernel(const uint4* __restrict__ buf, uint64_t nblocks, uint64_t n_access, uint32_t seed, uint32_t* __restrict__ sink) {
    uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    uint32_t acc = 0;
    // 4 independent gathers in flight per thread (the search kernels get their memory-level parallelism from
    // occupancy; the denominator should not be handicapped by a dependent loop)
    #pragma unroll 4
    for (uint64_t g = gid; g < n_access * LANES; g += stride) {
        uint64_t i = g / LANES; uint32_t lane4 = (uint32_t)(g % LANES);
        uint64_t h = (i + seed) * 0x9E3779B97F4A7C15ull; h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
        uint64_t blk = ((h >> 32) * nblocks) >> 32;         // nblocks < 2^32: multiply-shift instead of a 64-bit modulo
        uint4 v = __ldg(buf + blk * LANES + lane4);
        acc += v.x ^ v.y ^ v.z ^ v.w;
    }
    if (acc == 0x12345678u) sink[0] = acc;      // keep the loads alive
}

double gather_bench(int device, uint64_t buffer_bytes, uint64_t n_access, int iters, int granule) {
    CUDA_OK(cudaSetDevice(device));
    if (granule != 32 && granule != 64 && granule != 128) EDS_FAIL(EDSBWT_E_ARG, "gather_bench: granule must be 32, 64 or 128");
    uint64_t nblocks = buffer_bytes / granule;
    if (nblocks == 0 || n_access == 0) EDS_FAIL(EDSBWT_E_ARG, "gather_bench: empty buffer");
    DevBuf<uint4> buf; buf.ensure(nblocks * (granule / 16));
    DevBuf<uint32_t> sink; sink.ensure(4);
    CUDA_OK(cudaMemset(buf.p, 1, nblocks * granule));
    cudaEvent_t e0, e1; CUDA_OK(cudaEventCreate(&e0)); CUDA_OK(cudaEventCreate(&e1));
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    unsigned grid = (unsigned)sms * 8 * 4;
    float best = 1e30f;
    for (int it = 0; it < iters + 2; ++it) {
        CUDA_OK(cudaEventRecord(e0));
        if (granule == 32) gather_kernel<2><<<grid, 256>>>(buf.p, nblocks, n_access, (uint32_t)it * 7919u, sink.p);
        else if (granule == 64) gather_kernel<4><<<grid, 256>>>(buf.p, nblocks, n_access, (uint32_t)it * 7919u, sink.p);
        else gather_kernel<8><<<grid, 256>>>(buf.p, nblocks, n_access, (uint32_t)it * 7919u, sink.p);
        CUDA_OK(cudaEventRecord(e1));
        CUDA_OK(cudaEventSynchronize(e1));
        float ms = 0; CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
        if (it >= 2 && ms < best) best = ms;
    }
    CUDA_OK(cudaGetLastError());
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return (double)n_access * (double)granule / (best * 1e-3) / 1e9;
}

}  // namespace edsbwt
