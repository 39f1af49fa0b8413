// edsbwt_count_static: count-only search of a device-resident batch that ALLOCATES NOTHING.
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  It launches the step-loop search
// kernel (fused.cu, the one every other entry point uses) with all of its scratch -- cursors, the per-pattern interval
// list, the staged result intervals, the two overflow queues -- carved out of ONE file-scope device array of fixed size
// instead of the handle's growable workspace.  Limits: EDSBWT_STATIC_MAX_PATTERNS patterns per call,
// EDSBWT_STATIC_SCRATCH_BYTES of scratch per GPU that loads the library (held until the process exits).
//
// The array is shared by every call on a device and keeps the last call's contents: the cursors are cleared at the start of
// each call, everything else is written before it is read.  Calls are serialised by a per-process lock and return with the
// stream idle; anything else that overlapped on the GPU with the same array would race.
#include "internal.cuh"

using namespace edsbwt;

namespace {

constexpr size_t SCRATCH_BYTES = EDSBWT_STATIC_SCRATCH_BYTES;
constexpr uint32_t MAX_PATTERNS = EDSBWT_STATIC_MAX_PATTERNS;
// layout (bytes): [0,256) cursors | plist 8 B x MAX | ovf_ids 4 B x MAX | ovf_ids2 4 B x MAX | staging: the rest, 8 B per slot
constexpr size_t OFF_PLIST = 256;
constexpr size_t OFF_OVF1 = OFF_PLIST + (size_t)MAX_PATTERNS * 8;
constexpr size_t OFF_OVF2 = OFF_OVF1 + (size_t)MAX_PATTERNS * 4;
constexpr size_t OFF_STAGING = OFF_OVF2 + (size_t)MAX_PATTERNS * 4;
static_assert(OFF_STAGING % 256 == 0 && OFF_STAGING + (size_t)MAX_PATTERNS * 16 <= SCRATCH_BYTES, "scratch layout");
constexpr uint32_t STAGING_SLOTS = (uint32_t)((SCRATCH_BYTES - OFF_STAGING) / 8);

static __device__ __align__(256) unsigned char g_static_scratch[SCRATCH_BYTES];

// In host code g_static_scratch is a stand-in symbol, not a device address: the address is looked up once per device, with
// that device current, and cached.
std::mutex g_mu;
constexpr int MAX_DEVICES = 64;
unsigned char* g_dev_ptr[MAX_DEVICES] = {nullptr};

unsigned char* scratch_of(int device) {
    if (device < 0 || device >= MAX_DEVICES) EDS_FAIL(EDSBWT_E_ARG, "device %d out of range", device);
    if (!g_dev_ptr[device]) {
        CUDA_OK(cudaSetDevice(device));
        void* p = nullptr;
        CUDA_OK(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) EDS_FAIL(EDSBWT_E_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        g_dev_ptr[device] = static_cast<unsigned char*>(p);
    }
    return g_dev_ptr[device];
}

}  // namespace

extern "C" int edsbwt_count_static(edsbwt_index* idx, const uint8_t* d_pat_bytes, const uint64_t* d_pat_off, uint32_t npat,
                                   uint32_t* d_counts, uint32_t* n_unresolved) {
    if (n_unresolved) *n_unresolved = 0;
    if (!idx || (npat && (!d_pat_bytes || !d_pat_off || !d_counts))) { set_error("edsbwt_count_static: null argument"); return EDSBWT_E_ARG; }
    if (npat > MAX_PATTERNS) {
        set_error("edsbwt_count_static: %u patterns in one call, the static scratch holds at most %u (EDSBWT_STATIC_MAX_PATTERNS); split the batch or use edsbwt_search_device",
                  npat, MAX_PATTERNS);
        return EDSBWT_E_NOSPACE;
    }
    if (npat == 0) return EDSBWT_OK;
    try {
        Index& ix = idx->ix;
        std::lock_guard<std::mutex> all(g_mu);                 // one call at a time: the scratch is shared
        std::lock_guard<std::mutex> lock(ix.mu);               // ... and the handle's stream
        CUDA_OK(cudaSetDevice(ix.device));
        unsigned char* base = scratch_of(ix.device);
        cudaStream_t s = ix.stream;
        uint32_t* cursors = reinterpret_cast<uint32_t*>(base);
        uint2* plist = reinterpret_cast<uint2*>(base + OFF_PLIST);
        uint32_t* ovf1 = reinterpret_cast<uint32_t*>(base + OFF_OVF1);
        uint32_t* ovf2 = reinterpret_cast<uint32_t*>(base + OFF_OVF2);
        uint2* staging = reinterpret_cast<uint2*>(base + OFF_STAGING);
        CUDA_OK(cudaMemsetAsync(cursors, 0, 256, s));          // [0] staging slots used, [1] / [2] patterns overflowed in pass 1 / 2
        SeedView sv{};
        for (uint32_t L = 1; L <= ix.seed_k; ++L) { sv.off[L] = ix.seed[L]->off.p; sv.iv[L] = ix.seed[L]->iv.p; sv.cnt[L] = ix.seed[L]->cnt.p; }
        sv.first = ix.seed[ix.seed_k]->first.p; sv.first_level = ix.seed_k;
        FusedOut fo{plist, d_counts, staging, STAGING_SLOTS, cursors, cursors + 1, ovf1, nullptr};
        launch_fused_search(ix.view, sv, ix.seed_k, d_pat_bytes, d_pat_off, nullptr, npat, ix.lut.p, ix.fused_cap, fo, nullptr, ix.canonical_k, s);
        uint32_t h[4] = {0, 0, 0, 0};
        CUDA_OK(cudaMemcpyAsync(h, cursors, 16, cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));
        uint32_t left = h[1];
        if (left && ix.fused_cap < ix.fused_cap2) {
            // the patterns whose frontier outgrew the first pass: once more with the largest frontier one CTA can hold
            FusedOut fo2{plist, d_counts, staging, STAGING_SLOTS, cursors, cursors + 2, ovf2, nullptr};
            launch_fused_search(ix.view, sv, ix.seed_k, d_pat_bytes, d_pat_off, ovf1, left, ix.lut.p, ix.fused_cap2, fo2, nullptr, ix.canonical_k, s);
            CUDA_OK(cudaMemcpyAsync(h, cursors, 16, cudaMemcpyDeviceToHost, s));
            CUDA_OK(cudaStreamSynchronize(s));
            left = h[2];
        }
        CUDA_OK(cudaGetLastError());
        if (left) {
            // what the other entry points hand to the level-synchronous engine needs growable buffers: not available here
            if (n_unresolved) *n_unresolved = left;
            set_error("edsbwt_count_static: %u of %u patterns need more scratch than the static array holds (their counts read 0); use edsbwt_search_device for them",
                      left, npat);
            return EDSBWT_E_NOSPACE;
        }
        return EDSBWT_OK;
    }
    catch (const Failure& e) { return e.code; }
    catch (const std::exception& e) { set_error("edsbwt_count_static: %s", e.what()); return EDSBWT_E_IO; }
}
