/* edsbwt.h -- C ABI of libedsbwt_cuda.so: B200-native batched pattern search over the EDS-BWT index.
 *
 * The reference (riccardo-nozza/EDS-BWT) has no library API: its operator surface is two executables whose
 * constructors run the whole job.  Each entry point below names the reference code
 * it replaces, so that a maintainer can bind it from the reference's own main().
 * Plain pointers and sizes only; no C++/torch types.  All functions return 0 on success or a negative
 * EDSBWT_E_* code and never call exit() (the reference exits on every error: EDSBWTsearch.cpp:158-161,
 * 1562-1565).  There is NO CPU fallback: without a CUDA device every compute entry point fails with
 * EDSBWT_E_CUDA.
 *
 * Threads: the index data of a handle are immutable after open, but a handle owns ONE compute stream, one
 * workspace and one chunk pipeline, so calls on the same handle are serialised inside the library (safe from
 * any number of threads, not concurrent); edsbwt_set_* and edsbwt_last_stats refer to "the last call" and are
 * the caller's to order.  For concurrency open one handle per GPU (the reference object is not re-entrant at
 * all: per-pattern state lives in members, EDSBWTsearch.hpp:61-63).  edsbwt_last_error is thread-local.
 */
#ifndef EDSBWT_H
#define EDSBWT_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define EDSBWT_OK            0
#define EDSBWT_E_IO         -1   /* a file of the index / pattern set could not be opened or is truncated   */
#define EDSBWT_E_FORMAT     -2   /* inconsistent index files (sizes, tableOcc, _bv_ vs _bwt_, .bitvector)    */
#define EDSBWT_E_ALPHABET   -3   /* more than 7 letters besides the end-marker (2-bit / 3-bit Occ layouts)    */
#define EDSBWT_E_CUDA       -4   /* no device / CUDA runtime error / out of device memory                    */
#define EDSBWT_E_ARG        -5   /* bad argument                                                             */
#define EDSBWT_E_UNSUPPORTED -6
#define EDSBWT_E_NOSPACE     -7   /* caller's occurrence buffer too small; *nocc holds the size needed             */

typedef struct edsbwt_index edsbwt_index;     /* opaque, immutable after open; one handle per GPU */
typedef struct edsbwt_result edsbwt_result;   /* opaque, device-resident result of edsbwt_search_device */

/* One occurrence = one row of the reference's <patterns>output.csv (EDSBWTsearch.cpp:1320-1327):
 *   #Pat  $_i  D[i]  S_j  S_j[r]
 * pat is the 0-based index of the pattern inside the call (the csv prints pat+1 = the 1-based line number). */
typedef struct {
    uint32_t pat;         /* pattern index within the batch                                   */
    uint32_t str;         /* $_i : 0-based string index                                        */
    uint32_t seg;         /* D[i]: 1-based degenerate segment                                  */
    uint32_t idx_in_seg;  /* S_j : 0-based index of the string inside its segment              */
    uint32_t off;         /* S_j[r]: 0-based offset of the occurrence inside the string        */
} edsbwt_occ;

typedef struct {
    uint32_t n;                 /* eBWT length incl. end-markers (info.aux lengthTot_plus_eof) */
    uint32_t n_text;            /* number of strings                                           */
    uint32_t n_seg;             /* number of degenerate segments                               */
    uint32_t sigma;             /* alphabet size incl. the end-marker                          */
    uint8_t  alphabet[8];       /* alphaInverse[], alphabet[0] == '#'                          */
    uint32_t table_occ[25];     /* tableOcc[sigma][sigma], row-major (recoverInfo dump), sigma <= 5; zero otherwise */
    uint32_t seed_k;            /* depth of the seed table actually built                      */
    uint32_t locate_mode;       /* EDSBWT_LOCATE_*                                             */
    uint64_t device_bytes;      /* HBM held by the handle                                      */
    uint64_t seed_intervals;    /* total intervals stored in the seed table                    */
    uint32_t move_runs;         /* move table: runs of the eBWT (every end-marker its own run); 0 = table not built */
    uint32_t move_intervals;    /* move table: input intervals after a-balancing (Move-r's r')   */
    uint32_t move_a;            /* balancing parameter                                          */
    uint32_t move_rounds;       /* balancing rounds the builder needed                          */
    uint32_t from_file;         /* bit 0: the HBM image was read from <base>.edsbwt; bit 1: the move table from <base>_MLF.edsbwt */
    uint32_t engine;            /* EDSBWT_ENGINE_* in force (an alphabet of 5..7 letters never runs EDSBWT_ENGINE_ITEMS)           */
    uint32_t table_occ_wide[64];/* tableOcc[sigma][sigma], row-major, for every sigma <= 8                                          */
} edsbwt_info;

#define EDSBWT_LOCATE_SA    0   /* row -> (string, offset) table in HBM: one gather per occurrence           */
#define EDSBWT_LOCATE_WALK  1   /* LF-walk each row to its string start (the reference's a9/a13 arithmetic)  */
#define EDSBWT_ENGINE_FUSED      0  /* default: whole backward search in one launch, the kernel chosen by the form the batch    */
                                    /* arrives in -- bytes: the step-loop kernel (one thread per pattern, fused.cu); 2-bit     */
                                    /* words: the item kernel (items.cu).  The two are within 3 % of each other at cfg4; the    */
                                    /* item kernel would need a packing pass over bytes first.  Frontiers that  */
                                    /* outgrow the first pass go to the step-loop kernel with 192 slots and then to:            */
#define EDSBWT_ENGINE_LEVELSYNC  1  /* level-synchronous interval-parallel engine (also builds the seed table)                  */
#define EDSBWT_ENGINE_FUSED_STEP 2  /* step-loop kernel as the first pass for every batch (2-bit words are unpacked on the device) */
#define EDSBWT_ENGINE_ITEMS      3  /* item kernel as the first pass for every batch: persistent warps, one frontier item per   */
                                    /* lane per sweep, finished lanes refilled from a warp-wide pool (bytes are packed first)   */
#define EDSBWT_LF_RANK      0   /* EDSBWTsearch:      LF by blocked rank                                      */
#define EDSBWT_LF_MOVE      1   /* MOVE_EDSBWTSearch: LF by move table                                        */

#define EDSBWT_ROWS_GROUPED   0   /* tuples grouped by ascending pattern, order inside a pattern unspecified        */
#define EDSBWT_ROWS_REFERENCE 1   /* inside a pattern: the reference's csv order (EDSBWTsearch: round-major, then    */
                                  /* eBWT position of the string-start row; MOVE: frontier order, ascending row)     */

typedef struct {
    int32_t  seed_k;            /* -1 = auto: floor(log4 n), at most 14; values below 1 are raised to 1        */
    int32_t  locate_mode;       /* EDSBWT_LOCATE_SA / _WALK                                                   */
    int32_t  verify;            /* 1 = cross-check _bv_<p>.aux (rrr) and tableOcc against the pile bytes      */
    int32_t  build_move;        /* 1 = also build the move table (EDSBWT_LF_MOVE searches)                    */
    int32_t  engine;            /* EDSBWT_ENGINE_FUSED (default) / _LEVELSYNC / _FUSED_STEP / _ITEMS          */
    int32_t  fused_cap;         /* frontier slots per pattern of the first pass (0 = default 13)                */
    int32_t  fused_cap2;        /* slots of the second pass over the patterns that outgrew the first (0 = 192) */
    int32_t  move_a;            /* balancing parameter a of the move table (build_MLF's <a>); 0 = default 8 (or whatever a valid <base>_MLF.edsbwt was built with) */
    int32_t  cache;             /* 1 = keep the whole HBM image in <base>.edsbwt: read it when it is there and was written for these very index */
                                /* files (sizes, times and sampled contents are fingerprinted), otherwise build and write it.  The file is as large */
                                /* as the image (gigabytes: seed and locate tables are derived data) -- worth it when it stays in the page cache   */
} edsbwt_open_opts;

/* Canonical byte-model counters of DESIGN.md + engine statistics of the last search. */
typedef struct {
    uint64_t I, S, E, L, O;     /* seed intervals, interval-steps, end-markers enumerated, locate row-steps, occurrences */
    uint64_t steps;             /* level-synchronous backward steps executed                                   */
    uint64_t kernel_launches;   /* kernels of this library launched by the last search                         */
    uint64_t lf_intervals;      /* intervals LF-mapped by the rank kernel (all steps, incl. those inside the seed range) */
    double   lf_kernel_ms;      /* CUDA-event time summed over the launches of the LF rank kernel              */
    double   locate_kernel_ms;  /* CUDA-event time of the locate kernel                                        */
    double   total_ms;          /* CUDA-event time of the whole search on the library's stream                 */
    uint64_t frontier_peak;     /* largest frontier (intervals) over all steps of the level-synchronous engine  */
    uint64_t occ_blocks;        /* 64-byte Occ blocks fetched by the search kernels (collect_counters only)     */
    uint64_t overflow_patterns; /* patterns whose frontier outgrew the first pass of the fused kernel          */
    uint64_t levelsync_patterns;/* patterns finally run by the level-synchronous engine                        */
    double   search_kernel_ms;  /* CUDA-event time of the fused search kernel                                   */
} edsbwt_stats;

/* --- index load: replaces EDSBWT::recoverInfo (EDSBWTsearch.cpp:1557-1656), EDSBWT::computeVectorOcc
 *     (:1338-1442) and the .bitvector rank/select set-up of the constructor (:150-168).
 *     base is the reference's <indexBase>: reads base_info.aux, base_bwt_<p>.aux, base_bv_<p>.aux, base.bitvector. */
int  edsbwt_index_open(const char* base, int device, edsbwt_index** out);
int  edsbwt_index_open_ex(const char* base, int device, const edsbwt_open_opts* opts, edsbwt_index** out);
int  edsbwt_index_info(const edsbwt_index* idx, edsbwt_info* info);
void edsbwt_index_close(edsbwt_index* idx);

/* --- search: replaces the per-pattern loop of the EDSBWT constructor (EDSBWTsearch.cpp:207-241) with one
 *     batched call: EDSBWT::backwardSearch (:619-772), link (:1212-1307), updateIntervals (:956-1157) and,
 *     when locate != 0, findMultipleDollarsBackward + indices_of_dollars_in_interval (:300-616, 1310-1332);
 *     lf_variant = EDSBWT_LF_MOVE is MOVE_EDSBWT::backwardSearch (MOVE_EDSBWTSearch.cpp:228-374).
 *     Pattern i is pat_bytes[pat_off[i] .. pat_off[i+1]) (pat_off has npat+1 entries); HOST pointers.
 *     counts[i] = number of distinct starting positions ("Found N ...", :739-746).
 *     locate != 0: *occ receives a malloc'ed array of *nocc tuples grouped by ascending pat (order inside a
 *     pattern unspecified; the occurrence SET equals the reference's); release with edsbwt_free.
 *     Patterns holding a byte outside the index alphabet (or the end-marker) yield 0 (reference: undefined
 *     behaviour); empty patterns yield 0. */
int  edsbwt_search(edsbwt_index* idx, const uint8_t* pat_bytes, const uint64_t* pat_off, uint32_t npat,
                   int locate, int lf_variant, uint32_t* counts, edsbwt_occ** occ, uint64_t* nocc);

/* The same search into CALLER-OWNED host buffers (no allocation inside the call): counts[npat], occ[occ_cap].
 * The batch is cut into chunks that are copied in, searched and copied out on three streams, so that PCIe
 * traffic overlaps the kernels; buffers from edsbwt_host_alloc (pinned) make the copies truly asynchronous.
 * Returns EDSBWT_E_NOSPACE when more than occ_cap occurrences exist (*nocc = number needed, counts valid). */
int  edsbwt_search_into(edsbwt_index* idx, const uint8_t* pat_bytes, const uint64_t* pat_off, uint32_t npat,
                        int locate, int lf_variant, uint32_t* counts, edsbwt_occ* occ, uint64_t occ_cap, uint64_t* nocc);
/* The same for a batch of EQUAL-LENGTH patterns stored back to back (a k-mer file such as sample/test/kmers.txt without
 * its newlines): pattern i = pat_bytes[i*pat_len .. (i+1)*pat_len).  No offset array crosses PCIe. */
int  edsbwt_search_into_fixed(edsbwt_index* idx, const uint8_t* pat_bytes, uint32_t pat_len, uint32_t npat,
                              int locate, int lf_variant, uint32_t* counts, edsbwt_occ* occ, uint64_t occ_cap, uint64_t* nocc);
/* The same for patterns the caller already holds as 2-bit codes (a quarter of the bytes over PCIe: 8 B instead of 32 B
 * per length-32 pattern).  Layout (the kernel's own, internal.cuh PackedPats): pattern i occupies W = ceil(pat_len / 32)
 * consecutive uint64 words; word w holds characters [max(0, pat_len - 32(w+1)), pat_len - 32w) of the pattern, the LAST
 * of them in bits 62..63 and each earlier one two bits lower (a short head is padded with zero bits below); code =
 * rank of the letter in the index alphabet (edsbwt_index_info: alphabet[1 + code]).  For pat_len <= 32 that is one
 * word per pattern with character j at bits 2(j + 32 - pat_len).  invalid (may be NULL): bit i & 31 of word i >> 5 set =
 * pattern i holds a byte outside the alphabet and yields 0 occurrences.  edsbwt_pack_patterns produces both from bytes
 * on the host (the std::getline loop of the reference's constructor, EDSBWTsearch.cpp:207-241, is where a caller would
 * do it: once per pattern file, not once per search). */
int  edsbwt_search_into_packed(edsbwt_index* idx, const uint64_t* pat_words, const uint32_t* invalid, uint32_t pat_len, uint32_t npat,
                               int locate, int lf_variant, uint32_t* counts, edsbwt_occ* occ, uint64_t occ_cap, uint64_t* nocc);
/* Host-side packer for the call above (AVX2 + BMI2 when the CPU has them, scalar otherwise; nthreads <= 1: the calling
 * thread only).  words: npat * ceil(pat_len / 32) uint64; invalid: (npat + 31) / 32 uint32; *n_invalid (may be NULL)
 * receives the number of patterns flagged.  Needs the index only for its alphabet; no device work. */
int  edsbwt_pack_patterns(const edsbwt_index* idx, const uint8_t* pat_bytes, uint32_t pat_len, uint32_t npat, int nthreads,
                          uint64_t* words, uint32_t* invalid, uint32_t* n_invalid);
/* The same without a handle: alphabet[0] = end-marker, alphabet[1..sigma) = the letters (edsbwt_info.alphabet / sigma). */
int  edsbwt_pack_patterns_alphabet(const uint8_t* alphabet, uint32_t sigma, const uint8_t* pat_bytes, uint32_t pat_len, uint32_t npat,
                                   int nthreads, uint64_t* words, uint32_t* invalid, uint32_t* n_invalid);
/* Count-only search that also reports where each not-found pattern died, for the RECOVERBW=0 stderr line
 * "Pattern not found, interrupted in position <posSymb>" (EDSBWTsearch.cpp:714-717): interrupted[i] = that posSymb
 * (the frontier became empty when character posSymb-1 of the pattern was prepended), 0 = no such line (pattern
 * found, single-character pattern, or pattern rejected for a byte outside the alphabet). */
int  edsbwt_search_trace(edsbwt_index* idx, const uint8_t* pat_bytes, const uint64_t* pat_off, uint32_t npat,
                         int lf_variant, uint32_t* counts, uint32_t* interrupted);
int  edsbwt_host_alloc(void** p, uint64_t bytes);      /* page-locked host memory for the buffers above */
void edsbwt_host_free(void* p);
int  edsbwt_set_chunk(edsbwt_index* idx, uint32_t patterns_per_chunk);   /* pipeline granularity, default 2^21; 0 = one chunk */

/* Same search with the patterns (bytes + offsets) already resident in HBM and the results left there (the `value`
 * leg of bench.py).  d_* are DEVICE pointers on the handle's device.  total_bytes >= pat_off[npat] - pat_off[0] (it sizes
 * the 2-bit copy of the batch; a smaller value fails with EDSBWT_E_ARG); max_len = length of the longest pattern, or 0
 * when unknown (found on the device if needed).  The kernels read pattern bytes in aligned 16-byte units: the 16-byte
 * unit holding the last pattern byte must lie inside the allocation (true for any cudaMalloc'ed buffer, whose size is
 * rounded up to 256 bytes; a sub-allocated buffer needs 15 bytes of slack at its end). */
int  edsbwt_search_device(edsbwt_index* idx, const uint8_t* d_pat_bytes, const uint64_t* d_pat_off, uint32_t npat,
                          uint64_t total_bytes, uint32_t max_len, int locate, int lf_variant, edsbwt_result** res);
/* The same for one shard of a larger batch (multi-GPU: patterns sharded, index replicated): tuples carry pat + pat_base. */
int  edsbwt_search_device_at(edsbwt_index* idx, const uint8_t* d_pat_bytes, const uint64_t* d_pat_off, uint32_t npat,
                             uint64_t total_bytes, uint32_t max_len, int locate, int lf_variant, uint32_t pat_base, edsbwt_result** res);
int  edsbwt_result_sizes(const edsbwt_result* res, uint32_t* npat, uint64_t* nocc);
int  edsbwt_result_device_ptrs(const edsbwt_result* res, const uint32_t** d_counts, const edsbwt_occ** d_occ);
int  edsbwt_result_fetch(const edsbwt_result* res, uint32_t* counts, edsbwt_occ* occ); /* D2H into host buffers */
void edsbwt_result_free(edsbwt_result* res);

/* --- multi-GPU, one process (replaces the single EDSBWT object on a single thread of mainEDS-BWT.cpp:66-70):
 *     the index is replicated on every listed device (devices == NULL / ndev == 0: all visible devices), patterns are sharded
 *     in contiguous blocks, every shard is searched in chunks_per_gpu chunks (<= 0: 4) and the per-pattern counts and the
 *     occurrence tuples of chunk c are gathered on devices[0] while chunk c + 1 is searched: ncclSend / ncclRecv inside one
 *     ncclGroupStart/End per chunk (NCCL is bound at run time; without libnccl, or with EDSBWT_MULTI_P2P=1, peer copies over
 *     NVLink are used instead).  pat_off == NULL: npat patterns of fixed_len bytes back to back.  HOST pointers.
 *     The result lives in the HBM of devices[0]: counts[npat] and one tuple segment per (shard, chunk), in ascending pattern
 *     order; tuples carry batch-global pattern numbers. */
typedef struct edsbwt_multi edsbwt_multi;
typedef struct edsbwt_multi_result edsbwt_multi_result;
int  edsbwt_multi_open(const char* base, const int* devices, int ndev, const edsbwt_open_opts* opts, edsbwt_multi** out);
int  edsbwt_multi_info(const edsbwt_multi* m, int* ndev, int* devices, int devices_cap, int* nccl /* 1: NCCL in use */);
edsbwt_index* edsbwt_multi_handle(edsbwt_multi* m, int i);      /* the i-th device's handle (edsbwt_set_row_order, edsbwt_set_stats ...) */
void edsbwt_multi_close(edsbwt_multi* m);
int  edsbwt_search_multi(edsbwt_multi* m, const uint8_t* pat_bytes, const uint64_t* pat_off, uint32_t fixed_len, uint32_t npat,
                         int locate, int lf_variant, int chunks_per_gpu, edsbwt_multi_result** res);
/* search_ms: largest per-device sum of search times (CUDA events); gather_ms: first gather issued -> last gather landed
 * on devices[0] (CUDA events on its communication stream; overlaps the searches); wall_ms: the whole call;
 * gather_bytes: bytes that crossed NVLink. */
int  edsbwt_multi_result_info(const edsbwt_multi_result* r, uint32_t* npat, uint64_t* nocc, uint32_t* nseg, double* search_ms,
                              double* gather_ms, double* wall_ms, uint64_t* gather_bytes);
int  edsbwt_multi_result_device_ptrs(const edsbwt_multi_result* r, const uint32_t** d_counts, const edsbwt_occ** seg_occ,
                                     uint64_t* seg_nocc, uint32_t* seg_pat_lo, uint32_t seg_cap);
int  edsbwt_multi_result_fetch(const edsbwt_multi_result* r, uint32_t* counts, edsbwt_occ* occ);   /* D2H, ascending pattern order */
int  edsbwt_multi_result_write_csv(edsbwt_multi* m, const edsbwt_multi_result* r, const char* path, int move, int append, uint64_t* bytes);
void edsbwt_multi_result_free(edsbwt_multi_result* r);

/* --- multi-GPU, one process per GPU (torchrun and the like): the gather over peer memory.  The root process allocates an arena
 *     in its HBM (edsbwt_device_alloc) and exports it (edsbwt_ipc_export, a 64-byte CUDA IPC handle that travels by any means);
 *     the others map it (edsbwt_ipc_open) and push every finished chunk of their shard -- counts and tuples of a device-resident
 *     result -- to its place in the arena with asynchronous peer copies over NVLink (edsbwt_result_push: copy engines, a stream of
 *     its own, returns at once; the result must stay alive until edsbwt_push_wait).  No sizes are exchanged first and the root
 *     posts nothing, so the transfer of chunk c runs under the search of chunk c + 1 (eds-bwt_b200/sharded.py PeerGather). */
int  edsbwt_device_alloc(int device, uint64_t bytes, void** d_ptr);
void edsbwt_device_free(int device, void* d_ptr);
int  edsbwt_ipc_export(int device, const void* d_ptr, uint8_t handle[64]);
int  edsbwt_ipc_open(int device, const uint8_t handle[64], void** d_ptr);
int  edsbwt_ipc_close(int device, void* d_ptr);
int  edsbwt_result_push(edsbwt_index* idx, const edsbwt_result* res, uint32_t* d_counts_dst, edsbwt_occ* d_occ_dst);
int  edsbwt_push_wait(edsbwt_index* idx);

int  edsbwt_set_engine(edsbwt_index* idx, int engine);      /* EDSBWT_ENGINE_* for every later search on this handle (A/B measurements) */
int  edsbwt_last_stats(const edsbwt_index* idx, edsbwt_stats* st);
int  edsbwt_set_stats(edsbwt_index* idx, int collect_counters, uint32_t canonical_k);

/* --- move table (replaces build_MLF.cpp:53-164 + MOVE_EDSBWT::retrieve_MLF, MOVE_EDSBWTSearch.cpp:178-218): built on the
 *     GPU from the pile files, at open (opts.build_move) or by the first EDSBWT_LF_MOVE search.  edsbwt_move_table_fetch
 *     copies the columns to host arrays of `cap` entries (test / inspection hook): p = first row of input interval i,
 *     q = LF(p), dest = input interval containing q, sym = alphabet byte of the run.  *r = number of intervals. */
int  edsbwt_move_table_build(edsbwt_index* idx);
/* Writes the table to path (NULL: <base>_MLF.edsbwt, our documented format next to the reference's <base>_MLF.aux: csrc/cache.cu);
 * every later open of the same index files loads it instead of building (MOVE_EDSBWT::retrieve_MLF, MOVE_EDSBWTSearch.cpp:186-192).
 * The executable build_MLF <index> <a> is this call (build_MLF.cpp:53-164). */
int  edsbwt_move_table_save(edsbwt_index* idx, const char* path);
int  edsbwt_move_table_fetch(edsbwt_index* idx, uint32_t* p, uint32_t* q, uint32_t* dest, uint8_t* sym, uint32_t cap, uint32_t* r);

/* --- output: replaces the csv writing of EDSBWT's constructor (:179-193) and of
 *     indices_of_dollars_in_interval (:1327).  Creates/truncates path, writes the header
 *     "#Pat\t$_i\tD[i]\tS_j\tS_j[r]\n" (move != 0: the MOVE header ending "S_j[r] \n",
 *     MOVE_EDSBWTSearch.cpp:55-59) and one row per tuple with pat + pat_base + 1. */
int  edsbwt_write_csv(const char* path, const edsbwt_occ* occ, uint64_t nocc, uint32_t pat_base, int move, int append);
/* Row order of every later search on this handle (EDSBWT_ROWS_*).  EDSBWT_ROWS_REFERENCE reproduces the order in which
 * findMultipleDollarsBackward (EDSBWTsearch.cpp:349-603) reaches indices_of_dollars_in_interval (:1310-1332), so that the
 * csv equals the reference's byte for byte without sorting; costs one segmented sort of the tuples on the device. */
int  edsbwt_set_row_order(edsbwt_index* idx, int order);
/* The csv rows of a device-resident result, formatted to decimal text ON THE DEVICE (one thread per row) and appended to /
 * written at path (append = 0 writes the header first).  *bytes = bytes of row text written. */
int  edsbwt_result_write_csv(edsbwt_index* idx, const edsbwt_result* res, const char* path, int move, int append, uint64_t* bytes);
/* Whole job of the reference's constructor loop (:207-241) for HOST patterns: search + locate in reference row order,
 * rows formatted on the device and appended to csv_path (which must already hold the header: edsbwt_write_csv with
 * nocc = 0, append = 0), counts[npat] filled.  The batch is processed in chunks (edsbwt_set_chunk). */
int  edsbwt_search_to_csv(edsbwt_index* idx, const uint8_t* pat_bytes, const uint64_t* pat_off, uint32_t npat,
                          int lf_variant, uint32_t* counts, const char* csv_path, uint64_t* nocc);

/* --- pattern file: std::getline loop of the constructor (:200-241): one pattern per '\n'-separated line,
 *     last line may lack '\n'.  *bytes / *off are malloc'ed (edsbwt_free). */
int  edsbwt_read_patterns(const char* path, uint8_t** bytes, uint64_t** off, uint32_t* npat);

/* --- index construction (replaces eds_to_fasta.cpp + gsufsort + da_to_everything.cpp):
 *     GPU suffix sort of the string collection, writes the reference's files
 *     out_base_info.aux, _bwt_<p>.aux, _bv_<p>.aux (sdsl rrr_vector<63>), .bitvector. */
int  edsbwt_build_from_eds(const char* eds_path, const char* out_base, int device);
/* text = S_0 S_1 ... (letters only, no separators), str_off[nstr+1], seg_first[nstr] = 1 for the first string
 * of each segment.  HOST pointers. */
int  edsbwt_build_from_collection(const uint8_t* text, const uint64_t* str_off, uint32_t nstr,
                                  const uint8_t* seg_first, const char* out_base, int device);

/* --- EDS text hygiene (replaces stringCheck, stringCheck.cpp:11-84): adds the curly brackets around the
 *     non-degenerate stretches of in_path, removes <...> comments, writes out_path (the caller appends ".eds" like the
 *     reference's main, :99-106).  EDSBWT_E_FORMAT when the text holds the reserved empty-word character 'Z'
 *     (Parameters.h:34; the reference exits).  Host-only text processing, as in the reference. */
int  edsbwt_string_check(const char* in_path, const char* out_path);

/* --- roofline denominator: 64-byte random gather over buffer_bytes of HBM (kernel K7). */
int  edsbwt_gather_bench(int device, uint64_t buffer_bytes, uint64_t n_access, int iters, double* gb_per_s);
int  edsbwt_gather_bench_ex(int device, uint64_t buffer_bytes, uint64_t n_access, int iters, int granule_bytes /* 32, 64 */, double* gb_per_s);
/* The full sweep space (csrc/gather_bench.cu): one thread fetches a whole block with the search kernels' own 256-bit loads;
 * ilp = independent blocks in flight per thread (1, 2, 4, 8); ctas_per_sm = resident CTAs of 256 threads (0: as many as fit);
 * dependent = 1: every chain's next address comes out of the loaded data (what a backward search does); mix_bytes / mix_frac:
 * a second buffer and the fraction of accesses that also gather 16 bytes from it (the seed / link / locate gathers).
 * *gb_per_s counts granule_bytes per block + 16 per side gather; *req_per_s (may be NULL) the requests. */
typedef struct {
    uint64_t buffer_bytes, n_access;
    int32_t  iters, granule_bytes, ilp, ctas_per_sm, dependent, reserved;
    uint64_t mix_bytes; double mix_frac;
} edsbwt_gather_cfg;
int  edsbwt_gather_bench_cfg(int device, const edsbwt_gather_cfg* cfg, double* gb_per_s, double* req_per_s);

/* Count-only search of a device-resident batch that ALLOCATES NOTHING (csrc/static_scratch.cu).  No counterpart in the
 * reference: added for callers that cannot call cudaMalloc (CUDA graph capture, tight
 * serving loops, harnesses that forbid allocation).  Same kernel and same answers as edsbwt_search_device(locate = 0), but its
 * scratch lives in one file-scope device array of EDSBWT_STATIC_SCRATCH_BYTES per GPU (held until the process exits) instead
 * of the handle's growable workspace.  d_pat_bytes / d_pat_off (npat + 1 offsets) / d_counts (npat) are DEVICE pointers of the
 * handle's device.  npat > EDSBWT_STATIC_MAX_PATTERNS -> EDSBWT_E_NOSPACE, nothing launched.  Patterns whose frontier needs
 * more than the array holds (the ones the other entry points hand to the level-synchronous engine) -> EDSBWT_E_NOSPACE,
 * *n_unresolved = how many, their counts read 0, all other counts valid.  Returns with the handle's stream idle.  The array is
 * shared by all calls and handles on a device: calls are serialised inside the library; do not overlap them on the GPU with
 * anything else that uses this entry point from another process-wide context.  The first call on a device looks the array's
 * address up (cudaGetSymbolAddress); no call uses cudaMalloc / cudaFree. */
#define EDSBWT_STATIC_SCRATCH_BYTES (8u << 20)
#define EDSBWT_STATIC_MAX_PATTERNS  65536u
int  edsbwt_count_static(edsbwt_index* idx, const uint8_t* d_pat_bytes, const uint64_t* d_pat_off, uint32_t npat,
                         uint32_t* d_counts, uint32_t* n_unresolved);

void        edsbwt_free(void* p);
const char* edsbwt_strerror(int code);
const char* edsbwt_last_error(void);   /* thread-local detail of the last failure */
const char* edsbwt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* EDSBWT_H */
