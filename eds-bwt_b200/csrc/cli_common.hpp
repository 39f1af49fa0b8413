// Shared body of the two drop-in executables (EDSBWTsearch, MOVE_EDSBWTSearch): C++ host code over the C ABI of
// libedsbwt_cuda.so only (include/edsbwt.h).  Mirrors the operator surface of the reference's
//   mainEDS-BWT.cpp:39-83 + EDSBWT::EDSBWT (EDSBWTsearch.cpp:56-294)               -> run_cli(..., move = false)
//   mainMove_EDSBWT.cpp:17-62 + MOVE_EDSBWT::MOVE_EDSBWT (MOVE_EDSBWTSearch.cpp:23-176) -> run_cli(..., move = true)
// argv, stdout/stderr lines scraped by the authors' scripts (launch.sh:93, launch_COVID.sh:106-108), csv naming and
// header, RECOVERBW semantics and the exit status (1 on success!) are the reference's; the search is one batched GPU call.
#pragma once
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "edsbwt.h"

#ifndef RECOVERBW
#define RECOVERBW 1            // makefile:6 of the reference
#endif

namespace edsbwt_cli {

// "Found <N> starting positions of occurrences of pattern number <i>\n" (EDSBWTsearch.cpp:746) without snprintf: ten million of
// these lines are a visible part of a batch's wall time.  Writes at most 80 characters at w, returns the new end.
inline char* put_u32(char* w, uint32_t v) {
    char d[10]; int n = 0;
    do { d[n++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (n) *w++ = d[--n];
    return w;
}
inline char* put_found(char* w, uint32_t count, uint32_t number) {
    memcpy(w, "Found ", 6); w = put_u32(w + 6, count);
    memcpy(w, " starting positions of occurrences of pattern number ", 53); w = put_u32(w + 53, number);
    *w++ = '\n';
    return w;
}


inline int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}
[[noreturn]] inline void die(const char* what, int rc) {
    std::cerr << what << ": " << edsbwt_strerror(rc) << " (" << edsbwt_last_error() << ")" << std::endl;
    exit(1);                   // every error path of the reference is exit(1) / EXIT_FAILURE
}

inline int run_cli(int argc, char** argv, bool move) {
    if (argc != 3) {           // mainEDS-BWT.cpp:42-48 / mainMove_EDSBWT.cpp:19-25
        std::cerr << "usage: " << argv[0] << " inputEBWTfile inputPATTERNfile" << std::endl;
        std::cerr << "where:" << std::endl;
        std::cerr << "  inputEBWTfile is the BWT filename without the extension .ebwt (and .ebwt.qs for the QS string)" << std::endl;
        std::cerr << "  inputPATTERNfile is the pattern file" << std::endl;
        exit(1);
    }
    // RECOVERBW is a compile-time flag in the reference; here also a runtime override
    const int recover = env_int("EDSBWT_RECOVERBW", RECOVERBW) != 0 || move;      // MOVE always locates (MOVE_EDSBWTSearch.cpp:326-369)
    const int device = env_int("EDSBWT_DEVICE", 0);
    const std::string fileInput = argv[1], filePattern = argv[2];

    std::cout << "BCR_eds: " << argv[0] << std::endl;
    std::cout << "BCR_eds: The input ebwt file is " << argv[1] << std::endl;
    std::cout << "BCR_eds: The pattern file is " << argv[2] << std::endl;
    std::cerr << "Backward Search\n";
    std::cout << "DEBUG: " << 0 << std::endl;

    // ---- index load (recoverInfo + computeVectorOcc + .bitvector)
    edsbwt_open_opts opts;
    memset(&opts, 0, sizeof opts);
    opts.seed_k = env_int("EDSBWT_SEED_K", -1);
    opts.locate_mode = env_int("EDSBWT_LOCATE_WALK", 0) ? EDSBWT_LOCATE_WALK : EDSBWT_LOCATE_SA;
    opts.verify = env_int("EDSBWT_VERIFY", 1);
    opts.build_move = move ? 1 : 0;
    opts.cache = env_int("EDSBWT_CACHE", 0);               // 1: keep / reuse the HBM image in <index>.edsbwt
    edsbwt_index* idx = nullptr;
    int rc = edsbwt_index_open_ex(fileInput.c_str(), device, &opts, &idx);
    if (rc) die(("Error opening " + fileInput + "_info.aux / index files").c_str(), rc);
    edsbwt_info info;
    edsbwt_index_info(idx, &info);
    const std::string fnInfoFile = fileInput + "_info.aux";
    std::cout << "\nFrom " << fnInfoFile << " file:\n";                       // EDSBWTsearch.cpp:1577-1584
    std::cout << "\tNumber of sequences: " << info.n_text << "\n";
    std::cout << "\tTotal length (with $): " << info.n << "\n";
    std::cout << "\tSize alpha: " << (int)info.sigma << "\n";
    std::cout << "\tAlphabet: ";
    for (uint32_t i = 0; i < info.sigma; ++i) std::cout << (char)info.alphabet[i] << "\t";
    std::cout << "\n";
    std::cout << "\nFrom " << fnInfoFile << " file (TableOcc):\n";           // :1647-1652
    for (uint32_t j = 0; j < info.sigma; ++j) {
        for (uint32_t h = 0; h < info.sigma; ++h) std::cout << info.table_occ_wide[j * info.sigma + h] << "\t";
        std::cout << std::endl;
    }
    if (!move) {
        for (uint32_t x = 0; x < info.sigma; ++x) {                             // computeVectorOcc, :1341-1347 (freq[] is never filled: prints 0)
            uint64_t pile = 0;
            for (uint32_t h = 0; h < info.sigma; ++h) pile += info.table_occ_wide[x * info.sigma + h];
            std::cerr << "computeVectorUnbuildBCR: freq[" << x << "]= " << 0 << " and numBlocksInPartialBWT[" << x << "]= " << (pile + 1023) / 1024 << "\n";
        }
    } else {
        std::cout << "The bitvector is in: " << fileInput << ".bitvector" << std::endl;
    }

    // ---- output file (created and truncated before the search, like the reference: :179-193 / MOVE :55-60)
    const std::string csv = filePattern + (move ? "output_M_LF.csv" : "output.csv");
    if (recover) {
        rc = edsbwt_write_csv(csv.c_str(), nullptr, 0, 0, move ? 1 : 0, 0);
        if (rc) { std::cerr << "ERROR opening file " << csv << " to write output\n"; exit(1); }
    }

    // ---- patterns (std::getline loop, :207)
    uint8_t* bytes = nullptr; uint64_t* off = nullptr; uint32_t npat = 0;
    rc = edsbwt_read_patterns(filePattern.c_str(), &bytes, &off, &npat);
    if (rc) die(("Error opening " + filePattern).c_str(), rc);

    const auto t0 = std::chrono::steady_clock::now();
    std::vector<uint32_t> counts(npat ? npat : 1, 0), interrupted;
    edsbwt_occ* occ = nullptr; uint64_t nocc = 0;
    const int lf = move ? EDSBWT_LF_MOVE : EDSBWT_LF_RANK;
    // locate: rows come out in the reference's csv order and are formatted to text on the device
    // EDSBWT_DEVICES = all | <n>: the batch is sharded over that many GPUs of the box, results gathered on the first one over
    // NCCL / NVLink (edsbwt_search_multi) and written from there; the handle opened above stays the one that reports the index
    const char* devs = getenv("EDSBWT_DEVICES");
    const int want_dev = (devs && *devs) ? (strcmp(devs, "all") == 0 ? 0 : atoi(devs)) : 1;
    if (recover && want_dev != 1) {
        edsbwt_index_close(idx); idx = nullptr;          // the multi handle opens its own replicas
        std::vector<int> list;
        for (int d = 0; d < want_dev; ++d) list.push_back(d);
        edsbwt_multi* mh = nullptr;
        rc = edsbwt_multi_open(fileInput.c_str(), want_dev ? list.data() : nullptr, want_dev, &opts, &mh);
        if (rc) die("multi-GPU open failed", rc);
        int nd = 0; edsbwt_multi_info(mh, &nd, nullptr, 0, nullptr);
        for (int d = 0; d < nd; ++d) edsbwt_set_row_order(edsbwt_multi_handle(mh, d), EDSBWT_ROWS_REFERENCE);
        edsbwt_multi_result* mr = nullptr;
        rc = edsbwt_search_multi(mh, bytes, off, 0, npat, 1, lf, env_int("EDSBWT_CHUNKS", 4), &mr);
        if (rc) die("search failed", rc);
        rc = edsbwt_multi_result_fetch(mr, counts.data(), nullptr);
        if (!rc) rc = edsbwt_multi_result_write_csv(mh, mr, csv.c_str(), move ? 1 : 0, 1, nullptr);
        edsbwt_multi_result_free(mr);
        edsbwt_multi_close(mh);
    }
    else if (recover) rc = edsbwt_search_to_csv(idx, bytes, off, npat, lf, counts.data(), csv.c_str(), &nocc);
    else { interrupted.assign(npat ? npat : 1, 0); rc = edsbwt_search_trace(idx, bytes, off, npat, lf, counts.data(), interrupted.data()); }
    if (rc) die("search failed", rc);
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();

    // ---- per-pattern report (:746 / MOVE :113,371)
    uint64_t count_found = 0, count_not_found = 0;
    std::string out;                                     // MOVE: the whole report, written after the loop
    if (move) out.reserve((size_t)npat * 72 + 64);
    std::vector<char> chunk(move ? 0 : (1u << 22));      // EDSBWTsearch: "Found ..." lines leave in 4 MB pieces
    char* w = chunk.data();
    std::string err;
    char line[160];
    for (uint32_t i = 0; i < npat; ++i) {
        if (move) {
            out.append("Pattern: "); out.append((const char*)bytes + off[i], (size_t)(off[i + 1] - off[i]));
            // found: "num occ N" (:371); not found: one of the "pattern non presente" early returns (:250-253, 281-297)
            if (counts[i]) snprintf(line, sizeof line, " of length %llu\nnum occ %u\n", (unsigned long long)(off[i + 1] - off[i]), counts[i]);
            else snprintf(line, sizeof line, " of length %llu\npattern non presente\n", (unsigned long long)(off[i + 1] - off[i]));
            std::cerr << "OCCORRENZA DI: ";
            std::cerr.write((const char*)bytes + off[i], (std::streamsize)(off[i + 1] - off[i]));
            std::cerr << (counts[i] ? " TROVATA" : " NON TROVATA") << std::endl;                // :124,129
            out.append(line);
        } else {
            if (!recover && interrupted[i]) {                                                   // :714-717; stderr is unbuffered: one write per 4 MB, not per line
                char num[12]; char* e = put_u32(num, interrupted[i]); *e++ = '\n';
                err.append("Pattern not found, interrupted in position ", 43); err.append(num, (size_t)(e - num));
                if (err.size() > (1u << 22)) { fwrite(err.data(), 1, err.size(), stderr); err.clear(); }
            }
            w = put_found(w, counts[i], i + 1);
            if ((size_t)(w - chunk.data()) + 96 > chunk.size()) { fwrite(chunk.data(), 1, (size_t)(w - chunk.data()), stdout); w = chunk.data(); }
        }
        if (counts[i] > 0) ++count_found; else ++count_not_found;
    }
    if (!err.empty()) fwrite(err.data(), 1, err.size(), stderr);
    if (move) fwrite(out.data(), 1, out.size(), stdout);
    else if (w != chunk.data()) fwrite(chunk.data(), 1, (size_t)(w - chunk.data()), stdout);
    fflush(stdout);
    std::cout << "bs took:" << (float)secs;                                   // :244 (no newline)
    std::cout.flush();
    std::cerr << std::endl;
    std::cerr << "count_found = " << count_found << std::endl;
    std::cerr << "count_not_found = " << count_not_found << std::endl;

    edsbwt_free(occ); edsbwt_free(bytes); edsbwt_free(off);
    if (recover) std::cerr << "\nThe csv file is ready! \n";                 // mainEDS-BWT.cpp:72-76
    else std::cerr << "\nThe search is finished! \n";
    if (idx) edsbwt_index_close(idx);
    std::cerr << "The End!\n";
    return 1;                  // mainEDS-BWT.cpp:82, mainMove_EDSBWT.cpp:61: success IS exit status 1
}

}  // namespace edsbwt_cli
