/* TEST INFRASTRUCTURE (oracle) -- never linked into, imported by or executed from the product path.
 *
 * Plain-C CPU restatement of the reference's pattern-search hot path:
 * pile-relative 1-based intervals, rank = sampled per-1024-block counts + an in-block byte scan,
 * link() through the end-marker array, locate by rounds of LF over all symbols.  Each function names the
 * reference lines it follows.  Differences from the reference that do not change results: piles are kept in
 * RAM instead of being fseek/fread per rank; the '#' rank comes from the pile bytes (the reference reads the
 * same bitmap from <base>_bv_<p>.aux, whose content is cross-checked by oracle/eds_index.py:read_index);
 * patterns with a character outside the alphabet (or the end-marker) yield 0 instead of the reference's
 * uninitialised-table behaviour; a single-segment EDS yields plain matches instead of the
 * reference's sdsl assertion abort (App. C.3).
 *
 * usage: edsbwt_oracle <indexBase> <patternsFile> [--csv FILE] [--counts FILE] [--locate 0|1]
 *                      [--threads T] [--k K] [--limit N]
 *   --csv     write "#Pat\t$_i\tD[i]\tS_j\tS_j[r]" rows in the reference's emission order
 *   --counts  write one count per pattern line
 *   --k       canonical seed depth for the I,S,E,L,O byte-model counters, default 12
 * stdout: one JSON line {"patterns":..,"found":..,"seconds":..,"I":..,"S":..,"E":..,"L":..,"O":..}
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define DIMBLOCK 1024u          /* EDSBWTsearch.hpp:44 */
#define MAXSIGMA 64

typedef struct { uint32_t s, e; } Iv;                 /* rangeElement, Sorting.h:38-47 (1-based, inclusive) */
typedef struct { uint8_t pile; uint32_t s, e; } IvBW; /* rangeElementBW, Sorting.h:49-57 */
typedef struct { Iv *a; size_t n, cap; } IvVec;
typedef struct { IvBW *a; size_t n, cap; } IvBWVec;
typedef struct { uint32_t *a; size_t n, cap; } U32Vec;
typedef struct { uint32_t pat, w, seg, j, off; } Row;
typedef struct { Row *a; size_t n, cap; } RowVec;

static void *xrealloc(void *p, size_t sz) { void *q = realloc(p, sz ? sz : 1); if (!q) { fprintf(stderr, "oracle: out of memory\n"); exit(2); } return q; }
#define PUSH(v, x) do { if ((v)->n == (v)->cap) { (v)->cap = (v)->cap ? (v)->cap * 2 : 16; (v)->a = xrealloc((v)->a, (v)->cap * sizeof(*(v)->a)); } (v)->a[(v)->n++] = (x); } while (0)

typedef struct {
    uint32_t n, nText; int sigma;
    uint8_t alphaInv[MAXSIGMA]; int alpha[256];
    uint8_t *pile[MAXSIGMA]; uint32_t pileLen[MAXSIGMA];
    uint32_t *eofId[MAXSIGMA]; uint32_t nEof[MAXSIGMA];
    uint32_t tableOcc[MAXSIGMA][MAXSIGMA];
    uint32_t *vocc[MAXSIGMA];      /* vocc[p][blk*sigma+c] = #c in pile p up to the END of block blk (computeVectorOcc) */
    uint32_t Cq[MAXSIGMA][MAXSIGMA]; /* Cq[q][c] = sum_{g<q} tableOcc[g][c] */
    uint32_t *strSeg;              /* rb_1(w+1): 1-based segment of string w */
    uint32_t *segStart;            /* bsel_1(k): first string of segment k (1-based k), segStart[nSeg+1] = nText */
    uint32_t nSeg, firstSymbolIndex;
} Index;

static uint8_t *slurp(const char *path, size_t *len) {
    FILE *f = fopen(path, "rb");
    if (!f) { fprintf(stderr, "Error opening %s.\n", path); exit(1); }
    fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
    uint8_t *b = xrealloc(NULL, (size_t)sz + 1);
    if (sz && fread(b, 1, (size_t)sz, f) != (size_t)sz) { fprintf(stderr, "short read %s\n", path); exit(1); }
    fclose(f); *len = (size_t)sz; return b;
}

/* recoverInfo (EDSBWTsearch.cpp:1557-1656) + computeVectorOcc (:1338-1442) + ctor bitvector part (:150-168) */
static void index_load(Index *ix, const char *base) {
    char path[4096]; size_t len;
    memset(ix, 0, sizeof *ix);
    snprintf(path, sizeof path, "%s_info.aux", base);
    uint8_t *info = slurp(path, &len);
    memcpy(&ix->n, info, 4); memcpy(&ix->nText, info + 4, 4); ix->sigma = info[8];
    if (ix->sigma > MAXSIGMA) { fprintf(stderr, "oracle: alphabet too large\n"); exit(1); }
    for (int i = 0; i < 256; i++) ix->alpha[i] = -1;
    for (int i = 0; i < ix->sigma; i++) { ix->alphaInv[i] = info[9 + i]; ix->alpha[info[9 + i]] = i; }
    size_t off = 9 + (size_t)ix->sigma;
    for (int p = 0; p < ix->sigma; p++) {
        snprintf(path, sizeof path, "%s_bwt_%d.aux", base, p);
        size_t pl; ix->pile[p] = slurp(path, &pl); ix->pileLen[p] = (uint32_t)pl;
        uint32_t nblk = (uint32_t)((pl + DIMBLOCK - 1) / DIMBLOCK);
        ix->vocc[p] = calloc((size_t)(nblk ? nblk : 1) * ix->sigma, 4);
        uint32_t run[MAXSIGMA] = {0};
        for (uint32_t i = 0; i < pl; i++) {
            int c = ix->alpha[ix->pile[p][i]];
            if (c < 0) { fprintf(stderr, "oracle: pile %d holds a byte outside the alphabet\n", p); exit(1); }
            run[c]++;
            if ((i + 1) % DIMBLOCK == 0 || i + 1 == pl)
                for (int z = 0; z < ix->sigma; z++) ix->vocc[p][(size_t)(i / DIMBLOCK) * ix->sigma + z] = run[z];
        }
        ix->nEof[p] = run[0];
    }
    for (int p = 0; p < ix->sigma; p++) {           /* EOF_ID, :1612-1626 */
        ix->eofId[p] = xrealloc(NULL, (size_t)ix->nEof[p] * 4);
        memcpy(ix->eofId[p], info + off, (size_t)ix->nEof[p] * 4); off += (size_t)ix->nEof[p] * 4;
    }
    for (int p = 0; p < ix->sigma; p++) for (int c = 0; c < ix->sigma; c++) { memcpy(&ix->tableOcc[p][c], info + off, 4); off += 4; }
    if (off != len) { fprintf(stderr, "oracle: %s_info.aux has %zu bytes, expected %zu\n", base, len, off); exit(1); }
    for (int c = 0; c < ix->sigma; c++) { uint32_t acc = 0; for (int q = 0; q < ix->sigma; q++) { ix->Cq[q][c] = acc; acc += ix->tableOcc[q][c]; } }
    free(info);
    /* <base>.bitvector: sdsl bit_vector = u64 size + words (App. A.3) */
    snprintf(path, sizeof path, "%s.bitvector", base);
    uint8_t *bv = slurp(path, &len);
    uint64_t nbits; memcpy(&nbits, bv, 8);
    if (nbits != ix->nText) { fprintf(stderr, "oracle: .bitvector has %llu bits, nText=%u\n", (unsigned long long)nbits, ix->nText); exit(1); }
    ix->strSeg = xrealloc(NULL, (size_t)ix->nText * 4);
    ix->segStart = xrealloc(NULL, ((size_t)ix->nText + 2) * 4);
    uint32_t seg = 0;
    for (uint32_t w = 0; w < ix->nText; w++) {
        uint64_t word; memcpy(&word, bv + 8 + 8 * (w >> 6), 8);
        if ((word >> (w & 63)) & 1) { seg++; ix->segStart[seg] = w; }
        ix->strSeg[w] = seg;
    }
    ix->nSeg = seg; ix->segStart[seg + 1] = ix->nText;
    /* first_symbol_index = bsel_1(2)-1 (:168); the reference aborts on a single segment (App. C.3) */
    ix->firstSymbolIndex = (seg >= 2) ? ix->segStart[2] - 1 : ix->nText - 1;
    free(bv);
}

/* counts of every symbol in pile p[0, pos): vectorOcc of the preceding blocks (findBlockToRead :1444-1458)
 * plus the byte loop of rankManySymbols (:1160-1209) */
static void rank_all(const Index *ix, int p, uint32_t pos, uint32_t *cnt) {
    uint32_t blk = pos ? (pos - 1) / DIMBLOCK : 0;
    if (blk > 0) memcpy(cnt, ix->vocc[p] + (size_t)(blk - 1) * ix->sigma, (size_t)ix->sigma * 4);
    else memset(cnt, 0, (size_t)ix->sigma * 4);
    const uint8_t *b = ix->pile[p];
    for (uint32_t r = blk * DIMBLOCK; r < pos; r++) cnt[ix->alpha[b[r]]]++;
}
static uint32_t rank_one(const Index *ix, int p, int c, uint32_t pos) {
    uint32_t blk = pos ? (pos - 1) / DIMBLOCK : 0;
    uint32_t r0 = blk > 0 ? ix->vocc[p][(size_t)(blk - 1) * ix->sigma + c] : 0;
    const uint8_t *b = ix->pile[p]; uint8_t ch = ix->alphaInv[c];
    for (uint32_t r = blk * DIMBLOCK; r < pos; r++) r0 += (b[r] == ch);
    return r0;
}
/* EOFpos_rank[p](i): end-markers in pile p[0,i) */
static uint32_t eof_rank(const Index *ix, int p, uint32_t i) { return rank_one(ix, p, 0, i); }

typedef struct { uint64_t I, S, E, L, O; } Counters;

static int cmp_u32(const void *a, const void *b) { uint32_t x = *(const uint32_t *)a, y = *(const uint32_t *)b; return x < y ? -1 : x > y; }

/* dollars_in_interval, EDSBWTsearch.cpp:1265-1281 */
static void dollars_in_interval(const Index *ix, int p, U32Vec *d, uint32_t i, uint32_t j) {
    uint32_t l = eof_rank(ix, p, i - 1), u = eof_rank(ix, p, j);
    for (uint32_t k = l; k < u; k++) { uint32_t w = ix->eofId[p][k]; if (w > ix->firstSymbolIndex) PUSH(d, w); }
}

/* link + preceding_dollars_finder, EDSBWTsearch.cpp:1212-1262, 1287-1307 */
static void link_step(const Index *ix, int p, const IvVec *F, IvVec *Dol, U32Vec *d) {
    d->n = 0; Dol->n = 0;
    for (size_t t = 0; t < F->n; t++) if (F->a[t].s <= F->a[t].e) dollars_in_interval(ix, p, d, F->a[t].s, F->a[t].e);
    qsort(d->a, d->n, 4, cmp_u32);
    while (d->n) {
        uint32_t w = d->a[--d->n];
        uint32_t a = ix->strSeg[w];                       /* rb_1(i+1) */
        Iv cur = { ix->segStart[a - 1] + 1, ix->segStart[a] - 1 + 1 };
        if (Dol->n && cur.e + 1 >= Dol->a[Dol->n - 1].s) { if (cur.s < Dol->a[Dol->n - 1].s) Dol->a[Dol->n - 1].s = cur.s; }
        else PUSH(Dol, cur);
        dollars_in_interval(ix, 0, d, cur.s, cur.e);
    }
    for (size_t i = 0, j = Dol->n; i + 1 < j; i++, j--) { Iv t = Dol->a[i]; Dol->a[i] = Dol->a[j - 1]; Dol->a[j - 1] = t; }
}

/* updateIntervals + MergeAndRemove, EDSBWTsearch.cpp:956-1157, 911-953:
 * start' = C_q[c] + rank_c(q,start-1) + 1 ; end' = C_q[c] + rank_c(q,end); drop empty; merge start'==prev.end'+1 */
static void update_intervals(const Index *ix, IvVec *v, int c, int q) {
    size_t out = 0;
    for (size_t t = 0; t < v->n; t++) {
        Iv x = v->a[t];
        if (x.s > x.e) continue;
        Iv y = { ix->Cq[q][c] + rank_one(ix, q, c, x.s - 1) + 1, ix->Cq[q][c] + rank_one(ix, q, c, x.e) };
        if (y.s > y.e) continue;
        if (out && y.s == v->a[out - 1].e + 1) v->a[out - 1].e = y.e; else v->a[out++] = y;
    }
    v->n = out;
}

static int cmp_bw(const void *a, const void *b) {   /* cmpSortEl, Sorting.cpp:8-13 */
    const IvBW *x = a, *y = b;
    if (x->pile != y->pile) return x->pile < y->pile ? -1 : 1;
    return x->s < y->s ? -1 : x->s > y->s;
}

/* findMultipleDollarsBackward + indices_of_dollars_in_interval, EDSBWTsearch.cpp:300-616, 1310-1332 */
static void locate(const Index *ix, IvBWVec *R, IvBWVec *R2, uint32_t patNo, RowVec *rows, Counters *ct) {
    uint32_t occPos = 0, cs[MAXSIGMA], ce[MAXSIGMA];
    while (R->n) {
        qsort(R->a, R->n, sizeof(IvBW), cmp_bw);
        R2->n = 0;
        for (size_t t = 0; t < R->n; t++) {
            IvBW x = R->a[t];
            if (x.s > x.e) continue;
            rank_all(ix, x.pile, x.s - 1, cs);
            rank_all(ix, x.pile, x.e, ce);
            if (ce[0] > cs[0]) {
                for (uint32_t k = cs[0]; k < ce[0]; k++) {
                    uint32_t w = ix->eofId[x.pile][k], seg = ix->strSeg[w];
                    Row r = { patNo, w, seg, w - ix->segStart[seg], occPos };
                    PUSH(rows, r); ct->O++; ct->L += (uint64_t)occPos + 1;
                }
            }
            for (int i = 1; i < ix->sigma; i++) if (ce[i] > cs[i]) {
                IvBW y = { (uint8_t)i, ix->Cq[x.pile][i] + cs[i] + 1, ix->Cq[x.pile][i] + ce[i] };
                PUSH(R2, y);
            }
        }
        occPos++;
        IvBWVec t = *R; *R = *R2; *R2 = t;
    }
}

typedef struct { IvVec F, Dol; U32Vec d; IvBWVec R, R2; } Scratch;

/* backwardSearch, EDSBWTsearch.cpp:619-772 */
static uint32_t backward_search(const Index *ix, const uint8_t *P, uint32_t m, uint32_t patNo, int do_locate,
                                uint32_t kseed, Scratch *sc, RowVec *rows, Counters *ct, uint32_t *interrupted) {
    *interrupted = 0;
    if (m == 0) return 0;
    for (uint32_t i = 0; i < m; i++) if (ix->alpha[P[i]] <= 0) return 0;      /* App. C.2 deviation */
    int p = ix->alpha[P[m - 1]];
    sc->F.n = 0; Iv all = { 1, ix->pileLen[p] }; PUSH(&sc->F, all);
    uint32_t suffix_len = 1;
    if (suffix_len == (m < kseed ? m : kseed)) ct->I += sc->F.n;
    for (uint32_t pos = m - 1; pos > 0 && sc->F.n; pos--) {
        int c = ix->alpha[P[pos - 1]];
        int counted = suffix_len >= kseed;          /* this step lies beyond the canonical seed */
        link_step(ix, p, &sc->F, &sc->Dol, &sc->d);
        if (counted) {
            ct->S += sc->F.n + sc->Dol.n;
            for (size_t t = 0; t < sc->F.n; t++) ct->E += eof_rank(ix, p, sc->F.a[t].e) - eof_rank(ix, p, sc->F.a[t].s - 1);
            for (size_t t = 0; t < sc->Dol.n; t++) ct->E += eof_rank(ix, 0, sc->Dol.a[t].e) - eof_rank(ix, 0, sc->Dol.a[t].s - 1);
        }
        if (sc->Dol.n) update_intervals(ix, &sc->Dol, c, 0);
        update_intervals(ix, &sc->F, c, p);
        for (size_t t = 0; t < sc->F.n; t++) PUSH(&sc->Dol, sc->F.a[t]);          /* Dol' ++ F' (:696) */
        IvVec t = sc->F; sc->F = sc->Dol; sc->Dol = t;
        p = c; suffix_len++;
        if (suffix_len == (m < kseed ? m : kseed)) ct->I += sc->F.n;
        if (!sc->F.n) *interrupted = pos;     /* "Pattern not found, interrupted in position", :714-717 (RECOVERBW=0) */
    }
    uint32_t res = 0;
    for (size_t t = 0; t < sc->F.n; t++) res += sc->F.a[t].e - sc->F.a[t].s + 1;
    if (do_locate && sc->F.n) {
        sc->R.n = 0;
        for (size_t t = sc->F.n; t-- > 0;) { IvBW y = { (uint8_t)p, sc->F.a[t].s, sc->F.a[t].e }; PUSH(&sc->R, y); }   /* :757-765 */
        locate(ix, &sc->R, &sc->R2, patNo, rows, ct);
    }
    return res;
}

static double now_s(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }

int main(int argc, char **argv) {
    if (argc < 3) { fprintf(stderr, "usage: %s indexBase patternsFile [--csv F] [--counts F] [--interrupted F] [--locate 0|1] [--threads T] [--k K] [--limit N]\n", argv[0]); return 1; }
    const char *csv = NULL, *cntf = NULL, *intf = NULL; int do_locate = 1, threads = 1; uint32_t kseed = 12; long limit = -1;
    for (int i = 3; i + 1 < argc; i += 2) {
        if (!strcmp(argv[i], "--csv")) csv = argv[i + 1];
        else if (!strcmp(argv[i], "--counts")) cntf = argv[i + 1];
        else if (!strcmp(argv[i], "--interrupted")) intf = argv[i + 1];
        else if (!strcmp(argv[i], "--locate")) do_locate = atoi(argv[i + 1]);
        else if (!strcmp(argv[i], "--threads")) threads = atoi(argv[i + 1]);
        else if (!strcmp(argv[i], "--k")) kseed = (uint32_t)atoi(argv[i + 1]);
        else if (!strcmp(argv[i], "--limit")) limit = atol(argv[i + 1]);
        else { fprintf(stderr, "unknown option %s\n", argv[i]); return 1; }
    }
    Index ix; index_load(&ix, argv[1]);
    size_t plen; uint8_t *pbuf = slurp(argv[2], &plen);
    /* std::getline semantics (:207): split on '\n', last line may lack it, a trailing '\n' adds no line */
    size_t npat = 0, cap = 1024; size_t *pstart = xrealloc(NULL, cap * sizeof(size_t)); uint32_t *pl = xrealloc(NULL, cap * 4);
    for (size_t i = 0; i < plen;) {
        size_t j = i; while (j < plen && pbuf[j] != '\n') j++;
        if (npat == cap) { cap *= 2; pstart = xrealloc(pstart, cap * sizeof(size_t)); pl = xrealloc(pl, cap * 4); }
        pstart[npat] = i; pl[npat] = (uint32_t)(j - i); npat++;
        i = j + 1;
    }
    if (limit >= 0 && (size_t)limit < npat) npat = (size_t)limit;
    uint32_t *counts = calloc(npat ? npat : 1, 4);
    uint32_t *interr = calloc(npat ? npat : 1, 4);
    RowVec *prow = calloc(npat ? npat : 1, sizeof(RowVec));
    Counters tot = {0, 0, 0, 0, 0};
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
#else
    (void)threads;
#endif
    double t0 = now_s();
#pragma omp parallel
    {
        Scratch sc; memset(&sc, 0, sizeof sc); Counters ct = {0, 0, 0, 0, 0};
#pragma omp for schedule(dynamic, 16)
        for (long i = 0; i < (long)npat; i++)
            counts[i] = backward_search(&ix, pbuf + pstart[i], pl[i], (uint32_t)i + 1, do_locate, kseed, &sc, &prow[i], &ct, &interr[i]);
#pragma omp critical
        { tot.I += ct.I; tot.S += ct.S; tot.E += ct.E; tot.L += ct.L; tot.O += ct.O; }
        free(sc.F.a); free(sc.Dol.a); free(sc.d.a); free(sc.R.a); free(sc.R2.a);
    }
    double t1 = now_s();
    size_t found = 0; for (size_t i = 0; i < npat; i++) found += counts[i] > 0;
    if (csv) {
        FILE *f = fopen(csv, "w"); if (!f) { fprintf(stderr, "ERROR opening file %s to write output\n", csv); return 1; }
        fprintf(f, "#Pat\t$_i\tD[i]\tS_j\tS_j[r]\n");                       /* :186 */
        for (size_t i = 0; i < npat; i++) for (size_t r = 0; r < prow[i].n; r++) { Row *q = &prow[i].a[r]; fprintf(f, "%u\t%u\t%u\t%u\t%u\n", q->pat, q->w, q->seg, q->j, q->off); }
        fclose(f);
    }
    if (intf) { FILE *f = fopen(intf, "w"); if (!f) return 1; for (size_t i = 0; i < npat; i++) fprintf(f, "%u\n", interr[i]); fclose(f); }
    if (cntf) { FILE *f = fopen(cntf, "w"); if (!f) return 1; for (size_t i = 0; i < npat; i++) fprintf(f, "%u\n", counts[i]); fclose(f); }
    printf("{\"patterns\": %zu, \"found\": %zu, \"seconds\": %.6f, \"threads\": %d, \"k\": %u, \"I\": %llu, \"S\": %llu, \"E\": %llu, \"L\": %llu, \"O\": %llu}\n",
           npat, found, t1 - t0, threads, kseed, (unsigned long long)tot.I, (unsigned long long)tot.S, (unsigned long long)tot.E,
           (unsigned long long)tot.L, (unsigned long long)tot.O);
    return 0;
}
