"""TEST INFRASTRUCTURE (oracle) -- never imported by the product path.

CPU restatement (numpy) of the reference's *index construction* so that tests can
create index files in the reference's on-disk formats without gsufsort/BCR/sdsl:

  * parse_eds            restates eds_to_fasta.cpp:55-167  (string numbering, the
                         ``.bitvector`` first-of-segment flags, empty strings)
  * build_index          restates the net effect of gsufsort + da_to_everything.cpp:113-290,
                         371-486: suffixes of the string collection sorted with end-markers
                         ordered by string index; the BWT char of a whole-string row is the
                         end-marker; an empty string contributes exactly one row (pile 0, '#')
  * write_index          file layouts of SURVEY.md App. A.1-A.4
  * read_index           the loader side (EDSBWTsearch.cpp:1557-1656, 150-168)
  * brute_force          independent matcher of SURVEY.md App. B.1 (the mathematical
                         definition the reference binary was validated against)

gsufsort / BCR_LCP_GSA are third-party and absent from /root/reference; only their
output convention matters and it is pinned by running the
prebuilt reference binary on files written here (tests/test_oracle_vs_reference.py).
"""
import os
import struct
from dataclasses import dataclass, field
from functools import lru_cache

import numpy as np

from . import sdsl_io

TERMINATE_CHAR = ord("#")     # Parameters.h:30 (second #define wins)
EMPTY_CHAR_EDS = ord("E")     # Parameters.h:33


# ----------------------------------------------------------------------------
# EDS text -> segments
# ----------------------------------------------------------------------------
def parse_eds(text) -> list:
    """EDS text -> list of segments, each a list of ``bytes`` strings.

    Follows eds_to_fasta.cpp:55-140: '{' opens a segment and its first string, ','
    opens the next string, '}' is skipped, EMPTY_CHAR_EDS stands for the empty word.
    Whitespace (a trailing newline) is ignored here; the reference would copy it into
    the FASTA where gsufsort drops it.
    """
    if isinstance(text, str):
        text = text.encode()
    segs = []
    cur = None
    if not text or text[0:1] != b"{":
        raise ValueError("the string does not start with {")   # eds_to_fasta.cpp:81
    for ch in text:
        if ch == ord("{"):
            cur = [bytearray()]
            segs.append(cur)
        elif ch == ord("}"):
            pass
        elif ch == ord(","):
            cur.append(bytearray())
        elif ch == EMPTY_CHAR_EDS or ch in b" \t\r\n":
            pass
        else:
            cur[-1].append(ch)
    return [[bytes(s) for s in seg] for seg in segs]


def format_eds(segs) -> bytes:
    return b"".join(b"{" + b",".join(seg) + b"}" for seg in segs)


# ----------------------------------------------------------------------------
# index
# ----------------------------------------------------------------------------
@dataclass
class Index:
    n: int                      # eBWT length incl. end-markers
    n_text: int                 # number of strings
    alphabet: bytes             # ascending, alphabet[0] == '#'
    piles: list                 # list of bytes objects, one per alphabet symbol
    eof_id: list                # list of np.uint32 arrays, one per pile
    table_occ: np.ndarray       # [sigma, sigma] uint32
    seg_first: np.ndarray       # uint8 0/1 per string: first string of its segment
    # extras (not on disk), kept for tests
    sa_str: np.ndarray = field(default=None, repr=False)   # row -> string index
    sa_off: np.ndarray = field(default=None, repr=False)   # row -> offset inside string
    segs: list = field(default=None, repr=False)

    @property
    def ebwt(self) -> bytes:
        return b"".join(self.piles)


def flatten_collection(segs):
    """segments -> (T, str_start, str_len, seg_first): T = S_0 $ S_1 $ ... with $ coded 0."""
    strings = [s for seg in segs for s in seg]
    seg_first = np.zeros(len(strings), dtype=np.uint8)
    k = 0
    for seg in segs:
        seg_first[k] = 1
        k += len(seg)
    lens = np.array([len(s) for s in strings], dtype=np.int64)
    starts = np.zeros(len(strings) + 1, dtype=np.int64)
    np.cumsum(lens + 1, out=starts[1:])
    T = np.zeros(int(starts[-1]), dtype=np.uint8)
    joined = np.frombuffer(b"\x00".join(strings) + b"\x00", dtype=np.uint8)
    T[:] = joined
    return T, starts[:-1].copy(), lens, seg_first


def suffix_order(T: np.ndarray, str_start: np.ndarray, str_len: np.ndarray) -> np.ndarray:
    """Suffix array of the collection by prefix doubling.

    Every end-marker is unique and ordered by string index (gsufsort convention,
    SURVEY 8(c)), so a suffix is fully determined once its prefix reaches an end-marker.
    """
    n = len(T)
    n_text = len(str_start)
    end_pos = str_start + str_len
    rank = T.astype(np.int64) + n_text          # letters above all end-markers
    rank[end_pos] = np.arange(n_text)
    # compress initial ranks
    order = np.argsort(rank, kind="stable")
    sr = rank[order]
    newr = np.zeros(n, dtype=np.int64)
    newr[1:] = np.cumsum(sr[1:] != sr[:-1])
    rank = np.empty(n, dtype=np.int64)
    rank[order] = newr
    h = 1
    while int(newr[-1]) != n - 1:
        r2 = np.zeros(n, dtype=np.int64)
        r2[: n - h] = rank[h:]
        key = rank * n + r2
        order = np.argsort(key, kind="stable")
        sk = key[order]
        newr = np.zeros(n, dtype=np.int64)
        newr[1:] = np.cumsum(sk[1:] != sk[:-1])
        rank = np.empty(n, dtype=np.int64)
        rank[order] = newr
        h *= 2
    sa = np.empty(n, dtype=np.int64)
    sa[rank] = np.arange(n)
    return sa


def build_index(segs) -> Index:
    T, str_start, str_len, seg_first = flatten_collection(segs)
    n = len(T)
    n_text = len(str_start)
    sa = suffix_order(T, str_start, str_len)
    # position -> string index / offset
    pos_str = np.repeat(np.arange(n_text), str_len + 1)
    pos_off = np.arange(n) - str_start[pos_str]
    # BWT: preceding char inside the string, or the end-marker for a whole-string row
    bw = np.where(pos_off[sa] == 0, TERMINATE_CHAR, T[np.maximum(sa - 1, 0)]).astype(np.uint8)
    first = np.where(T[sa] == 0, TERMINATE_CHAR, T[sa]).astype(np.uint8)
    letters = sorted(set(T[T != 0].tolist()))
    alphabet = bytes([TERMINATE_CHAR] + letters)
    assert all(c > TERMINATE_CHAR for c in letters), "letters must sort above the end-marker"
    sigma = len(alphabet)
    code = np.zeros(256, dtype=np.int64)
    for i, c in enumerate(alphabet):
        code[c] = i
    piles, eof_id = [], []
    table = np.zeros((sigma, sigma), dtype=np.uint32)
    da = pos_str[sa]
    fc = code[first]
    for p in range(sigma):
        sel = fc == p
        pb = bw[sel]
        piles.append(pb.tobytes())
        eof_id.append(da[sel][pb == TERMINATE_CHAR].astype(np.uint32))
        table[p] = np.bincount(code[pb], minlength=sigma).astype(np.uint32)
    return Index(n=n, n_text=n_text, alphabet=alphabet, piles=piles, eof_id=eof_id, table_occ=table,
                 seg_first=seg_first, sa_str=da.astype(np.int64), sa_off=pos_off[sa].astype(np.int64), segs=segs)


def info_bytes(ix: Index) -> bytes:
    """<base>_info.aux, App. A.1 (da_to_everything.cpp:149-163,221,249-254)."""
    out = struct.pack("<IIB", ix.n, ix.n_text, len(ix.alphabet)) + ix.alphabet
    for e in ix.eof_id:
        out += np.asarray(e, dtype="<u4").tobytes()
    out += np.asarray(ix.table_occ, dtype="<u4").tobytes()
    return out


def write_index(ix: Index, base: str, rrr: bool = True) -> None:
    """Write <base>_info.aux, <base>_bwt_<p>.aux, <base>_bv_<p>.aux, <base>.bitvector.

    rrr=False writes the ``_bv_`` files as plain sdsl bit_vectors, which only the
    shim-built oracle/_ref binaries understand (SURVEY 8(c) build 2).
    """
    with open(base + "_info.aux", "wb") as f:
        f.write(info_bytes(ix))
    for p, pile in enumerate(ix.piles):
        with open(f"{base}_bwt_{p}.aux", "wb") as f:
            f.write(pile)
        bm = (np.frombuffer(pile, dtype=np.uint8) == TERMINATE_CHAR).astype(np.uint8)
        with open(f"{base}_bv_{p}.aux", "wb") as f:
            f.write(sdsl_io.write_rrr_vector(bm) if rrr else sdsl_io.write_bit_vector(bm))
    with open(base + ".bitvector", "wb") as f:
        f.write(sdsl_io.write_bit_vector(ix.seg_first))


def read_index(base: str) -> Index:
    """Loader side: EDSBWTsearch.cpp:1557-1656 (info, bv, EOF_ID, tableOcc) + :150-168."""
    with open(base + "_info.aux", "rb") as f:
        buf = f.read()
    n, n_text, sigma = struct.unpack_from("<IIB", buf, 0)
    alphabet = buf[9: 9 + sigma]
    off = 9 + sigma
    piles, bvs = [], []
    for p in range(sigma):
        with open(f"{base}_bwt_{p}.aux", "rb") as f:
            piles.append(f.read())
        with open(f"{base}_bv_{p}.aux", "rb") as f:
            bvs.append(sdsl_io.read_rrr_vector(f.read()))
    eof_id = []
    for p in range(sigma):
        cnt = int(bvs[p].sum())
        eof_id.append(np.frombuffer(buf, dtype="<u4", count=cnt, offset=off).astype(np.uint32))
        off += 4 * cnt
    table = np.frombuffer(buf, dtype="<u4", count=sigma * sigma, offset=off).reshape(sigma, sigma).astype(np.uint32)
    with open(base + ".bitvector", "rb") as f:
        seg_first, _ = sdsl_io.read_bit_vector(f.read())
    for p in range(sigma):
        bm = (np.frombuffer(piles[p], dtype=np.uint8) == TERMINATE_CHAR).astype(np.uint8)
        assert np.array_equal(bm, bvs[p]), f"_bv_{p} disagrees with _bwt_{p}"
    return Index(n=n, n_text=n_text, alphabet=alphabet, piles=piles, eof_id=eof_id, table_occ=table,
                 seg_first=seg_first)


# ----------------------------------------------------------------------------
# brute force (definition B.1)
# ----------------------------------------------------------------------------
def brute_force(segs, pattern: bytes):
    """All (w, seg, j, r) at which ``pattern`` occurs, per SURVEY App. B.1, sorted."""
    nseg = len(segs)
    P = bytes(pattern)
    m = len(P)

    @lru_cache(maxsize=None)
    def can(s: int, q: int) -> bool:
        """Can P[q:] (non-empty) be matched starting with a string of segment s (0-based)?"""
        if s >= nseg:
            return False
        rest = m - q
        for X in segs[s]:
            if len(X) >= rest:
                if X[:rest] == P[q:]:
                    return True
            elif P[q: q + len(X)] == X and can(s + 1, q + len(X)):
                return True
        return False

    out = []
    w = 0
    for s, seg in enumerate(segs):
        for j, X in enumerate(seg):
            for r in range(len(X)):
                tail = len(X) - r
                if tail >= m:
                    ok = X[r: r + m] == P
                else:
                    ok = X[r:] == P[:tail] and can(s + 1, tail)
                if ok:
                    out.append((w, s + 1, j, r))
            w += 1
    return sorted(out)


# ----------------------------------------------------------------------------
# small helpers shared by tests / bench
# ----------------------------------------------------------------------------
def write_patterns(path: str, patterns) -> None:
    with open(path, "wb") as f:
        for p in patterns:
            f.write(bytes(p) + b"\n")


def read_csv_rows(path: str):
    """<patterns>output.csv -> sorted list of 5-int tuples (header skipped)."""
    rows = []
    with open(path, "r") as f:
        header = f.readline()
        for line in f:
            line = line.strip()
            if line:
                rows.append(tuple(int(x) for x in line.split("\t")))
    return header, rows


def random_eds(rng, n_seg=30, max_var=4, max_len=6, p_empty=0.15, alphabet=b"ACGT", nondeg_len=(1, 12)):
    """Random EDS alternating non-degenerate stretches and degenerate segments."""
    segs = []
    for s in range(n_seg):
        if s % 2 == 0:
            ln = int(rng.integers(nondeg_len[0], nondeg_len[1] + 1))
            segs.append([bytes(rng.choice(list(alphabet), size=ln).astype(np.uint8))])
        else:
            v = int(rng.integers(2, max_var + 1))
            seen, seg = set(), []
            has_empty = False
            for _ in range(v):
                if not has_empty and rng.random() < p_empty:
                    s_ = b""
                    has_empty = True
                else:
                    ln = int(rng.integers(1, max_len + 1))
                    s_ = bytes(rng.choice(list(alphabet), size=ln).astype(np.uint8))
                if s_ in seen:
                    continue
                seen.add(s_)
                seg.append(s_)
            segs.append(seg)
    return segs


def sample_patterns(rng, segs, n, lens, p_path=0.7):
    """Patterns drawn from random EDS paths (hits) mixed with uniform random ones."""
    strings = [s for seg in segs for s in seg]
    seg_of = [i for i, seg in enumerate(segs) for _ in seg]
    first = np.cumsum([0] + [len(s) for s in segs])
    pats = []
    for _ in range(n):
        m = int(rng.choice(lens))
        if rng.random() < p_path:
            while True:
                w = int(rng.integers(len(strings)))
                if len(strings[w]):
                    break
            r = int(rng.integers(len(strings[w])))
            out = bytearray(strings[w][r:])
            s = seg_of[w] + 1
            while len(out) < m and s < len(segs):
                out += segs[s][int(rng.integers(len(segs[s])))]
                s += 1
            out = out[:m]
            if len(out) < m:
                out += bytes(rng.choice(list(b"ACGT"), size=m - len(out)).astype(np.uint8))
            if rng.random() < 0.1:
                k = int(rng.integers(m))
                out[k] = int(rng.choice(list(b"ACGT")))
            pats.append(bytes(out))
        else:
            pats.append(bytes(rng.choice(list(b"ACGT"), size=m).astype(np.uint8)))
    return pats
