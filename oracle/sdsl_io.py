"""TEST INFRASTRUCTURE (oracle) -- never imported by the product path.

Reader/writer for the two sdsl-lite serialisations the EDS-BWT index uses
(SURVEY.md App. A.3 / A.4):

  * ``bit_vector``  (= int_vector<1>)       -- <base>.bitvector
        written by eds_to_fasta.cpp:151-153, read by EDSBWTsearch.cpp:158
  * ``rrr_vector<63, int_vector<>, 32>``    -- <base>_bv_<p>.aux
        written by da_to_everything.cpp:234-236, read by EDSBWTsearch.cpp:1598

sdsl-lite itself is not in /root/reference (third-party, un-vendored, v2.1.x
behaviour); the layout below restates its published rrr_vector construction and
is pinned by feeding files written here to the prebuilt reference binary
(tests/test_oracle_vs_reference.py) and by the byte dumps of SURVEY App. D.4.
"""
import struct
from math import comb

import numpy as np

T_BS = 63  # block size of rrr_vector<>   (sdsl default t_bs=63)
T_K = 32   # blocks per superblock        (sdsl default t_k=32)

_BINOM = [[comb(n, k) for k in range(64)] for n in range(64)]
_SPACE = [0 if comb(T_BS, k) == 1 else (comb(T_BS, k).bit_length()) for k in range(T_BS + 1)]
# space(k) = floor(log2 C(63,k)) + 1 == bit_length(C(63,k)) for C > 1


def hi(x: int) -> int:
    """sdsl bits::hi -- index of the most significant set bit, hi(0) = 0."""
    return x.bit_length() - 1 if x > 0 else 0


# ----------------------------------------------------------------------------
# int_vector helpers
# ----------------------------------------------------------------------------
def _pack_bits_to_words(bits: np.ndarray) -> np.ndarray:
    """bits (0/1 uint8, LSB-first) -> little-endian u64 words, ceil(len/64) words."""
    nb = len(bits)
    nwords = (nb + 63) // 64
    padded = np.zeros(nwords * 64, dtype=np.uint8)
    padded[:nb] = bits
    by = np.packbits(padded, bitorder="little")
    return by.view("<u8")


def write_bit_vector(bits) -> bytes:
    """sdsl int_vector<1>::serialize: u64 size-in-bits, then ceil(size/64) u64 words."""
    bits = np.asarray(bits, dtype=np.uint8)
    return struct.pack("<Q", len(bits)) + _pack_bits_to_words(bits).tobytes()


def read_bit_vector(buf: bytes, off: int = 0):
    (nbits,) = struct.unpack_from("<Q", buf, off)
    nwords = (nbits + 63) // 64
    words = np.frombuffer(buf, dtype="<u8", count=nwords, offset=off + 8)
    bits = np.unpackbits(words.view(np.uint8), bitorder="little")[:nbits]
    return bits.astype(np.uint8), off + 8 + 8 * nwords


def _write_int_vector(values, width: int) -> bytes:
    """sdsl int_vector<0>::serialize: u64 size-in-bits, u8 width, packed payload."""
    nbits = len(values) * width
    acc = 0
    for j, v in enumerate(values):
        acc |= int(v) << (j * width)
    nwords = (nbits + 63) // 64
    return struct.pack("<QB", nbits, width) + acc.to_bytes(nwords * 8, "little")


def _read_int_vector(buf: bytes, off: int):
    nbits, width = struct.unpack_from("<QB", buf, off)
    nwords = (nbits + 63) // 64
    acc = int.from_bytes(buf[off + 9: off + 9 + 8 * nwords], "little")
    cnt = nbits // width if width else 0
    mask = (1 << width) - 1
    vals = [(acc >> (j * width)) & mask for j in range(cnt)]
    return vals, width, off + 9 + 8 * nwords


# ----------------------------------------------------------------------------
# rrr_vector<63>
# ----------------------------------------------------------------------------
def _bin_to_nr(bin_: int) -> int:
    """Combinatorial-number-system rank of a 63-bit block (sdsl rrr_helper::bin_to_nr)."""
    k = bin(bin_).count("1")
    if k == 0 or k == T_BS:
        return 0
    nr = 0
    pos = 0
    while bin_:
        if bin_ & 1:
            nr += _BINOM[T_BS - pos - 1][k]
            k -= 1
        bin_ >>= 1
        pos += 1
    return nr


def _nr_to_bin(k: int, nr: int) -> int:
    if k == 0:
        return 0
    if k == T_BS:
        return (1 << T_BS) - 1
    out = 0
    for pos in range(T_BS):
        c = _BINOM[T_BS - pos - 1][k]
        if nr >= c:
            out |= 1 << pos
            nr -= c
            k -= 1
            if k == 0:
                break
    return out


def write_rrr_vector(bits) -> bytes:
    """Serialise a bitmap as sdsl rrr_vector<63,int_vector<>,32>."""
    bits = np.asarray(bits, dtype=np.uint8)
    m_size = len(bits)
    nblocks = m_size // T_BS + 1            # includes a dummy block when m_size % 63 == 0
    nsb = (nblocks + T_K - 1) // T_K
    big = int.from_bytes(np.packbits(bits, bitorder="little").tobytes(), "little") if m_size else 0
    mask63 = (1 << T_BS) - 1
    blocks = [(big >> (T_BS * j)) & mask63 for j in range(nblocks)]
    bt = [bin(b).count("1") for b in blocks]
    total_code_bits = sum(_SPACE[k] for k in bt)
    total_ones = sum(bt)
    btnr = 0
    btnr_pos = 0
    btnrp = [0] * nsb
    n_rank = nsb + (1 if (m_size % (T_K * T_BS)) > 0 else 0)
    rank = [0] * n_rank
    invert = [0] * nsb
    sum_rank = 0
    inv = False
    # number of *real* blocks: full ones plus the partial tail (the dummy block is never coded)
    n_real = m_size // T_BS + (1 if m_size % T_BS else 0)
    for i in range(n_real):
        full = (i + 1) * T_BS <= m_size
        if i % T_K == 0:
            btnrp[i // T_K] = btnr_pos
            rank[i // T_K] = sum_rank
            inv = False
            # sdsl considers a superblock for inversion when all 32 m_bt entries exist
            # (partial tail and dummy block included).
            if full and i + T_K <= nblocks:
                gt_half = sum(1 for j in range(i, i + T_K) if bt[j] > T_BS // 2)
                if gt_half > T_K // 2:
                    invert[i // T_K] = 1
                    for j in range(i, i + T_K):
                        bt[j] = T_BS - bt[j]
                    inv = True
        x = bt[i]
        sp = _SPACE[x]
        sum_rank += (T_BS - x) if inv else x
        if sp:
            # PROBED against the prebuilt reference binary: "inversion" only flips the stored
            # block type (m_bt = 63 - popcount); the block code is that of the ORIGINAL block.
            # (SURVEY App. A.4 says the block is complemented; that only holds for the
            # all-ones/all-zeros blocks its probes used, where no code bits are stored.)
            btnr |= _bin_to_nr(blocks[i]) << btnr_pos
        btnr_pos += sp
    # When the dummy block opens a superblock of its own (m_size % 2016 == 0) sdsl never
    # visits it: its m_btnrp entry stays 0 and its m_rank entry is the "finishing" element.
    rank[n_rank - 1] = sum_rank
    assert btnr_pos == total_code_bits and sum_rank == total_ones
    out = struct.pack("<Q", m_size)
    out += _write_int_vector(bt, hi(T_BS) + 1)
    nb = max(total_code_bits, 64)
    out += struct.pack("<Q", nb) + btnr.to_bytes(((nb + 63) // 64) * 8, "little")
    out += _write_int_vector(btnrp, hi(total_code_bits) + 1)
    out += _write_int_vector(rank, hi(total_ones) + 1)
    out += write_bit_vector(np.array(invert, dtype=np.uint8))
    return out


def read_rrr_vector(buf: bytes) -> np.ndarray:
    """Decode an rrr_vector<63> file to a plain 0/1 array (reader procedure of App. A.4)."""
    (m_size,) = struct.unpack_from("<Q", buf, 0)
    off = 8
    bt, _, off = _read_int_vector(buf, off)
    (nb,) = struct.unpack_from("<Q", buf, off)
    nw = (nb + 63) // 64
    btnr = int.from_bytes(buf[off + 8: off + 8 + 8 * nw], "little")
    off += 8 + 8 * nw
    btnrp, _, off = _read_int_vector(buf, off)
    rank, _, off = _read_int_vector(buf, off)
    invert, off = read_bit_vector(buf, off)
    out = np.zeros(m_size, dtype=np.uint8)
    n_real = (m_size + T_BS - 1) // T_BS
    pos = 0
    mask63 = (1 << T_BS) - 1
    for j in range(n_real):
        if j % T_K == 0:
            pos = btnrp[j // T_K]
        k = bt[j]
        ln = _SPACE[k]
        nr = (btnr >> pos) & ((1 << ln) - 1)
        pos += ln
        if invert[j // T_K]:
            k = T_BS - k          # stored type is flipped, the code is of the original block
        b = _nr_to_bin(k, nr)
        nbits = min(T_BS, m_size - T_BS * j)
        for t in range(nbits):
            out[T_BS * j + t] = (b >> t) & 1
    if len(rank):
        assert int(out.sum()) == rank[-1], "rrr: m_rank tail does not equal the decoded popcount"
    return out
