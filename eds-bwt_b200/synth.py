"""Seeded synthetic workloads of BASELINE.json / SURVEY.md 8(d): elastic-degenerate strings and pattern sets.

Host-side numpy only (no compute path): the outputs are the HOST arrays handed to
edsbwt_build_from_collection / edsbwt_search through the C ABI.  An EDS is kept as a flat collection

    letters   uint8   all strings back to back, no separators
    str_off   uint64  [nstr+1] string w = letters[str_off[w]:str_off[w+1]]
    seg_first uint8   [nstr]   1 = first string of its degenerate segment (the reference's .bitvector,
                               eds_to_fasta.cpp:69,98,118)
"""
import numpy as np

ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)


class Collection:
    def __init__(self, letters, str_off, seg_first, name=""):
        self.letters = np.ascontiguousarray(letters, dtype=np.uint8)
        self.str_off = np.ascontiguousarray(str_off, dtype=np.uint64)
        self.seg_first = np.ascontiguousarray(seg_first, dtype=np.uint8)
        self.name = name

    @property
    def nstr(self):
        return len(self.seg_first)

    @property
    def n_rows(self):
        return int(self.str_off[-1]) + self.nstr

    @property
    def n_seg(self):
        return int(self.seg_first.sum())

    def lens(self):
        return np.diff(self.str_off.astype(np.int64))

    def segments(self):
        """-> list of lists of bytes (small inputs only)."""
        segs = []
        lb = self.letters.tobytes()
        for w in range(self.nstr):
            s = lb[int(self.str_off[w]):int(self.str_off[w + 1])]
            if self.seg_first[w]:
                segs.append([s])
            else:
                segs[-1].append(s)
        return segs

    def eds_text(self):
        return b"".join(b"{" + b",".join(seg) + b"}" for seg in self.segments())


def from_segments(segs, name=""):
    strings = [s for seg in segs for s in seg]
    seg_first = np.zeros(len(strings), dtype=np.uint8)
    k = 0
    for seg in segs:
        seg_first[k] = 1
        k += len(seg)
    off = np.zeros(len(strings) + 1, dtype=np.uint64)
    off[1:] = np.cumsum([len(s) for s in strings], dtype=np.uint64)
    letters = np.frombuffer(b"".join(strings), dtype=np.uint8)
    return Collection(letters, off, seg_first, name)


def _interleave(nondeg_len, v, allele_len):
    """strings in file order: nondeg_0, alleles of site 0, nondeg_1, ..., nondeg_Nd  -> (lens, seg_first)"""
    nd = len(v)
    cv = np.zeros(nd + 1, dtype=np.int64)
    np.cumsum(v, out=cv[1:])
    total = nd + 1 + int(cv[-1])
    lens = np.empty(total, dtype=np.int64)
    seg_first = np.zeros(total, dtype=np.uint8)
    idx_nondeg = np.arange(nd + 1, dtype=np.int64) + cv
    is_nondeg = np.zeros(total, dtype=bool)
    is_nondeg[idx_nondeg] = True
    lens[is_nondeg] = nondeg_len
    lens[~is_nondeg] = allele_len
    seg_first[idx_nondeg] = 1
    seg_first[idx_nondeg[:-1] + 1] = 1          # first allele of every site
    return lens, seg_first


def chromosome_eds(mbp=250.0, seed=4):
    """cfg4: non-degenerate stretches geometric with mean 100 (min 1) totalling ~mbp*1e6 bases,
    one variant site between consecutive stretches ("1 % degenerate" = one site per 100 bp), v in U{2..8}
    alleles per site with lengths 70 % 1, 15 % 0 (at most one empty allele per site), 15 % U{2..8};
    letters i.i.d. uniform over ACGT.  First and last segments are non-degenerate."""
    rng = np.random.default_rng(seed)
    nd = max(1, int(round(mbp * 1e6 / 100.0)) - 1)
    nondeg_len = rng.geometric(0.01, nd + 1).astype(np.int64)
    v = rng.integers(2, 9, nd).astype(np.int64)
    ns = int(v.sum())
    u = rng.random(ns)
    allele_len = np.where(u < 0.70, 1, np.where(u < 0.85, 0, rng.integers(2, 9, ns))).astype(np.int64)
    # at most one empty allele per site: every later one becomes a length-1 allele
    z = (allele_len == 0).astype(np.int64)
    cz = np.cumsum(z)
    site_start = np.zeros(nd, dtype=np.int64)
    np.cumsum(v[:-1], out=site_start[1:])
    before_site = np.repeat(cz[site_start] - z[site_start], v)
    allele_len[(z == 1) & (cz - before_site > 1)] = 1
    lens, seg_first = _interleave(nondeg_len, v, allele_len)
    off = np.zeros(len(lens) + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:].view(np.int64))
    letters = ACGT[rng.integers(0, 4, int(off[-1]), dtype=np.uint8)]
    return Collection(letters, off, seg_first, "chromosome_%gMbp_seed%d" % (mbp, seed))


def sarscov2_like_eds(n_sites=5000, seed=3):
    """cfg3: n_sites degenerate segments alternating with n_sites+1 non-degenerate ones of
    geometric length (mean 6, min 1); each site has v in U{2,3,4} distinct alleles: a length-1 reference
    allele, the others 75 % SNP, 15 % empty (at most one), 10 % insertion of length U{2..6}."""
    rng = np.random.default_rng(seed)
    segs = []
    for s in range(n_sites + 1):
        ln = int(rng.geometric(1.0 / 6.0))
        segs.append([ACGT[rng.integers(0, 4, ln)].tobytes()])
        if s == n_sites:
            break
        v = int(rng.integers(2, 5))
        ref = int(rng.integers(0, 4))
        alleles = [bytes([ACGT[ref]])]
        snps = [c for c in range(4) if c != ref]
        rng.shuffle(snps)
        has_empty = False
        while len(alleles) < v:
            u = rng.random()
            if u < 0.75 and snps:
                alleles.append(bytes([ACGT[snps.pop()]]))
            elif u < 0.90 and not has_empty:
                alleles.append(b"")
                has_empty = True
            else:
                ins = ACGT[rng.integers(0, 4, int(rng.integers(2, 7)))].tobytes()
                if ins not in alleles:
                    alleles.append(ins)
        segs.append(alleles)
    return from_segments(segs, "sarscov2_like_%d_seed%d" % (n_sites, seed))


def repetitive_eds(unit_len=30000, copies=100, divergence=0.001, site_every=100, seed=7):
    """A repetitive EDS for the MOVE_EDSBWTSearch variant: the non-degenerate backbone is `copies` copies
    of one random unit of unit_len bases, every copy carrying its own substitutions at rate `divergence` (a pangenome laid end
    to end / segmental duplications -- what makes the eBWT run-compressible; an EDS built from an alignment of near-identical
    genomes is NOT repetitive by itself, the alignment has already collapsed the copies), cut by one variant site every
    ~site_every bases with 2-4 alleles of length 1 (75 %), 0 (15 %, at most one) or 2-6 (10 %), like cfg3/cfg4."""
    rng = np.random.default_rng(seed)
    unit = ACGT[rng.integers(0, 4, unit_len, dtype=np.uint8)]
    backbone = np.tile(unit, copies)
    nmut = int(round(len(backbone) * divergence))
    if nmut:
        where = rng.integers(0, len(backbone), nmut)
        backbone[where] = ACGT[(np.searchsorted(ACGT, backbone[where]) + rng.integers(1, 4, nmut)) % 4]
    total = len(backbone)
    nd = max(1, total // site_every - 1)
    cuts = np.sort(rng.choice(np.arange(1, total), nd, replace=False))
    nondeg_len = np.diff(np.concatenate([[0], cuts, [total]])).astype(np.int64)
    v = rng.integers(2, 5, nd).astype(np.int64)
    ns = int(v.sum())
    u = rng.random(ns)
    allele_len = np.where(u < 0.75, 1, np.where(u < 0.90, 0, rng.integers(2, 7, ns))).astype(np.int64)
    z = (allele_len == 0).astype(np.int64)
    cz = np.cumsum(z)
    site_start = np.zeros(nd, dtype=np.int64)
    np.cumsum(v[:-1], out=site_start[1:])
    before_site = np.repeat(cz[site_start] - z[site_start], v)
    allele_len[(z == 1) & (cz - before_site > 1)] = 1
    lens, seg_first = _interleave(nondeg_len, v, allele_len)
    off = np.zeros(len(lens) + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:].view(np.int64))
    letters = ACGT[rng.integers(0, 4, int(off[-1]), dtype=np.uint8)]
    # the non-degenerate strings carry the backbone, in order
    is_nondeg = np.zeros(len(lens), dtype=bool)
    cv = np.zeros(nd + 1, dtype=np.int64)
    np.cumsum(v, out=cv[1:])
    is_nondeg[np.arange(nd + 1, dtype=np.int64) + cv] = True
    dst = np.repeat(off[:-1].astype(np.int64)[is_nondeg], lens[is_nondeg]) + (np.arange(total, dtype=np.int64) - np.repeat(np.cumsum(lens[is_nondeg]) - lens[is_nondeg], lens[is_nondeg]))
    letters[dst] = backbone
    return Collection(letters, off, seg_first, "repetitive_%dx%d_div%g_seed%d" % (copies, unit_len, divergence, seed))


def haplotype(col, rng):
    """One random path through the EDS (uniform allele choice per segment) as a letter array."""
    lens = col.lens()
    first = np.flatnonzero(col.seg_first)
    count = np.diff(np.append(first, col.nstr))
    pick = first + np.minimum((rng.random(len(first)) * count).astype(np.int64), count - 1)
    sl = lens[pick]
    out_off = np.zeros(len(pick) + 1, dtype=np.int64)
    np.cumsum(sl, out=out_off[1:])
    src0 = col.str_off.astype(np.int64)[pick]
    idx = np.repeat(src0 - out_off[:-1], sl) + np.arange(int(out_off[-1]), dtype=np.int64)
    return col.letters[idx]


def sample_patterns(col, npat, m, seed, frac_path=0.9, frac_mut=0.1, n_hap=3):
    """npat patterns of length m as one uint8 array [npat, m]: frac_path of them substrings of random EDS
    paths (uniform start, uniform allele choice), frac_mut of THOSE given one random substitution, the rest
    uniform random ACGT (SURVEY 8(d) "90/10 recipe")."""
    rng = np.random.default_rng(seed)
    out = ACGT[rng.integers(0, 4, (npat, m), dtype=np.uint8)]
    n_path = int(round(npat * frac_path))
    per = [n_path // n_hap + (1 if i < n_path % n_hap else 0) for i in range(n_hap)]
    ar = np.arange(m, dtype=np.int64)
    row = 0
    for cnt in per:
        if cnt == 0:
            continue
        H = haplotype(col, rng)
        if len(H) < m:
            row += cnt
            continue
        st = rng.integers(0, len(H) - m + 1, cnt)
        out[row:row + cnt] = H[st[:, None] + ar[None, :]]
        row += cnt
    n_mut = int(round(n_path * frac_mut))
    if n_mut:
        who = rng.choice(n_path, n_mut, replace=False)
        where = rng.integers(0, m, n_mut)
        cur = out[who, where]
        code = np.searchsorted(ACGT, cur)
        out[who, where] = ACGT[(code + rng.integers(1, 4, n_mut)) % 4]
    perm = rng.permutation(npat)
    return np.ascontiguousarray(out[perm])


def pack_fixed(pats):
    """[npat, m] uint8 -> (bytes array, uint64 offsets) of the C ABI"""
    npat, m = pats.shape
    off = np.arange(npat + 1, dtype=np.uint64) * np.uint64(m)
    return np.ascontiguousarray(pats).reshape(-1), off
