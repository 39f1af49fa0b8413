"""TEST INFRASTRUCTURE (oracle) -- CPU restatement of the MOVE_EDSBWTSearch path.  Pure Python / numpy: small cases only.

PARITY STATUS.  The arithmetic of the move structure lives in the un-vendored dependency Move-r
(git@github.com:riccardo-nozza/Move-r.git @ 418a11f7fab3cdff14c991ae2617df9ba6017ef0, plus its patched-files overlay,
build.sh:3-4), which is absent from /root/reference (empty submodule), and the reference holds no test or golden output
for MOVE_EDSBWTSearch.  So for the `_MLF.aux` byte format and for the interval boundaries after balancing: PARITY
UNPINNED.  What IS pinned: (1) the run / LF-pair inputs of the table follow the reference's own writer
(da_to_everything.cpp:62-109, 435-463) literally; (2) the move structure is restated from its published definition
(Nishimoto & Tabei, "Optimal-time queries on BWT-runs compressed indexes", ICALP 2021: disjoint input intervals
[p_i, p_{i+1}) mapped to [q_i, q_i + d_i), move(i, x) = (q_x + i - p_x, interval containing it), a-balanced = no output
interval holds >= 2a input starts); (3) the search restated here from MOVE_EDSBWTSearch.cpp:228-625 must emit exactly the
occurrence tuples EDSBWTsearch emits (a10 and a13 print the same five columns for the same rows) -- tests/test_move.py
checks this restatement against every golden vector produced by the unmodified EDSBWTsearch.
Only tests/ may import this module.
"""
import bisect

import numpy as np


# ---- inputs of the table: runs and (run head, LF(run head)) pairs ------------------------------------------------
def runs_of(ebwt: bytes, term=ord("#")):
    """X_runs.aux (da_to_everything.cpp:435-463): a new run starts when the character changes OR the character is the
    end-marker (every end-marker is its own run).  -> (heads, symbols)"""
    heads, syms = [], []
    prev = None
    for i, ch in enumerate(ebwt):
        if prev is None or ch != prev or ch == term:
            heads.append(i)
            syms.append(ch)
        prev = ch
    return heads, syms


def lf_pairs(ebwt: bytes, heads, syms, alphabet: bytes):
    """X_runs.txt (build_ilf, da_to_everything.cpp:62-109): LF(head) = start of the pile of its symbol + number of that
    symbol before the head (running per-symbol counts advanced by the run lengths)."""
    start, acc = {}, 0
    for a in alphabet:                      # get_char_start_pos: pile starts in alphabet order
        start[a] = acc
        acc += ebwt.count(bytes([a]))
    out = []
    n = len(ebwt)
    for j, (h, s) in enumerate(zip(heads, syms)):
        out.append(start[s])
        nxt = heads[j + 1] if j + 1 < len(heads) else n
        start[s] += nxt - h
    return out


# ---- the move structure ------------------------------------------------------------------------------------------
class MoveTable:
    """p[i] = first row of input interval i (p[r] = n), q[i] = LF(p[i]), dest[i] = input interval containing q[i],
    L[i] = symbol of the interval."""

    def __init__(self, ebwt: bytes, alphabet: bytes, a=8):
        self.n = len(ebwt)
        heads, syms = runs_of(ebwt, alphabet[0])
        q = lf_pairs(ebwt, heads, syms, alphabet)
        self.r_runs = len(heads)
        # a-balancing, restated sequentially: while some output interval holds >= 2a input starts, cut it after the a-th
        if a is not None:
            changed = True
            while changed:
                changed = False
                i = 0
                while i < len(heads):
                    d = (heads[i + 1] if i + 1 < len(heads) else self.n) - heads[i]
                    lo = bisect.bisect_left(heads, q[i])
                    hi = bisect.bisect_left(heads, q[i] + d)
                    if hi - lo >= 2 * a:
                        cut = heads[lo + a] - q[i]              # offset of the a-th contained start inside the output interval
                        pos = bisect.bisect_left(heads, heads[i] + cut)
                        heads.insert(pos, heads[i] + cut); syms.insert(pos, syms[i]); q.insert(pos, q[i] + cut)
                        changed = True
                    else:
                        i += 1
        self.p = heads + [self.n]
        self.q = q
        self.L = syms
        self.r = len(heads)
        self.dest = [bisect.bisect_right(self.p, x, 0, self.r) - 1 for x in q]
        self.a = a

    def move(self, i, x):
        """M_LF.move: (row, its interval) -> (LF(row), its interval)."""
        i2 = self.q[x] + (i - self.p[x])
        x2 = self.dest[x]
        while i2 >= self.p[x2 + 1]:
            x2 += 1
        return i2, x2

    def interval_of(self, row):
        return bisect.bisect_right(self.p, row, 0, self.r) - 1

    def check(self, ebwt: bytes, alphabet: bytes):
        """Invariants any valid table satisfies (used on the table the GPU builds as well)."""
        assert self.p[0] == 0 and all(self.p[i] < self.p[i + 1] for i in range(self.r)) and self.p[self.r] == self.n
        term = alphabet[0]
        for i in range(self.r):
            seg = ebwt[self.p[i]:self.p[i + 1]]
            assert seg == bytes([self.L[i]]) * len(seg), ("interval not inside one run", i)
            assert self.L[i] != term or len(seg) == 1, ("end-marker run longer than 1", i)
            assert self.p[self.dest[i]] <= self.q[i] < self.p[self.dest[i] + 1]
        # q = LF(p): piles in alphabet order, rank of the symbol before the row
        start, acc = {}, 0
        for a in alphabet:
            start[a] = acc
            acc += ebwt.count(bytes([a]))
        seen = {a: 0 for a in alphabet}
        j = 0
        for row, ch in enumerate(ebwt):
            if j < self.r and self.p[j] == row:
                assert self.q[j] == start[ch] + seen[ch], ("q != LF(p)", j)
                j += 1
            seen[ch] += 1
        if self.a is not None:
            for i in range(self.r):
                d = self.p[i + 1] - self.p[i]
                cnt = bisect.bisect_left(self.p, self.q[i] + d, 0, self.r) - bisect.bisect_left(self.p, self.q[i], 0, self.r)
                assert cnt < 2 * self.a, ("output interval holds %d >= 2a starts" % cnt, i)


# ---- the search (MOVE_EDSBWTSearch.cpp:228-625) ---------------------------------------------------------------------
class MoveSearch:
    def __init__(self, ix, a=8):
        """ix: oracle.eds_index.Index"""
        self.alphabet = bytes(ix.alphabet)
        self.term = self.alphabet[0]
        self.ebwt = b"".join(ix.piles)
        self.n = len(self.ebwt)
        self.M = MoveTable(self.ebwt, self.alphabet, a)
        self.r = self.M.r
        # _RS_L_: rank/select over L' (one symbol per input interval), MOVE_EDSBWTSearch.cpp:51-53
        self.pos = {c: [i for i, s in enumerate(self.M.L) if s == c] for c in self.alphabet}
        # EOF_ID_Copy: EOF_ID flattened in pile order (:730-736)
        self.eof_copy = [int(w) for pile in ix.eof_id for w in pile]
        bits = [int(b) for b in ix.seg_first]
        self.n_text = len(bits)
        self.sel1 = [i for i, b in enumerate(bits) if b] + [self.n_text]      # select(1, k) = sel1[k-1]
        self.rank1 = np.concatenate([[0], np.cumsum(bits)]).tolist()          # rank(1, i) = ones in [0, i)
        self.fsi = (self.sel1[1] - 1) if len(self.sel1) > 2 else self.n_text - 1     # first_symbol_index (:94)
        # M_LF_Dollar_Input_interval (retrieve_MLF / findInputInterval, :200-218)
        self.dollar_ii = [self.M.interval_of(i) for i in range(self.n_text)]

    def _rank(self, c, i):
        return bisect.bisect_left(self.pos[c], i)

    def _update_single(self, c, iv):
        """updateSingleInterval (:424-510) -> new (b, e, b_, e_) or None"""
        b, e, b_, e_ = iv
        if not self.pos.get(c):
            return None
        if c != self.M.L[b_]:
            k = self._rank(c, b_)
            if k == len(self.pos[c]):
                return None
            b_ = self.pos[c][k]
            if b_ > e_:
                return None
            b = self.M.p[b_]
        if c != self.M.L[e_]:
            e_ = self.pos[c][self._rank(c, e_) - 1]
            e = self.M.p[e_ + 1] - 1
        if b > e:
            return None
        if b_ == e_:
            diff = e - b
            b, b_ = self.M.move(b, b_)
            e, e_ = b + diff, b_
            while e >= self.M.p[e_ + 1]:
                e_ += 1
        else:
            b, b_ = self.M.move(b, b_)
            e, e_ = self.M.move(e, e_)
        return (b, e, b_, e_)

    def _step(self, c, ranges):
        """backward_search_step (:376-422); returns (bool as the reference computes it, surviving ranges)"""
        valid = [x for x in (self._update_single(c, iv) for iv in ranges) if x is not None]
        return (len(ranges) > 0), valid                  # quirk: tests emptiness BEFORE the swap

    def _dollars(self, d, b_, e_):
        l, u = self._rank(self.term, b_), self._rank(self.term, e_ + 1)
        for k in range(l, u):
            w = self.eof_copy[k]
            if w > self.fsi:
                d.append(w)

    def _link(self, F):
        d = []
        for (_, _, b_, e_) in F:
            if b_ <= e_:
                self._dollars(d, b_, e_)
        d.sort()
        dol = []
        while d:
            i = d.pop()
            a = self.rank1[min(i + 1, self.n_text)]
            start, end = self.sel1[a - 2], self.sel1[a - 1] - 1
            cur = (start, end, self.dollar_ii[start], self.dollar_ii[end])
            if dol and cur[1] + 1 >= dol[-1][0]:
                if cur[0] < dol[-1][0]:
                    dol[-1] = (cur[0], dol[-1][1], cur[2], dol[-1][3])
            else:
                dol.append(cur)
            self._dollars(d, cur[2], cur[3])
        dol.reverse()
        return dol

    def search(self, pattern: bytes, pat_no=1):
        """-> (num_occ, rows in the order MOVE_EDSBWTSearch writes them)"""
        m = len(pattern)
        if m == 0 or any(ch not in self.alphabet[1:] for ch in pattern):
            return 0, []                                  # same deviation as the EDSBWTsearch restatement
        ok, F = self._step(pattern[m - 1], [(0, self.n - 1, 0, self.r - 1)])
        if not ok:
            return 0, []
        for pos in range(m - 1, 0, -1):
            dol = self._link(F)
            c = pattern[pos - 1]
            found = False
            if dol:
                ok, dol = self._step(c, dol)
                found = found or ok
            ok, F = self._step(c, F)
            found = found or ok
            if not found:
                return 0, []
            F = dol + F
            if len(F) > 1:                                # merge adjacent over the whole list (:309-324)
                merged = [F[0]]
                for iv in F[1:]:
                    last = merged[-1]
                    if iv[0] == last[1] + 1:
                        merged[-1] = (last[0], iv[1], last[2], iv[3])
                    else:
                        merged.append(iv)
                F = merged
        rows, num = [], 0
        for (start, end, sii, _) in F:                    # locate (:328-369)
            num += end - start + 1
            ii_copy = sii
            for j in range(start, end + 1):
                if not (self.M.p[ii_copy + 1] - j > 0):
                    ii_copy += 1
                position, ii, off = j, ii_copy, 0
                while self.M.L[ii] != self.term:
                    position, ii = self.M.move(position, ii)
                    off += 1
                w = self.eof_copy[self._rank(self.term, ii)]
                seg = self.rank1[w + 1]
                rows.append((pat_no, w, seg, w - self.sel1[seg - 1], off))
        return num, rows
