"""Synthetic inputs of SURVEY.md 8(d): residues i.i.d. from Swiss-Prot background frequencies over
the 20 standard letters, log-normal lengths, Philox counter-based PRNG (seeded, reproducible)."""
from __future__ import annotations

import numpy as np

LETTERS = "ARNDCQEGHILKMFPSTWYV"
# Swiss-Prot amino-acid composition (percent), release-independent to the digit shown
FREQ = np.array([8.25, 5.53, 4.06, 5.45, 1.37, 3.93, 6.75, 7.07, 2.27, 5.96,
                 9.66, 5.84, 2.42, 3.86, 4.70, 6.56, 5.34, 1.08, 2.92, 6.87])
FREQ = FREQ / FREQ.sum()
CODES = np.array([ord(c) - 64 for c in LETTERS], dtype=np.int8)   # letters_to_index: A=1 .. Z=26
C2_QUERY_LENGTHS = [144, 189, 222, 375, 464, 567, 657, 729, 850, 1000, 1500, 2005, 2504, 3005, 3564, 4061,
                    4548, 4743, 5147, 5478]


def rng(seed):
    return np.random.Generator(np.random.Philox(seed))


def residues(g, n, freq=FREQ, codes=CODES):
    return codes[g.choice(len(codes), size=int(n), p=freq)]


def lognormal_lengths(g, n, median=290.0, sigma=0.68, lo=20, hi=35000):
    return np.clip(np.floor(g.lognormal(np.log(median), sigma, size=n)), lo, hi).astype(np.int64)


def make_db(seed, n, median=290.0, sigma=0.68, lo=20, hi=35000, total=None, force_max=None, sort=True):
    """Returns (codes int8[sum], offsets int64[n+1]); sorted longest first like the reference demands."""
    g = rng(seed)
    lens = lognormal_lengths(g, n, median, sigma, lo, hi)
    if total is not None:                      # rescale to a residue budget (C2: ~205 M)
        lens = np.clip(np.floor(lens * (total / lens.sum())), lo, hi).astype(np.int64)
    if force_max is not None:
        lens[0] = force_max
    if sort:
        lens = np.sort(lens)[::-1].copy()
    off = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lens, out=off[1:])
    return residues(g, off[-1]), off


def make_query(seed, length):
    return residues(rng(seed), length)


def config1():
    """C1: 360-aa query x 10 000 proteins, seed 42."""
    codes, off = make_db(42, 10000)
    return make_query(1042, 360), codes, off


def config2(n=570000, total=205_000_000):
    """C2: Swiss-Prot-shaped DB, seed 43 (one target forced to 35 213)."""
    return make_db(43, n, lo=2, hi=35213, total=total, force_max=35213)


def config3(n_each=1024, lq=6000):
    """C3: seed 44; 16-bit overflow stress for PAM250, -10/-1.  Query of `lq` residues from a
    low-complexity generator (W 0.4, C 0.2, Y 0.2, F 0.2: an exact copy of length L scores ~13 L and crosses 32 767
    near L ~ 2 520).  4 x n_each targets: (A) exact substrings with L in [3 000, 6 000] -- far overflow; (B) exact
    substrings with L in [2 300, 2 700] -- true scores straddling 32 767; (C) copies of (A)/(B) with 1-5 %
    substitutions and one deletion; (D) C1-style background targets.  Returns (query, codes, offsets), targets
    sorted longest first (the reference demands it)."""
    g = rng(44)
    lc = np.array([23, 3, 25, 6], dtype=np.int8)               # W C Y F
    q = lc[g.choice(4, size=lq, p=[0.4, 0.2, 0.2, 0.2])]
    parts = []
    for _ in range(n_each):
        L = int(g.integers(lq // 2, lq + 1)); s = int(g.integers(0, lq - L + 1)); parts.append(q[s:s + L])
    for _ in range(n_each):
        L = int(g.integers(2300, 2701)); s = int(g.integers(0, lq - L + 1)); parts.append(q[s:s + L])
    for k in range(n_each):
        src = parts[k * 2].copy()
        mut = g.random(src.size) < g.uniform(0.01, 0.05)
        src[mut] = residues(g, int(mut.sum()))
        cut = int(g.integers(10, src.size - 10))
        parts.append(np.concatenate([src[:cut], src[cut + 2:]]))
    bg, bo = make_db(45, n_each)
    parts += [bg[bo[i]:bo[i + 1]] for i in range(n_each)]
    parts.sort(key=lambda p: -p.size)
    off = np.zeros(len(parts) + 1, dtype=np.int64)
    np.cumsum([p.size for p in parts], out=off[1:])
    return q, np.concatenate(parts), off


def config4(n=1024, lq=35213):
    """C4: seed 45; titin-length query; 16 targets at 35 213 aa, 48 log-uniform in [8 k, 35 k],
    the rest C1-distributed; target 1 is a 90 %-identity copy of the query (overflows 16 bits).
    Returns (query, codes, offsets), targets sorted longest first."""
    g = rng(45)
    q = make_query(4500, lq)
    lens = [35213] * 16 + [int(x) for x in np.exp(g.uniform(np.log(8000), np.log(35000), size=48))]
    lens += [int(x) for x in lognormal_lengths(g, max(n - len(lens), 0))]
    lens = sorted(lens[:n], reverse=True)
    parts = [residues(g, L) for L in lens]
    cp = q[:lens[1]].copy()
    mut = g.random(cp.size) < 0.10
    cp[mut] = residues(g, int(mut.sum()))
    parts[1] = cp
    off = np.zeros(len(parts) + 1, dtype=np.int64)
    np.cumsum([p.size for p in parts], out=off[1:])
    return q, np.concatenate(parts), off


def to_fasta(path, codes, off, prefix="sp"):
    letters = np.frombuffer(b"?ABCDEFGHIJKLMNOPQRSTUVWXYZ????*", dtype=np.uint8)
    txt = letters[np.asarray(codes, dtype=np.int64)]
    with open(path, "wb") as f:
        for i in range(len(off) - 1):
            f.write(b">%s|%07d len=%d\n" % (prefix.encode(), i, off[i + 1] - off[i]))
            f.write(txt[off[i]:off[i + 1]].tobytes())
            f.write(b"\n")


def load_matrix(path):
    """Whitespace-separated matrix file -> (int8[32,32] table, defined_diag uint8[32]); mirrors
    src/alignment_scoring_load.c:98-151 + scoring_add_mutation (zero-initialised table)."""
    tab = np.zeros((32, 32), dtype=np.int8)
    cols = None
    def idx(ch):
        if ch == "*":
            return 31
        return ord(ch.upper()) - 64
    with open(path) as f:
        for line in f:
            if line.startswith("#") or not line.strip():
                continue
            if cols is None:
                cols = [idx(c) for c in line.split()]
                continue
            parts = line.split()
            r = idx(parts[0])
            for c, v in zip(cols, parts[1:]):
                tab[r, c] = int(v)
    diag = np.array([1 if (i in cols) else 0 for i in range(32)], dtype=np.uint8)
    return tab, diag


# ---- device-generated databases (swgpu_db_create_synthetic, include/swgpu.h) ----------------
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def synth_spec(freq=FREQ, codes=CODES):
    """(cdf uint32[32], codes int8[32], n_letters): cumulative thresholds out of 65536 for the 16-bit draw."""
    c = np.floor(np.cumsum(freq) * 65536.0 + 0.5).astype(np.int64)
    c[-1] = 65536
    cdf = np.zeros(32, dtype=np.uint32); cdf[: len(c)] = c
    cd = np.zeros(32, dtype=np.int8); cd[: len(codes)] = codes
    return cdf, cd, len(codes)


def _mix64(x):
    x = x ^ (x >> np.uint64(30)); x = x * np.uint64(0xBF58476D1CE4E5B9)
    x = x ^ (x >> np.uint64(27)); x = x * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def synth_target(seed, entry, length, spec=None):
    """Host mirror of the device generator: the residues of the target with global entry id `entry`."""
    cdf, cd, n = spec or synth_spec()
    with np.errstate(over="ignore"):
        key = np.uint64(seed) ^ (np.uint64(entry) * np.uint64(0x9E3779B97F4A7C15))
        j = np.arange(1, int(length) + 1, dtype=np.uint64)
        u = (_mix64(key + j * np.uint64(0xBF58476D1CE4E5B9)) >> np.uint64(48)).astype(np.int64)
    i = np.minimum(np.searchsorted(cdf[:n].astype(np.int64), u, side="right"), n - 1)
    return cd[i]


C5_SEED = 46
C5_TOTAL_RESIDUES = 10_000_000_000


def c5_lengths(scale=1.0, seed=C5_SEED):
    """C5: UniRef50-shaped lengths, LogNormal(ln 230, 0.75) clipped to [11, 35 213] (mean ~305), ~32.8 M targets for
    10 G residues at scale 1; entry 0 is forced to the maximum when the database is big enough to carry it."""
    n = max(64, int(round(C5_TOTAL_RESIDUES * scale / 304.9)))
    lens = lognormal_lengths(rng(seed), n, median=230.0, sigma=0.75, lo=11, hi=35213).astype(np.int32)
    if scale >= 0.001:
        lens[0] = 35213
    return lens


def c5_queries(lens, n_queries=64, seed=C5_SEED, spec=None, min_len=200, identity=0.9):
    """64 queries with lengths cycling through the C2 list.  Query i carries a mutated copy (90 % identity) of the
    first min(Lq, Lt) residues of one database target (entry planted[i], the first target of >= min_len residues at
    or after i * n / n_queries), so every query has a known best hit whatever the size of the database."""
    n = len(lens)
    qs, planted = [], []
    for i in range(n_queries):
        lq = C2_QUERY_LENGTHS[i % len(C2_QUERY_LENGTHS)]
        g = rng(4600 + i)
        q = residues(g, lq)
        e = (i * n) // n_queries
        while lens[e] < min_len:
            e = (e + 1) % n
        t = synth_target(seed, e, int(lens[e]), spec)
        m = min(lq, len(t))
        a = int(g.integers(0, lq - m + 1))
        seg = t[:m].copy()
        mut = g.random(m) >= identity
        seg[mut] = residues(g, int(mut.sum()))
        q[a:a + m] = seg
        qs.append(q); planted.append(e)
    return qs, planted
