"""Seeded synthetic workloads of BASELINE.json's configs.

Everything is integer-valued and stored as float64, like the reference's parser produces
(``int()`` at week-2/diet/simplex.py:712-717).  Problems are generated in chunks of
``CHUNK`` with ``numpy.random.default_rng([seed, chunk])`` so problem ``p`` has the same
data whatever the batch size or the rank that generates it (ranks generate only their
shard; the golden vectors in tests/golden use the first problems of chunk 0).
"""
from __future__ import annotations

import numpy as np

CHUNK = 256
SINGLE_COLS = 1024  # column-block granularity of single_lp


def _chunks(start: int, count: int):
    p = start
    end = start + count
    while p < end:
        ch = p // CHUNK
        lo = p - ch * CHUNK
        hi = min(CHUNK, end - ch * CHUNK)
        yield ch, lo, hi
        p = ch * CHUNK + hi


def simplex_batch(dist: str, count: int, n: int = 100, m: int = 100, seed: int = 2, start: int = 0):
    """Config 4 -- batched simplex.  Returns (A[count,n,m], b[count,n], c[count,m]) float64.

    D1 "feasible-bounded": A ~ U{0..100},    b ~ U{1..10^6}, c ~ U{1..100}
    D2 "mixed":            A ~ U{-100..100}, b ~ U{1..10^6}, c ~ U{-100..100}
    Both have the origin feasible (b > 0), so the reference goes straight to phase 2.
    D3 "phase1":           D2 with b ~ U{-10^5..10^6} \\ {0}: about 9 % negative b, phase 1 runs.
    """
    A = np.empty((count, n, m), dtype=np.float64)
    b = np.empty((count, n), dtype=np.float64)
    c = np.empty((count, m), dtype=np.float64)
    tag = {"D1": 1, "D2": 2, "D3": 3}[dist]
    o = 0
    for ch, lo, hi in _chunks(start, count):
        rng = np.random.default_rng([seed, tag, n, m, ch])
        if dist == "D1":
            Ac = rng.integers(0, 101, size=(CHUNK, n, m))
            bc = rng.integers(1, 10**6 + 1, size=(CHUNK, n))
            cc = rng.integers(1, 101, size=(CHUNK, m))
        else:
            Ac = rng.integers(-100, 101, size=(CHUNK, n, m))
            if dist == "D2":
                bc = rng.integers(1, 10**6 + 1, size=(CHUNK, n))
            else:
                bc = rng.integers(-10**5, 10**6 + 1, size=(CHUNK, n))
                bc[bc == 0] = 1
            cc = rng.integers(-100, 101, size=(CHUNK, m))
        k = hi - lo
        A[o:o + k] = Ac[lo:hi]
        b[o:o + k] = bc[lo:hi]
        c[o:o + k] = cc[lo:hi]
        o += k
    return A, b, c


def stress_batch(count: int, seed: int = 0, zero_b_frac: float = 0.07, nmax: int = 7):
    """Config 2 -- the distribution of week-2/diet/stress.py:23-26 (n, m ~ U{1..7},
    A ~ U{-100..99}, b ~ U{-10^6..10^6-1}, c ~ U{-100..99}) with a fraction of b forced
    to 0 to reach the degenerate / crash paths.  Returns a list of (A, b, c) float64."""
    rng = np.random.default_rng([seed, 77])
    out = []
    for _ in range(count):
        n, m = (int(v) for v in rng.integers(1, nmax + 1, size=2))
        A = rng.integers(-100, 100, size=(n, m)).astype(np.float64)
        b = rng.integers(-1000000, 1000000, size=n).astype(np.float64)
        b[rng.random(n) < zero_b_frac] = 0.0
        c = rng.integers(-100, 100, size=m).astype(np.float64)
        out.append((A, b, c))
    return out


def gauss_jordan_batch(count: int, n: int = 512, seed: int = 1, swap_frac: float = 0.01, start: int = 0):
    """Config 3 -- batched Gauss-Jordan.  a = U(-1,1) + n*I (diagonally dominant: no zero
    pivots, well conditioned), b = U(-1,1).  A fraction ``swap_frac`` of the systems gets
    an exact zero planted at a[k][k] for one k < n-1 (rows k and k+1 are first exchanged so
    the system stays well conditioned) to exercise the reference's row swap
    (energy_values.cc:100-110).  Returns (a[count,n,n], b[count,n]) float64."""
    a = np.empty((count, n, n), dtype=np.float64)
    b = np.empty((count, n), dtype=np.float64)
    gchunk = 16
    o = 0
    p = start
    end = start + count
    while p < end:
        ch = p // gchunk
        lo = p - ch * gchunk
        hi = min(gchunk, end - ch * gchunk)
        rng = np.random.default_rng([seed, 33, n, ch])
        ac = rng.uniform(-1.0, 1.0, size=(gchunk, n, n))
        bc = rng.uniform(-1.0, 1.0, size=(gchunk, n))
        ac[:, np.arange(n), np.arange(n)] += float(n)
        plant = rng.random(gchunk) < swap_frac
        ks = rng.integers(0, max(n - 1, 1), size=gchunk)
        if n >= 2:
            for s in np.nonzero(plant)[0]:
                k = int(ks[s])
                ac[s, [k, k + 1]] = ac[s, [k + 1, k]]   # big entries now at (k,k+1) and (k+1,k)
                ac[s, k, k] = 0.0
        k2 = hi - lo
        a[o:o + k2] = ac[lo:hi]
        b[o:o + k2] = bc[lo:hi]
        o += k2
        p = ch * gchunk + hi
    return a, b


def single_lp(n: int = 16383, m: int = 32767, seed: int = 5, dist: str = "D1", col_lo: int = 0,
              col_cnt: int | None = None):
    """Config 5 -- one large dense LP (condensed tableau (n+1) x (m+1)); D1-style data so the
    origin is feasible.  A is generated in blocks of ``SINGLE_COLS`` columns, each from its own
    seeded stream, so a rank of the column-sharded solver can generate just ITS columns
    (``col_lo``, ``col_cnt``) and still hold exactly the data of the full matrix.  Returns
    (A[n, col_cnt], b[n], c[m]); b and c are always complete."""
    if col_cnt is None:
        col_cnt = m - col_lo
    lo_a = 0 if dist == "D1" else -100
    A = np.empty((n, col_cnt), dtype=np.float64)
    blk0 = col_lo // SINGLE_COLS
    blk1 = (col_lo + col_cnt + SINGLE_COLS - 1) // SINGLE_COLS if col_cnt > 0 else blk0
    for blk in range(blk0, blk1):
        k0 = blk * SINGLE_COLS
        w = min(SINGLE_COLS, m - k0)
        rng = np.random.default_rng([seed, 55, n, m, blk])
        part = rng.integers(lo_a, 101, size=(n, w), dtype=np.int16)
        a = max(col_lo, k0)
        z = min(col_lo + col_cnt, k0 + w)
        A[:, a - col_lo:z - col_lo] = part[:, a - k0:z - k0]
    rng = np.random.default_rng([seed, 56, n, m])
    b = rng.integers(1, 10**6 + 1, size=n).astype(np.float64)
    if dist == "D1":
        c = rng.integers(1, 101, size=m).astype(np.float64)
    else:
        c = rng.integers(-100, 101, size=m).astype(np.float64)
    return A, b, c
