"""Founder-mosaic synthetic MSA generator.

Every cell is a pure O(1) function of (seed, row, column), so the same matrix can be produced on the
host (numpy, here) and on the device (csrc/fsq_synth.cu, fsq_synth_rows) without being stored or
shipped.  The two implementations are bit-identical (tests/test_synth.py).

    founder f has symbol ACGT[h(seed,0,f,col) & 3] at every column;
    row r copies founder h(seed,1,r,(col+phi_r)//R) % F, phi_r = h(seed,2,r,0) % R  (recombination block R);
    a cell is replaced by ACGT[h(seed,5,r,col) & 3] when h(seed,3,r,col) < mu * 2^32;
    with gaps enabled a cell is '-' when h(seed,4,r,col//20) < gap * 2^32 (gap runs of 20 columns).
"""
import numpy as np

_M1 = np.uint64(0xFF51AFD7ED558CCD)
_M2 = np.uint64(0xC4CEB9FE1A85EC53)
_G = np.uint64(0x9E3779B97F4A7C15)
_G2 = np.uint64(0xD6E8FEB86659FD93)
_S33 = np.uint64(33)
_S32 = np.uint64(32)


def _mix(x):
    x = x ^ (x >> _S33)
    x = x * _M1
    x = x ^ (x >> _S33)
    x = x * _M2
    x = x ^ (x >> _S33)
    return x


def hash32(seed, stream, i, j):
    """h(seed, stream, i, j) -> uint32 (numpy broadcasting over i, j)."""
    with np.errstate(over="ignore"):
        s = np.uint64(seed) ^ (np.uint64(stream) * _G)
        x = _mix(s + np.asarray(i, dtype=np.uint64) * _G2)
        x = _mix(x ^ np.asarray(j, dtype=np.uint64))
    return (x >> _S32).astype(np.uint32)


ALPHABET = np.frombuffer(b"ACGT-", dtype=np.uint8)


def synth_rows(m, n, seed, founders, gaps=False, recomb_block=1000, mu=1e-3, gap=2e-3, col0=0):
    """Row-major uint8 matrix (m, n) of ASCII characters for columns [col0, col0+n)."""
    rows = np.arange(m, dtype=np.uint64)[:, None]
    out = np.empty((m, n), dtype=np.uint8)
    mu32 = np.uint32(int(mu * 2.0 ** 32))
    gap32 = np.uint32(int(gap * 2.0 ** 32))
    phi = (hash32(seed, 2, rows, 0) % np.uint32(recomb_block)).astype(np.uint64)
    chunk = max(1, (1 << 22) // max(1, m))
    for c0 in range(0, n, chunk):
        c1 = min(n, c0 + chunk)
        cols = np.arange(col0 + c0, col0 + c1, dtype=np.uint64)[None, :]
        f = hash32(seed, 1, rows, (cols + phi) // np.uint64(recomb_block)) % np.uint32(founders)
        sym = hash32(seed, 0, f, cols) & np.uint32(3)
        mut = hash32(seed, 3, rows, cols) < mu32
        sym = np.where(mut, hash32(seed, 5, rows, cols) & np.uint32(3), sym)
        if gaps:
            g = hash32(seed, 4, rows, cols // np.uint64(20)) < gap32
            sym = np.where(g, np.uint32(4), sym)
        out[:, c0:c1] = ALPHABET[sym]
    return out
