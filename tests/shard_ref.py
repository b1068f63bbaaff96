"""numpy restatement of the sample-sharded exchange layout (tests only; mirrors bvc_shard.cuh's dense record).

dense record of a site, B = 16 + 8*Qw u32 words, Qw = qhi - qlo + 1:
  [0..7]  indel entries by (strand << 2 | stale base)      [8] n_entries      [9..15] 0
  [16 + ((b*Qw + (q-qlo))*2 + strand)]  non-indel entries of base b, qual q, strand
`expand` turns merged records back into a per-read tile (entries grouped by class), so the per-read oracle can be
run on exchanged statistics: the histograms are sufficient for everything but the rank-sum features.
"""
import numpy as np

from basevarc_b200._abi import make_code
from basevarc_b200.pileup import PileupTile

HDR = 16


def dense_hist(tile, qlo, qhi):
    S, Qw = tile.n_sites, qhi - qlo + 1
    B = HDR + 8 * Qw
    out = np.zeros((S, B), np.uint32)
    site = np.repeat(np.arange(S), np.diff(tile.entry_off.astype(np.int64)))
    code = tile.code.astype(np.int64)
    base, strand, indel = code & 7, (code >> 3) & 1, (code >> 4) & 1
    q = tile.qual.astype(np.int64)
    ok = (indel == 0) & (base < 4)
    if (ok & ((q < qlo) | (q > qhi))).any():
        raise ValueError("qual outside the exchange window")
    np.add.at(out, (site[ok], HDR + ((base[ok] * Qw + (q[ok] - qlo)) * 2 + strand[ok])), 1)
    ind = (indel == 1) & (base < 4)
    np.add.at(out, (site[ind], (strand[ind] << 2) | base[ind]), 1)
    out[:, 8] = np.diff(tile.entry_off.astype(np.int64)).astype(np.uint32)
    return out


def vcf_hist(tile, site, ref, altmask):
    """[6][256]: (mapq, qual, rpr) x (ref, alt) over the non-indel entries of one site (src/BaseType.cpp:174-191)."""
    e0, e1 = int(tile.entry_off[site]), int(tile.entry_off[site + 1])
    code = tile.code[e0:e1].astype(np.int64)
    base, indel = code & 7, (code >> 4) & 1
    H = np.zeros((6, 256), np.uint32)
    for grp, sel in ((0, (indel == 0) & (base == ref)), (1, (indel == 0) & (base < 4) & (((altmask >> np.minimum(base, 3)) & 1) == 1) & (base != ref))):
        for f, arr in enumerate((tile.mapq, tile.qual, tile.rpr)):
            np.add.at(H[f * 2 + grp], arr[e0:e1][sel].astype(np.int64), 1)
    return H


def expand(dense, ref_base, n_samples, qlo, qhi):
    """A per-read tile whose (base, qual, strand) / indel statistics are exactly `dense` (mapq = rpr = 0)."""
    S, Qw = dense.shape[0], qhi - qlo + 1
    codes, quals, counts = [], [], np.zeros(S, np.int64)
    bins = np.arange(8 * Qw)
    b_of, q_of, s_of = bins // (2 * Qw), qlo + (bins % (2 * Qw)) // 2, bins % 2
    cls_code = make_code(b_of, s_of, 0)
    ind_code = make_code(np.arange(8) & 3, np.arange(8) >> 2, 1)
    for s in range(S):
        c = dense[s, HDR:].astype(np.int64)
        ic = dense[s, :8].astype(np.int64)
        codes.append(np.concatenate([np.repeat(cls_code, c), np.repeat(ind_code, ic)]))
        quals.append(np.concatenate([np.repeat(q_of, c), np.full(int(ic.sum()), qlo)]).astype(np.uint8))
        counts[s] = c.sum() + ic.sum()
        assert counts[s] == dense[s, 8]
    off = np.zeros(S + 1, np.int64)
    np.cumsum(counts, out=off[1:])
    code = np.concatenate(codes).astype(np.uint8) if S else np.zeros(0, np.uint8)
    qual = np.concatenate(quals) if S else np.zeros(0, np.uint8)
    z = np.zeros(code.size, np.uint8)
    return PileupTile(n_samples, np.ascontiguousarray(ref_base), off.astype(np.uint32), code, qual, z, z.copy(),
                      np.zeros(code.size, np.uint32))
