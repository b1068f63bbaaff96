"""Record-level comparison between two bvc_site_result arrays (tests only).

Integer / index fields: bit-exact.  FP64 fields: |a-b| <= RTOL*max(|a|,|b|) with RTOL = 1e-9 (the
tolerance BASELINE.json's north_star states), NaN == NaN, and an absolute floor of 1e-12 for values
that are differences of O(1..1e4) log-likelihood sums (chi, and var_qual derived from it).
"""
import os
import threading

import numpy as np

from basevarc_b200._abi import EXACT_FIELDS, EXACT_FIELDS_IF_OK, FLOAT_FIELDS, FLOAT_FIELDS_IF_OK

RTOL = 1e-9
PHRED_ATOL = 1e-12


def _close(a, b, rtol=RTOL, atol=0.0):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    both_nan = np.isnan(a) & np.isnan(b)
    with np.errstate(invalid="ignore"):
        ok = np.abs(a - b) <= np.maximum(rtol * np.maximum(np.abs(a), np.abs(b)), atol)
        ok |= (a == b)
    return ok | both_nan


# fields no FP64 threshold can move: counted straight from the entries (src/BaseVarC.cpp:548-608)
HARD_FIELDS = ("n_entries", "depth", "cvg_strand", "cvg_alt_base", "cvg_fs", "cvg_sor")
CHI_ATOL_PER_READ = 5e-11        # chi / var_qual: absolute floor per read of the site
# a kernel that flagged liberally would excuse itself: at most this share of the sites may carry a flag (C2-like windows: ~1.7 %,
# nearly all BVC_FLAG_ARGMIN_TIE -- two error bases with mirror-image evidence make two null models tie exactly) ...
MAX_FLAGGED_FRACTION = 0.05
MAX_EXCUSED_FRACTION = 0.002     # ... and at most this share may actually NEED the excuse (mismatch in a threshold-downstream field)


def compare_records(got, want, allow_flagged=True, max_flagged_fraction=MAX_FLAGGED_FRACTION):
    """Returns (n_mismatching_sites, report lines).  Sites whose GPU record carries a boundary flag
    (BVC_FLAG_*) are the stated threshold-boundary sites: a mismatch there is listed, not counted -- but ONLY in the
    fields downstream of a threshold (ALT set, AF, chi, QUAL, VCF statistics).  The fields counted straight from the
    entries (HARD_FIELDS) must match at flagged sites too, and the share of flagged sites is bounded."""
    assert got.shape == want.shape
    bad = np.zeros(got.shape[0], bool)
    bad_hard = np.zeros(got.shape[0], bool)
    lines = []

    def note(name, mask):
        nonlocal bad, bad_hard
        m = mask if mask.ndim == 1 else mask.any(axis=tuple(range(1, mask.ndim)))
        if m.any():
            idx = np.nonzero(m)[0]
            lines.append("%s differs at %d sites, first %s: got %r want %r" % (
                name, idx.size, idx[:5].tolist(), got[name][idx[0]].tolist(), want[name][idx[0]].tolist()))
            bad |= m
            if name in HARD_FIELDS:
                bad_hard |= m

    ok_site = (want["lrt_ok"] == 1) & (got["lrt_ok"] == 1)
    for f in EXACT_FIELDS:
        note(f, got[f] != want[f])
    for f in EXACT_FIELDS_IF_OK:
        m = got[f] != want[f]
        note(f, m & ok_site[:, None])
    # phred-scaled p-values (-10*log10 p): for p -> 1 the value -> 0 by cancellation, so a 1e-15 relative
    # difference in p shows as ~1e-15 ABSOLUTE in the phred value; PHRED_ATOL covers exactly that.
    for f in FLOAT_FIELDS:
        note(f, ~_close(got[f], want[f], atol=PHRED_ATOL))
    # chi = 2*(lr_alt - lr_null) cancels two sums of magnitude ~n*|log L|: absolute floor scaled by depth
    scale = np.maximum(1.0, want["depth"].sum(1).astype(np.float64)) * CHI_ATOL_PER_READ
    for f in FLOAT_FIELDS_IF_OK:
        a, b = got[f], want[f]
        if f in ("chi", "var_qual"):
            m = ~_close(a, b, atol=0.0) & ~(np.abs(a - b) <= scale)
        else:
            m = ~_close(a, b, atol=PHRED_ATOL if f in ("vcf_fs", "ranksum") else 0.0)
        m = m if m.ndim == 1 else m.any(1)
        note(f, m & ok_site)
    flagged = got["flags"] != 0
    n_sites = max(1, got.shape[0])
    if allow_flagged and flagged.sum() > max(3, max_flagged_fraction * n_sites):
        lines.append("%d of %d sites carry a threshold-boundary flag (more than %.1f %%): flags raised too liberally" % (
            int(flagged.sum()), n_sites, 100 * max_flagged_fraction))
        return int(max(bad.sum(), 1)), lines
    n_flagged_bad = int((bad & flagged & ~bad_hard).sum())
    if allow_flagged and n_flagged_bad:
        lines.append("%d mismatching sites carry a threshold-boundary flag (listed, not counted): %s" % (
            n_flagged_bad, np.nonzero(bad & flagged & ~bad_hard)[0][:20].tolist()))
        # degenerate test tiles (one qual for every read) pass a larger max_flagged_fraction: they may need the excuse more often
        allowed = max(3, MAX_EXCUSED_FRACTION * n_sites, 0.5 * max_flagged_fraction * n_sites if max_flagged_fraction > MAX_FLAGGED_FRACTION else 0)
        if n_flagged_bad > allowed:
            lines.append("too many sites need the threshold-boundary excuse (%d > %d)" % (n_flagged_bad, allowed))
            return int(bad.sum()), lines
        bad &= ~(flagged & ~bad_hard)
    return int(bad.sum()), lines


def oracle_threaded(oracle, tile, params, threads=None):
    """The per-read oracle over every site of `tile` on the host's cores (ctypes drops the GIL; sites are independent)."""
    threads = threads or min(32, os.cpu_count() or 1)
    out = np.zeros(tile.n_sites, got_dtype(oracle))
    bounds = np.linspace(0, tile.n_sites, threads + 1).astype(int)

    def work(a, b):
        if b > a:
            out[a:b] = oracle.call_sites(tile, params, int(a), int(b))

    ths = [threading.Thread(target=work, args=(bounds[i], bounds[i + 1])) for i in range(threads)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    return out


def got_dtype(oracle):
    from basevarc_b200._abi import SITE_RESULT_DTYPE
    return SITE_RESULT_DTYPE
