"""Record-level comparison between two bvc_site_result arrays (tests only).

Integer / index fields: bit-exact.  FP64 fields: |a-b| <= RTOL*max(|a|,|b|) with RTOL = 1e-9 (the
tolerance BASELINE.json's north_star states), NaN == NaN, and an absolute floor of 1e-12 for values
that are differences of O(1..1e4) log-likelihood sums (chi, and var_qual derived from it).
"""
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
MAX_FLAGGED_FRACTION = 0.02      # a kernel that flags liberally would excuse itself: at most this share of the sites may carry a flag


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
        bad &= ~(flagged & ~bad_hard)
    return int(bad.sum()), lines
