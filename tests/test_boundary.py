"""Threshold-boundary sites (north_star: "identical PASS/QUAL calls except at stated threshold-boundary sites").
A site is put EXACTLY on a hard FP64 threshold of the reference -- depth/total == min_af by construction, lrt_threshold
set to the site's own chi, em_epsilon set to a delta of the site's own EM -- and the GPU record must carry the matching
BVC_FLAG_* bit there; on ordinary sites every flag must be justified by the oracle's doubles (the "and only then")."""
import math

import numpy as np
import pytest

import basevarc_b200 as B
from basevarc_b200._abi import FLAG_ARGMIN_TIE, FLAG_EM_DELTA, FLAG_LRT_CHI, FLAG_MIN_AF, FLAG_QUAL60, make_code
from basevarc_b200.pileup import PileupTile, synth_tile
from em_ref import em_deltas
from parity import compare_records, oracle_threaded


def _site_tile(n_samples, sites):
    """sites: list of (ref, bases, quals) -> a PileupTile (forward strand, mapq 40, rpr 10, sample = running index)."""
    ref = np.array([s[0] for s in sites], np.uint8)
    off = np.concatenate([[0], np.cumsum([len(s[1]) for s in sites])]).astype(np.uint32)
    base = np.concatenate([np.asarray(s[1], np.uint8) for s in sites])
    qual = np.concatenate([np.asarray(s[2], np.uint8) for s in sites])
    R = base.size
    sample = np.concatenate([np.arange(len(s[1]), dtype=np.uint32) for s in sites])
    return PileupTile(n_samples, ref, off, make_code(base, (np.arange(R) & 1).astype(np.uint8), np.zeros(R, np.uint8)),
                      qual, np.full(R, 40, np.uint8), np.full(R, 10, np.uint8), sample)


def _em_site(seed=0, n=60):
    """Low base qualities: the full-model EM needs ~10 passes, so several deltas lie between 0.001 and the 0.0019 bvc_create allows."""
    rng = np.random.default_rng(seed)
    bases = np.where(rng.random(n) < 0.3, 2, 0).astype(int)
    quals = rng.integers(3, 12, n).astype(int)
    return bases.tolist(), quals.tolist()


def _boundary_delta(deltas):
    """index of a delta the default epsilon does not stop at and bvc_create accepts as an epsilon (<= 0.0019)"""
    ks = [k for k, d in enumerate(deltas[:-1]) if 0.001 < d <= 0.0019]
    return ks[0] if ks else None


def test_python_em_restatement_is_the_oracles_em(oracle):
    """tests/em_ref.py against the C oracle at the bit level: with em_epsilon EXACTLY a delta of the fit the oracle must run one
    more pass (delta < eps is false), with the next double above it the oracle must stop there."""
    n_checked = 0
    for seed in range(8):
        bases, quals = _em_site(seed)
        tile = _site_tile(1000, [(0, bases, quals)])
        deltas = em_deltas(bases, quals, [0, 2])
        assert len(deltas) >= 3 and deltas[-1] < 0.001
        prm = B.default_params(1000)
        for k in range(len(deltas) - 1):
            prm.em_epsilon = deltas[k]
            on = int(oracle.call_sites(tile, prm)["em_passes"][0])
            prm.em_epsilon = math.nextafter(deltas[k], math.inf)
            above = int(oracle.call_sites(tile, prm)["em_passes"][0])
            # full model: 1 + (k + 1) passes when it stops at delta k; the singles add 2 passes each whatever epsilon is
            assert above == (k + 2) + 4 and on == above + 1
            n_checked += 1
    assert n_checked > 40


@pytest.mark.gpu
def test_min_af_boundary_is_flagged(oracle):
    """depth_b / total == min_af exactly (1 read of 1000 at min_af = 0.001; src/BaseType.cpp:79): BVC_FLAG_MIN_AF there, and not
    at the neighbouring counts."""
    N = 1000
    sites = []
    for n_alt in (1, 2, 0):
        sites.append((0, [0] * (1000 - n_alt) + [3] * n_alt, [30] * 1000))
    sites.append((0, [0] * 998 + [3], [30] * 999))            # 1 / 999 > min_af: an ordinary candidate
    tile = _site_tile(N, sites)
    caller = B.BaseTypeCaller(N, max_sites=8, max_entries=tile.n_entries)
    got = caller.call_sites(tile).copy()
    caller.close()
    want = oracle.call_sites(tile, B.default_params(N))
    assert [bool(f & FLAG_MIN_AF) for f in got["flags"]] == [True, False, False, False]
    nbad, lines = compare_records(got, want, allow_flagged=False)
    assert nbad == 0, "\n".join(lines)                        # the values themselves agree: the decision was >= on equal doubles


@pytest.mark.gpu
def test_lrt_threshold_boundary_is_flagged(oracle):
    """lrt_threshold := the chi the oracle computes at a chosen site (src/BaseType.cpp:102 `chi_sqrt_t < LRT_THRESHOLD` is then
    decided on equal doubles): that site must carry BVC_FLAG_LRT_CHI, and sites whose chi is not within 1e-6 of it must not."""
    tile = synth_tile(300, 4000, 1.0, "highpoly", seed=17, indel_frac=0.01)
    prm = B.default_params(300)
    base = oracle.call_sites(tile, prm)
    cand = np.nonzero((base["em_fits"] >= 3) & (base["chi"] > 1.0) & (base["chi"] < 200.0))[0]
    assert cand.size > 50
    n_hit = 0
    for site in cand[:6]:
        thr = float(base["chi"][site])
        prm.lrt_threshold = thr
        want = oracle.call_sites(tile, prm)
        caller = B.BaseTypeCaller(300, max_sites=tile.n_sites, max_entries=tile.n_entries, lrt_threshold=thr)
        got = caller.call_sites(tile).copy()
        caller.close()
        on_thr = np.abs(want["chi"] - thr) <= 1e-9 * thr
        if on_thr[site]:                                      # (a changed threshold can end the ladder earlier at this site)
            n_hit += 1
            assert got["flags"][site] & FLAG_LRT_CHI
        far = np.abs(got["chi"] - thr) > 1e-6 * thr
        assert not (got["flags"][far & (got["em_fits"] <= 3)] & FLAG_LRT_CHI).any()     # one ladder step: its chi is the record's
        nbad, lines = compare_records(got, want)
        assert nbad == 0, "\n".join(lines)
    assert n_hit >= 3


@pytest.mark.gpu
def test_em_epsilon_boundary_is_flagged(oracle):
    """em_epsilon := a delta the reference's EM meets at this site (src/Algorithm.cpp:124 `delta < epsilon` on equal doubles):
    BVC_FLAG_EM_DELTA at the site, the records still within tolerance or listed; a neighbouring epsilon raises no flag."""
    seed = next(s for s in range(50) if _boundary_delta(em_deltas(*_em_site(s), [0, 2])) is not None)
    bases, quals = _em_site(seed)
    others = [_em_site(seed=100 + s, n=50 + s) for s in range(1, 40)]
    tile = _site_tile(1000, [(0, bases, quals)] + [(0, b, q) for b, q in others])
    deltas = em_deltas(bases, quals, [0, 2])
    k = _boundary_delta(deltas)
    for eps, expect in ((deltas[k], True), (deltas[k] * 1.003, False)):
        prm = B.default_params(1000)
        prm.em_epsilon = eps
        want = oracle.call_sites(tile, prm)
        caller = B.BaseTypeCaller(1000, max_sites=tile.n_sites, max_entries=tile.n_entries, em_epsilon=eps)
        got = caller.call_sites(tile).copy()
        caller.close()
        assert bool(got["flags"][0] & FLAG_EM_DELTA) == expect
        nbad, lines = compare_records(got, want, max_flagged_fraction=0.2)
        assert nbad == 0, "\n".join(lines)
        if not expect:
            assert (got["em_passes"] == want["em_passes"]).all()


@pytest.mark.gpu
def test_flags_only_at_boundaries_on_ordinary_sites(oracle):
    """>= 100,000 ordinary sites: every BVC_FLAG_LRT_CHI / QUAL60 / MIN_AF bit is justified by the ORACLE's doubles (the value
    within 1e-6 of its threshold), and EM_DELTA / ARGMIN_TIE (not visible in a record) stay rare."""
    tile = B.synth_config("C2", n_sites=110000, seed=99)
    prm = B.default_params(tile.n_samples)
    caller = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
    got = caller.call_sites(tile).copy()
    caller.close()
    want = oracle_threaded(oracle, tile, prm)
    fl = got["flags"]
    chi_near = np.abs(want["chi"] - prm.lrt_threshold) <= 1e-6 * prm.lrt_threshold
    assert not ((fl & FLAG_LRT_CHI) != 0)[~chi_near & (want["em_fits"] <= 3)].any()
    q_near = np.abs(want["var_qual"] - 60.0) <= 6e-5
    assert not (((fl & FLAG_QUAL60) != 0) & ~q_near).any()
    tot = np.maximum(want["depth"].sum(1), 1).astype(np.float64)
    fr = want["depth"] / tot[:, None]
    af_near = (np.abs(fr - prm.min_af) <= 1e-6 * prm.min_af).any(1)
    assert not (((fl & FLAG_MIN_AF) != 0) & ~af_near).any()
    # EM_DELTA / ARGMIN_TIE are not visible in a record.  Ties are common by construction (two error bases with mirror-image
    # evidence give two null models the same chi: ~1.7 % of C2's sites) but a tied site must then have >= 3 candidates, i.e. >= 4 fits
    assert ((fl & FLAG_EM_DELTA) != 0).sum() <= 30
    tie = (fl & FLAG_ARGMIN_TIE) != 0
    assert (want["em_fits"][tie] >= 4).all() and tie.mean() < 0.04
    nbad, lines = compare_records(got, want)
    assert nbad == 0, "\n".join(lines)
