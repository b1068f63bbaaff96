"""-m gpu: BASELINE.json's C2 at its full size (1,000,000 sites x 1,000 samples, 393 M entries) -- too large for the
per-read oracle as a whole, so: the oracle on a random sample of its sites, and size-independent properties on all of them."""
import numpy as np
import pytest

import basevarc_b200 as B
from parity import compare_records, oracle_threaded
from basevarc_b200._abi import FLAG_ARGMIN_TIE, FLAG_EM_DELTA

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def c2_full():
    tile = B.synth_config("C2", seed=1002)
    caller = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
    caller.resident_upload(tile)
    caller.resident_run(1, flush_l2=False)
    rec = caller.resident_download().copy()
    caller.close()
    return tile, rec


def test_full_c2_sampled_sites_against_the_oracle(oracle, c2_full):
    tile, rec = c2_full
    rng = np.random.default_rng(5)
    # every LRT-positive site of a stretch + random sites of the whole window
    pick = np.unique(np.concatenate([rng.integers(0, tile.n_sites, 1200), np.nonzero(rec["lrt_ok"][:200000] == 1)[0][:300]]))
    parts = [tile.slice_sites(int(s), int(s) + 1) for s in pick]
    off = np.concatenate([[0], np.cumsum([p.n_entries for p in parts])]).astype(np.uint32)
    sub = B.PileupTile(tile.n_samples, np.concatenate([p.ref_base for p in parts]), off,
                       *(np.concatenate([getattr(p, k) for p in parts]) for k in ("code", "qual", "mapq", "rpr", "sample")))
    want = oracle.call_sites(sub, B.default_params(tile.n_samples))
    nbad, lines = compare_records(rec[pick], want)
    assert nbad == 0, "\n".join(lines)
    assert want["lrt_ok"].sum() >= 100


def test_full_c2_invariants(c2_full):
    tile, rec = c2_full
    n = np.diff(tile.entry_off.astype(np.int64))
    indel = np.add.reduceat(((tile.code >> 4) & 1).astype(np.int64), tile.entry_off[:-1].astype(np.int64))
    indel[n == 0] = 0
    assert np.array_equal(rec["n_entries"], n)
    assert np.array_equal(rec["depth"].sum(1), n - indel)                      # every non-indel entry is counted once
    assert (rec["cvg_strand"].sum(1) <= n).all()
    ok = rec["lrt_ok"] == 1
    assert 1000 < ok.sum() < 20000                                             # ~1/200 variant sites, AF ~ Beta(0.3, 2)
    assert (rec["n_alt"][ok] >= 1).all() and (rec["n_alt"][~ok] == 0).all()
    af = rec["af_lrt"]
    assert ((af >= 0) & (af <= 1.0 + 1e-12)).all() and (af[~ok] == 0).all()
    assert (rec["vcf_strand"][ok].sum(1) <= rec["depth"][ok].sum(1)).all()
    assert np.isfinite(rec["cvg_fs"]).all() and (rec["cvg_fs"] >= 0).all()
    assert (rec["em_fits"] >= (n - indel > 0)).all() and (rec["em_passes"] <= 101 * rec["em_fits"]).all()


def test_full_c2_is_invariant_to_tiling_and_wire_form(c2_full):
    """The same window as 7 uneven sub-tiles through bvc_enqueue (compact wire form, two tickets in flight): no state
    crosses sites (src/BaseVarC.cpp:403) and the wire form carries everything the core reads, so every integer field is
    identical and every FP64 field agrees to the parity tolerance.  (Not bitwise: whether the last passes of a fit run in
    em_fit_kernel or in em_long_kernel -- sequential sums against 8-lane butterflies -- depends on the stage's size and on
    when its queue drains, which moves the last bits of chi / af_lrt.)"""
    tile, rec = c2_full
    bounds = [0, 1, 9999, 250000, 250001, 600000, 999999, tile.n_sites]
    caller = B.BaseTypeCaller(tile.n_samples, max_sites=400000, max_entries=int(400000 * 0.45 * tile.n_samples))
    outs, flight = [], []
    for a, b in zip(bounds[:-1], bounds[1:]):
        t = tile.slice_sites(a, b)
        assert t.pack8()
        out = np.zeros(t.n_sites, B.SITE_RESULT_DTYPE)
        flight.append((caller.enqueue(t, out), t, out))
        outs.append(out)
        if len(flight) == 2:
            caller.wait(flight.pop(0)[0])
    for tk, _, _ in flight:
        caller.wait(tk)
    caller.close()
    got = np.concatenate(outs)
    for f in ("n_entries", "depth", "cvg_strand", "vcf_strand", "em_fits", "cvg_alt_base", "lrt_ok", "n_alt", "alt_bases"):
        assert np.array_equal(got[f], rec[f]), f
    assert (got["em_passes"] == rec["em_passes"]).mean() > 0.9999
    nbad, lines = compare_records(got, rec)
    assert nbad == 0, "\n".join(lines)
    for f in ("cvg_fs", "cvg_sor", "vcf_fs", "vcf_sor"):                          # pure functions of the integer tables
        assert got[f].tobytes() == rec[f].tobytes(), f


@pytest.mark.parametrize("workload,n_sites", [("C2", 120000), ("C3", 100000), ("C5", 100000)])
def test_100k_sites_per_config_against_the_threaded_oracle(oracle, workload, n_sites):
    """>= 100,000 consecutive sites of each single-GPU shape BASELINE.json names (C2 1,000 x 0.5x, C3 10,000 x 0.3x, C5 2,000 x 5x
    high-polymorphism), EVERY site against the per-read oracle run on all host cores: integers bit-exact, FP64 within 1e-9, a pass
    count may differ only under BVC_FLAG_EM_DELTA, and the flagged sites are a bounded share (parity.compare_records)."""
    tile = B.synth_config(workload, n_sites=n_sites, seed=4242)
    assert tile.pack8()
    caller = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
    got = caller.call_sites(tile).copy()
    caller.close()
    want = oracle_threaded(oracle, tile, B.default_params(tile.n_samples))
    nbad, lines = compare_records(got, want)
    assert nbad == 0, "\n".join(lines)
    assert (got["em_fits"] == want["em_fits"]).all()
    diff = np.nonzero(got["em_passes"] != want["em_passes"])[0]
    # (or under BVC_FLAG_ARGMIN_TIE: the other branch of a tied ladder step runs different fits)
    unflagged = [int(i) for i in diff if not (int(got["flags"][i]) & (FLAG_EM_DELTA | FLAG_ARGMIN_TIE))]
    assert not unflagged, "em_passes differ without BVC_FLAG_EM_DELTA / ARGMIN_TIE at %s" % unflagged[:10]
    assert diff.size <= n_sites // 2000, "em_passes differ at %d sites" % diff.size
    assert want["lrt_ok"].sum() > 100


def test_c4_shape_sample_sharded_against_the_threaded_oracle(oracle):
    """BASELINE config C4 at its full WIDTH: 100,000 samples (~9,500 entries per site), 3,000 consecutive sites, every site
    against the per-read oracle on the whole cohort -- through the sample-sharded path (3 ranks, one-byte tiles, the per-tile
    exchange window) and through the plain path (the wide K1 form), which must also agree with each other bit for bit."""
    tile = B.synth_config("C4", n_sites=3000, seed=777)
    want = oracle_threaded(oracle, tile, B.default_params(tile.n_samples))
    shards = tile.shard_samples(3)
    assert all(t.pack8() for t in shards)
    grp = B.ShardGroup(tile.n_samples, 3, max_sites=tile.n_sites, max_entries=max(t.n_entries for t in shards))
    try:
        got = grp.call_sites(shards)
    finally:
        grp.close()
    nbad, lines = compare_records(got, want)
    assert nbad == 0, "\n".join(lines)
    assert (got["em_fits"] == want["em_fits"]).all() and want["lrt_ok"].sum() > 10
    caller = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
    plain = caller.call_sites(tile)
    caller.close()
    assert plain.tobytes() == got.tobytes()
