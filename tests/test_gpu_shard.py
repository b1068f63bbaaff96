"""-m gpu: the SAMPLE-SHARDED path through the C ABI against the per-read oracle on the WHOLE cohort.

On one GPU the W ranks are W contexts of this process with the in-process exchange (bvc_shard_call_group): the same
kernels and phases as the NCCL path, whose collectives are covered by tests/test_gpu_shard_nccl.py (2 GPUs)."""
import numpy as np
import pytest

import basevarc_b200 as B
from basevarc_b200.pileup import synth_tile
from parity import compare_records

pytestmark = pytest.mark.gpu


def _group_call(tile, world, qlo=1, qhi=63, one_byte=False):
    shards = tile.shard_samples(world)
    assert sum(t.n_entries for t in shards) == tile.n_entries
    if one_byte:                                              # every rank sends its tile in the one-byte form (its own dictionary)
        assert all(t.pack8() for t in shards if t.n_entries)
    grp = B.ShardGroup(tile.n_samples, world, qual_lo=qlo, qual_hi=qhi, max_sites=max(tile.n_sites, 1),
                       max_entries=max(max(t.n_entries for t in shards), 1))
    try:
        got = grp.call_sites(shards)
        stats = grp.stats(0)
        assert grp.kernel_launches > 0
    finally:
        grp.close()
    return got, stats


@pytest.mark.parametrize("name,N,S,depth,model,seed,world", [
    ("c4_like_w2", 100000, 48, 0.1, "highpoly", 4, 2),
    ("c4_like_w3", 30000, 50, 0.1, "highpoly", 14, 3),       # S not a multiple of W: ragged last slice
    ("c2_like_w4", 1000, 2001, 0.5, "lowpass", 2, 4),
    ("c5_like_w2", 2000, 900, 5.0, "highpoly", 5, 2),
    ("w1", 500, 700, 0.5, "lowpass", 6, 1),
    ("more_ranks_than_sites", 200, 3, 2.0, "highpoly", 8, 4),
])
def test_sharded_parity(oracle, name, N, S, depth, model, seed, world):
    tile = synth_tile(N, S, depth, model, seed=seed)
    got, stats = _group_call(tile, world)
    want = oracle.call_sites(tile, B.default_params(N))
    nbad, lines = compare_records(got, want)
    assert nbad == 0, "\n".join(lines)
    assert (got["em_fits"] == want["em_fits"]).all()
    assert stats["n_positive"] == int(want["lrt_ok"].sum())
    if S > 10:
        assert want["lrt_ok"].sum() > 0


def test_sharded_equals_unsharded_bitwise():
    """The exchanged statistics are integers, so the owner sees exactly the classes the unsharded K1 builds: every
    field of the record -- FP64 ones included -- is bit-identical to the single-GPU path."""
    tile = synth_tile(3000, 1500, 1.0, "highpoly", seed=31, indel_frac=0.02)
    q = tile.qual.copy(); q[::53] = 1; q[::59] = 63; tile.qual = q
    got, _ = _group_call(tile, 3)
    caller = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
    ref = caller.call_sites(tile)
    caller.close()
    for f in got.dtype.names:
        a, b = got[f], ref[f]
        same = (a == b) | ((a != a) & (b != b)) if a.dtype.kind == "f" else (a == b)
        assert same.all(), f


def test_sharded_qual_window_and_errors():
    tile = synth_tile(400, 300, 1.0, "lowpass", seed=3)
    got, _ = _group_call(tile, 2, qlo=20, qhi=40)             # the generator draws quals in 20..40
    caller = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
    ref = caller.call_sites(tile)
    caller.close()
    assert np.array_equal(got["depth"], ref["depth"]) and np.array_equal(got["af_lrt"], ref["af_lrt"])
    with pytest.raises(B.BvcError, match="outside the exchange window"):
        _group_call(tile, 2, qlo=25, qhi=40)
    with pytest.raises(B.BvcError, match="qual window"):
        _group_call(tile, 2, qlo=0, qhi=40)                   # qual 0 is not served in sharded mode
    c = B.BaseTypeCaller(tile.n_samples, max_sites=8, max_entries=8)
    with pytest.raises(B.BvcError):                           # not initialised for sharding
        c._check(c.lib.bvc_shard_run(c.ctx, None, None, None), "bvc_shard_run")
    c.close()


def test_exchange_window_is_fitted_to_the_tile(monkeypatch):
    """Per tile the ranks agree on the quals their entries carry (a 2-word MAX exchange) and the dense records shrink to that
    range: same records as with the window of bvc_shard_init, fewer bytes through the reduce-scatter."""
    tile = synth_tile(2000, 600, 1.0, "highpoly", seed=41, indel_frac=0.02)      # quals 20..40 inside the window 1..63
    fitted, st_fit = _group_call(tile, 3)
    monkeypatch.setenv("BVC_SHARD_FIXED_WINDOW", "1")
    fixed, st_fix = _group_call(tile, 3)
    monkeypatch.delenv("BVC_SHARD_FIXED_WINDOW")
    assert fitted.tobytes() == fixed.tobytes()
    # dense record: 16 + 8 words per qual of the window -> (16 + 8*21) / (16 + 8*63) = 0.35 of those bytes; the rank-sum histograms
    # of the LRT-positive sites (6 KB each, many on this tile) are exchanged as before
    assert st_fit["exchange_bytes"] < 0.6 * st_fix["exchange_bytes"], (st_fit["exchange_bytes"], st_fix["exchange_bytes"])
    # a rank without a single counted entry, and a tile without any: the window collapses, nothing breaks
    one = synth_tile(3, 40, 0.4, "lowpass", seed=42)
    got, _ = _group_call(one, 3)
    caller = B.BaseTypeCaller(one.n_samples, max_sites=one.n_sites, max_entries=max(one.n_entries, 1))
    ref = caller.call_sites(one)
    caller.close()
    assert got.tobytes() == ref.tobytes()


@pytest.mark.parametrize("world,S", [(3, 1501), (2, 7), (4, 1024)])
def test_sharded_one_byte_form_bitwise(monkeypatch, world, S):
    """Tiles that arrive in the one-byte form run phase A through site_hist8_kernel's streaming (four sites per warp, dense record
    out): bit-identical to the two-array path and to the unsharded GPU path; indel entries (stale base 0..3 and N), ragged sites."""
    tile = synth_tile(2400, S, 0.9, "highpoly", seed=51 + world, indel_frac=0.05)
    c = tile.code.copy()
    ind = np.nonzero((c & 16) != 0)[0]
    c[ind[::7]] = (c[ind[::7]] & 0x18) | 4                     # some indel entries whose stale base is N (qi 30)
    tile.code = c
    monkeypatch.setenv("BVC_SHARD_HIST8", "1")
    got, _ = _group_call(tile, world, one_byte=True)
    monkeypatch.setenv("BVC_SHARD_HIST_ARRAYS", "1")
    arrays, _ = _group_call(tile, world, one_byte=True)
    monkeypatch.delenv("BVC_SHARD_HIST_ARRAYS")
    assert got.tobytes() == arrays.tobytes()
    caller = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
    ref = caller.call_sites(tile)
    caller.close()
    assert got.tobytes() == ref.tobytes()
    with pytest.raises(B.BvcError, match="outside the exchange window"):
        _group_call(tile, world, qlo=25, qhi=40, one_byte=True)
