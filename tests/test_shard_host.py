"""CPU (-m "not gpu"): host side of the sample-sharded mode -- sample-range sharding of a tile, the dense exchange
layout, and the world_size-2 protocol over gloo with the oracle standing in for the owner's kernels."""
import os
import subprocess
import sys

import numpy as np

import basevarc_b200 as B
from basevarc_b200.pileup import synth_tile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_samples_partitions_every_site():
    tile = synth_tile(300, 400, 1.0, "highpoly", seed=12, indel_frac=0.02)
    for world in (1, 2, 3, 7):
        shards = tile.shard_samples(world)
        assert sum(t.n_entries for t in shards) == tile.n_entries
        for r, t in enumerate(shards):
            assert t.n_sites == tile.n_sites and np.array_equal(t.ref_base, tile.ref_base)
            lo, hi = (r * 300) // world, ((r + 1) * 300) // world
            assert ((t.sample >= lo) & (t.sample < hi)).all()
        # per site: the shards' entries, concatenated in rank order, are the site's entries in sample order
        for s in (0, 17, 399):
            cat = np.concatenate([t.code[t.entry_off[s]:t.entry_off[s + 1]] for t in shards])
            assert np.array_equal(cat, tile.code[tile.entry_off[s]:tile.entry_off[s + 1]])


def test_dense_layout_is_additive_and_sufficient(oracle):
    import shard_ref
    from parity import compare_records
    tile = synth_tile(500, 120, 2.0, "highpoly", seed=5, indel_frac=0.03)
    whole = shard_ref.dense_hist(tile, 1, 63)
    parts = sum(shard_ref.dense_hist(t, 1, 63).astype(np.int64) for t in tile.shard_samples(3))
    assert np.array_equal(whole.astype(np.int64), parts)
    assert whole.shape[1] == 16 + 8 * 63 and np.array_equal(whole[:, 8], np.diff(tile.entry_off.astype(np.int64)))
    got = oracle.call_sites(shard_ref.expand(whole, tile.ref_base, 500, 1, 63), B.default_params(500))
    want = oracle.call_sites(tile, B.default_params(500))
    got["ranksum"] = want["ranksum"]
    nbad, lines = compare_records(got, want)
    assert nbad == 0, "\n".join(lines)


def test_world2_gloo_protocol():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29537", os.path.join(ROOT, "tests", "shard_worker.py"), "gloo"]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert "shard_worker ok" in r.stdout
