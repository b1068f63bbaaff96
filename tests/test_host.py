"""CPU: host-side logic -- window partition, batch-file writer (BGZF), synthetic generator invariants."""
import gzip
import os

import numpy as np
import pytest

import basevarc_b200 as B
from basevarc_b200 import batchfile


def test_window_bounds_match_reference_formula():
    # src/BaseVarC.cpp:399-403: window = P%T + P/T, last thread runs to the end
    assert batchfile.window_bounds(10, 1) == [0, 10]
    assert batchfile.window_bounds(10, 3) == [0, 4, 8, 10]
    assert batchfile.window_bounds(78183, 4) == [0, 19548, 39096, 58644, 78183]
    with pytest.raises(ValueError):
        batchfile.window_bounds(5, 4)      # (T-1)*window > P: the reference would read past pv.end()


def test_bgzf_roundtrip_and_eof(tmp_path):
    p = str(tmp_path / "x.gz")
    w = batchfile.BgzfWriter(p)
    data = os.urandom(200000) + b"line\n" * 1000
    w.write(data)
    w.close()
    raw = open(p, "rb").read()
    assert raw[:4] == b"\x1f\x8b\x08\x04" and raw[12:14] == b"BC" and raw.endswith(batchfile._BGZF_EOF)
    assert gzip.open(p, "rb").read() == data


def test_batch_file_grammar(tmp_path):
    tile = B.synth_tile(23, 40, 1.0, "highpoly", seed=5, indel_frac=0.1, first_pos=11)
    paths, names = batchfile.write_batch_files(tile, str(tmp_path / "o"), threads=3, batch=10)
    assert len(paths) == 3 and len(paths[0]) == 3
    bounds = batchfile.window_bounds(40, 3)
    n_tok = 0
    for t in range(3):
        lines = [batchfile.read_bgzf_text(p).split("\n") for p in paths[t]]
        assert lines[0][0] == "".join(n + "\t" for n in names[:10])
        for ls in lines:
            assert len(ls) - 2 == bounds[t + 1] - bounds[t] and ls[-1] == ""
        for row in range(1, bounds[t + 1] - bounds[t] + 1):
            toks = sum((l[row].split(" ")[:-1] for l in lines), [])
            assert len(toks) == 23
            n_tok += sum(tk != "." for tk in toks)
    assert n_tok == tile.n_entries


def test_synth_tile_invariants():
    t = B.synth_tile(200, 3000, 0.5, "lowpass", seed=2)
    assert t.entry_off[0] == 0 and t.entry_off[-1] == t.n_entries and (np.diff(t.entry_off.astype(np.int64)) >= 0).all()
    assert t.qual.min() >= 20 and t.qual.max() <= 40 and t.ref_base.max() <= 3
    for s in range(0, 3000, 97):     # entries of a site are in ascending sample order, one per sample
        smp = t.sample[t.entry_off[s]:t.entry_off[s + 1]]
        assert (np.diff(smp.astype(np.int64)) > 0).all()
    assert abs(t.n_entries / (200 * 3000) - (1 - np.exp(-0.5))) < 0.01
    t2 = B.synth_tile(200, 3000, 0.5, "lowpass", seed=2)
    assert np.array_equal(t.code, t2.code) and np.array_equal(t.qual, t2.qual)
    sl = t.slice_sites(100, 200)
    assert sl.n_sites == 100 and sl.n_entries == int(t.entry_off[200] - t.entry_off[100])
    cat = B.PileupTile.concat([t.slice_sites(0, 100), t.slice_sites(100, 3000)])
    assert np.array_equal(cat.entry_off, t.entry_off) and np.array_equal(cat.code, t.code)
