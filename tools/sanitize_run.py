#!/usr/bin/env python
"""What compute-sanitizer wraps (one tool per gpurun call):
    compute-sanitizer --tool memcheck  python tools/sanitize_run.py
    compute-sanitizer --tool racecheck python tools/sanitize_run.py
smoke()'s tile and small C2- / C5- / C3-like tiles through every kernel of the tile path: the one-byte form (site_hist8_kernel),
the two-array form (site_hist4_kernel), the wide form (site_hist_kernel<G>, >= 65,536 samples), the split result, a qual-0 site
(the legacy warp-per-site kernel) and --group; every result is compared with the oracle so a clean sanitizer log belongs to a
CORRECT run."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import basevarc_b200 as B  # noqa: E402
import oracle_py  # noqa: E402
from parity import compare_records  # noqa: E402


def check(name, tile, packed, **kw):
    if packed:
        assert tile.pack8()
    c = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries, **kw)
    got = c.call_sites(tile).copy()
    core, vcf = c.call_sites_split(tile)
    c.resident_upload(tile)
    c.resident_run(1, False)
    res = c.resident_download().copy()
    launches = c.kernel_launches
    c.close()
    want = oracle_py.call_sites(tile, B.default_params(tile.n_samples))
    nbad, lines = compare_records(got, want, max_flagged_fraction=0.1)
    assert nbad == 0, "\n".join(lines)
    assert np.array_equal(res["depth"], got["depth"]) and np.array_equal(core["depth"], got["depth"])
    print("%-28s %6d sites %9d entries  %4d LRT-positive  %d kernel launches: ok" % (name, tile.n_sites, tile.n_entries, int(got["lrt_ok"].sum()), launches), flush=True)


def main():
    check("smoke (C2-like, packed8)", B.synth_tile(1000, 2000, 0.5, "lowpass", seed=7), True)
    check("C2-like, two-array form", B.synth_tile(1000, 1500, 0.5, "lowpass", seed=8, indel_frac=0.01), False)
    check("C5-like, packed8", B.synth_tile(2000, 300, 5.0, "highpoly", seed=9, indel_frac=0.01), True)
    check("C3-like, packed8", B.synth_tile(10000, 120, 0.3, "lowpass", seed=10), True)
    check("thin / ragged, packed8", B.synth_tile(60, 3000, 0.05, "lowpass", seed=11, indel_frac=0.2), True)
    check("wide (70,000 samples)", B.synth_tile(70000, 24, 0.1, "highpoly", seed=12), False)
    t = B.synth_tile(64, 400, 1.5, "highpoly", seed=13, indel_frac=0.1)
    q = t.qual.copy(); q[::17] = 0; q[::19] = 1; t.qual = q
    check("qual 0 / 1 (legacy kernel)", t, False)
    # --group (bvc_groups.cuh)
    t = B.synth_tile(300, 500, 1.0, "highpoly", seed=14)
    sg = (np.arange(300) % 4 - 1).astype(np.int32)            # groups 0..2 and -1 (no group)
    c = B.BaseTypeCaller(300, max_sites=t.n_sites, max_entries=t.n_entries, sample_group=sg, n_groups=3)
    got, gd, ga = c.call_sites_with_groups(t)
    c.close()
    want = oracle_py.call_sites(t, B.default_params(300))
    nbad, lines = compare_records(got, want, max_flagged_fraction=0.1)
    assert nbad == 0 and gd.sum() > 0, "\n".join(lines)
    print("--group, 3 groups: ok", flush=True)
    # sample-sharded mode, in-process exchange (bvc_shard.cuh), 3 ranks, ragged last slice
    t = B.synth_tile(30000, 50, 0.1, "highpoly", seed=15)
    shards = t.shard_samples(3)
    g = B.ShardGroup(t.n_samples, 3, max_sites=t.n_sites, max_entries=max(x.n_entries for x in shards))
    got = g.call_sites(shards)
    g.close()
    want = oracle_py.call_sites(t, B.default_params(t.n_samples))
    nbad, lines = compare_records(got, want, max_flagged_fraction=0.1)
    assert nbad == 0, "\n".join(lines)
    print("sample-sharded x3 (in-process exchange): ok", flush=True)
    # bvc_call_sites_noalloc (static scratch array), two calls in a row; device memory through cudart (bvc_host_alloc is pinned
    # HOST memory the kernels read in place: no torch needed under the sanitizer)
    import ctypes as C
    from basevarc_b200 import caller as bc
    from basevarc_b200._abi import SITE_RESULT_DTYPE
    for seed in (16, 17):
        t = B.synth_tile(200, 800, 0.8, "highpoly", seed=seed, indel_frac=0.02)
        bufs = {}
        for k in ("ref_base", "entry_off", "code", "qual", "mapq", "rpr"):
            a = getattr(t, k)
            bufs[k] = bc.PinnedArray(a.shape, a.dtype)
            bufs[k].array[...] = a
        out = bc.PinnedArray((t.n_sites,), SITE_RESULT_DTYPE)
        st = bc.PinnedArray((1,), np.uint32)
        st.array[0] = 99
        bc.call_sites_noalloc(B.default_params(200), t.n_sites, t.n_entries, *(bufs[k].ptr for k in ("ref_base", "entry_off", "code", "qual", "mapq", "rpr")),
                              out.ptr, st.ptr, None)
        rt = C.CDLL("/usr/local/cuda/lib64/libcudart.so.12")
        assert rt.cudaDeviceSynchronize() == 0
        c = B.BaseTypeCaller(200, max_sites=t.n_sites, max_entries=t.n_entries)
        ref = c.call_sites(t)
        c.close()
        assert int(st.array[0]) == 0
        nbad, lines = compare_records(out.array.copy(), ref, max_flagged_fraction=0.1)
        assert nbad == 0, "\n".join(lines)
    print("bvc_call_sites_noalloc x2: ok", flush=True)
    # the coded wire form (decode_codes_kernel writes the resident bytes 32 at a time: tails of every alignment)
    for n_sites, seed in ((2000, 18), (1777, 19), (4099, 20)):
        t = B.synth_tile(700, n_sites, 0.6, "lowpass", seed=seed, indel_frac=0.01)
        assert t.pack8()
        cod = B.CodedArrays(t)
        assert cod.bits == 6
        c = B.BaseTypeCaller(700, max_sites=t.n_sites, max_entries=t.n_entries)
        core = np.zeros(t.n_sites, B.SITE_CORE_DTYPE)
        vcf = np.zeros(t.n_sites, B.SITE_RESULT_DTYPE)
        c.wait_split(c.enqueue_split_coded(t, cod, core, vcf))
        ref = c.call_sites(t)
        c.close()
        assert np.array_equal(core["depth"], ref["depth"]) and np.array_equal(core["cvg_strand"], ref["cvg_strand"]) and np.array_equal(core["lrt_ok"], ref["lrt_ok"])
    print("coded wire form x3: ok", flush=True)
    print("sanitize_run: all tiles agree with the oracle")


if __name__ == "__main__":
    main()
