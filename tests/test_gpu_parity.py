"""-m gpu: the CUDA path, called through the C ABI, against the per-read oracle on the same inputs."""
import numpy as np
import pytest

import basevarc_b200 as B
from basevarc_b200._abi import FLAG_ARGMIN_TIE, FLAG_EM_DELTA
from basevarc_b200.pileup import PileupTile, synth_tile
from parity import compare_records

pytestmark = pytest.mark.gpu


def _run_both(tile, oracle, maf=0.001):
    caller = B.BaseTypeCaller(tile.n_samples, maf=maf, max_sites=max(tile.n_sites, 1), max_entries=max(tile.n_entries, 1))
    got = caller.call_sites(tile)
    assert caller.kernel_launches >= (1 if tile.n_sites else 0)
    caller.close()
    want = oracle.call_sites(tile, B.default_params(tile.n_samples, maf))
    return got, want


@pytest.mark.parametrize("name,N,S,depth,model,seed", [
    ("c2_like", 1000, 3000, 0.5, "lowpass", 2),
    ("c3_like", 10000, 300, 0.3, "lowpass", 3),
    ("c5_like", 2000, 1200, 5.0, "highpoly", 5),
    ("thin", 100, 5000, 0.086, "lowpass", 1),
    ("wide_c4_like", 100000, 40, 0.1, "highpoly", 4),
])
def test_synthetic_parity(oracle, name, N, S, depth, model, seed):
    tile = synth_tile(N, S, depth, model, seed=seed)
    got, want = _run_both(tile, oracle)
    nbad, lines = compare_records(got, want)
    assert nbad == 0, "\n".join(lines)
    # emitted site set + ALT alleles are bit-exact by the checks above; make sure the case is not vacuous
    assert want["lrt_ok"].sum() > 0
    assert (got["em_fits"] == want["em_fits"]).all()
    # a pass count may differ only where the convergence test delta < epsilon (src/Algorithm.cpp:124) was met within 1e-9:
    # every such site is listed and must carry BVC_FLAG_EM_DELTA
    diff = np.nonzero(got["em_passes"] != want["em_passes"])[0]
    unflagged = [int(i) for i in diff if not (int(got["flags"][i]) & (FLAG_EM_DELTA | FLAG_ARGMIN_TIE))]
    assert not unflagged, "em_passes differ without BVC_FLAG_EM_DELTA at sites %s (got %s want %s)" % (
        unflagged[:10], got["em_passes"][unflagged[:10]].tolist(), want["em_passes"][unflagged[:10]].tolist())
    assert diff.size <= max(1, tile.n_sites // 1000), "em_passes differ at %d sites: %s" % (diff.size, diff[:20].tolist())


def test_empty_and_ragged(oracle):
    tile = synth_tile(50, 400, 0.05, "lowpass", seed=9)      # most sites empty, some with 1-3 reads
    assert (np.diff(tile.entry_off.astype(np.int64)) == 0).any()
    got, want = _run_both(tile, oracle)
    nbad, lines = compare_records(got, want)
    assert nbad == 0, "\n".join(lines)
    empty = PileupTile(10, np.zeros(0, np.uint8), np.zeros(1, np.uint32), *(np.zeros(0, np.uint8) for _ in range(4)),
                       np.zeros(0, np.uint32))
    caller = B.BaseTypeCaller(10, max_sites=4, max_entries=4)
    assert caller.call_sites(empty).shape == (0,)
    caller.close()


def test_qual_edge_cases(oracle):
    """qual 0 (1-e == 0 -> the reference's NaN paths), qual 127, all four bases, indel-only sites."""
    rng = np.random.default_rng(11)
    S, N = 600, 64
    tile = synth_tile(N, S, 1.5, "highpoly", seed=11, indel_frac=0.1)
    q = tile.qual.copy()
    q[rng.random(q.size) < 0.05] = 0
    q[rng.random(q.size) < 0.05] = 127
    q[rng.random(q.size) < 0.05] = 1
    tile.qual = q
    got, want = _run_both(tile, oracle)
    nbad, lines = compare_records(got, want)
    assert nbad == 0, "\n".join(lines)


@pytest.mark.parametrize("lo,hi", [(100, 127), (2, 93), (28, 59), (31, 62), (64, 66), (5, 5)])
def test_qual_windows(oracle, lo, hi):
    """K1's emit has a 16-byte form for batches whose quals span <= 32 values from a 4-aligned start (clamped to 96 at the top
    of the scale) and a general form: windows on both sides of each boundary, same records as the oracle."""
    rng = np.random.default_rng(lo * 131 + hi)
    tile = synth_tile(300, 900, 0.8, "highpoly", seed=lo + hi, indel_frac=0.02)
    tile.qual = rng.integers(lo, hi + 1, tile.qual.size).astype(np.uint8)
    got, want = _run_both(tile, oracle)
    # one qual for every read (5, 5) makes the null models of equal-depth bases tie EXACTLY (BVC_FLAG_ARGMIN_TIE, ~3 % of the sites)
    nbad, lines = compare_records(got, want, max_flagged_fraction=0.1 if lo == hi else 0.05)
    assert nbad == 0, "\n".join(lines)
    assert want["lrt_ok"].sum() > 0


@pytest.mark.parametrize("N,S,depth,model,indel", [(1000, 2000, 0.5, "lowpass", 0.005), (200, 1500, 3.0, "highpoly", 0.05)])
def test_packed8_wire_form_same_records(oracle, N, S, depth, model, indel):
    """The compact wire form (include/bvc_gpu.h BVC_PACK8: one byte per entry + qual dictionary, expanded on the device)
    gives byte-identical records to the two-array form, through bvc_enqueue and as a resident tile."""
    tile = synth_tile(N, S, depth, model, seed=21, indel_frac=indel)
    caller = B.BaseTypeCaller(N, max_sites=S, max_entries=tile.n_entries)
    plain = caller.call_sites(tile).copy()
    l0 = caller.kernel_launches
    assert tile.pack8()
    packed = caller.call_sites(tile).copy()
    assert caller.kernel_launches - l0 > 0
    assert packed.tobytes() == plain.tobytes()
    caller.resident_upload(tile)
    caller.resident_run(1, flush_l2=False)
    assert caller.resident_download().tobytes() == plain.tobytes()
    # the dictionary need not be sorted (the host packer fills it in order of first appearance): same records
    perm = np.random.default_rng(5).permutation(tile.qual_dict.size)
    inv = np.argsort(perm)
    qi = (tile.packed8 >> 1) & 31
    shuffled = tile.packed8.copy()
    counted = qi < 30
    shuffled[counted] = (inv[qi[counted]].astype(np.uint8) << 1) | (tile.packed8[counted] & 0xc1)
    t2 = tile.slice_sites(0, tile.n_sites)
    t2.packed8, t2.qual_dict = shuffled, np.ascontiguousarray(tile.qual_dict[perm])
    assert caller.call_sites(t2).tobytes() == plain.tobytes()
    caller.close()
    want = oracle.call_sites(tile, B.default_params(N))
    nbad, lines = compare_records(packed, want)
    assert nbad == 0, "\n".join(lines)


def test_packed8_resident_kernel_equals_expanded_arrays(monkeypatch):
    """site_hist8_kernel (K1 on the one-byte form, the form a packed8 tile stays resident in) against the same tile expanded
    into the (code, qual) arrays behind the copy (BVC_EXPAND8=1 -> site_hist4_kernel): byte-identical records, on shapes that
    exercise the 16-byte-word corrections (tiny and empty sites, one site spanning many words, ragged ends)."""
    rng = np.random.default_rng(77)
    for N, S, depth, model, indel in [(1000, 4001, 0.5, "lowpass", 0.005), (60, 3000, 0.05, "lowpass", 0.2), (2000, 403, 5.0, "highpoly", 0.01),
                                      (40000, 37, 0.3, "lowpass", 0.005)]:
        tile = synth_tile(N, S, depth, model, seed=int(rng.integers(1 << 30)), indel_frac=indel)
        assert tile.pack8()
        monkeypatch.setenv("BVC_EXPAND8", "1")
        c1 = B.BaseTypeCaller(N, max_sites=S, max_entries=tile.n_entries)
        want = c1.call_sites(tile).copy()
        c1.close()
        monkeypatch.delenv("BVC_EXPAND8")
        c2 = B.BaseTypeCaller(N, max_sites=S, max_entries=tile.n_entries)
        got = c2.call_sites(tile).copy()
        c2.close()
        assert got.tobytes() == want.tobytes(), (N, S)


def test_device_rejects_what_the_abi_promises_to_reject():
    """include/bvc_gpu.h BVC_ERR_ARG: qual > 127, a non-indel base > 3, a packed8 dictionary index >= n_qual_dict.  A direct
    C-ABI caller (no host packer in front) gets the error from bvc_wait instead of silently different records."""
    tile = synth_tile(200, 500, 1.0, "lowpass", seed=3)
    caller = B.BaseTypeCaller(200, max_sites=500, max_entries=tile.n_entries)
    ok = caller.call_sites(tile).copy()
    bad = tile.slice_sites(0, tile.n_sites)
    bad.qual = tile.qual.copy(); bad.qual[tile.n_entries // 2] = 200
    with pytest.raises(B.BvcError, match="qual > 127"):
        caller.call_sites(bad)
    bad = tile.slice_sites(0, tile.n_sites)
    bad.code = tile.code.copy(); bad.code[7] = 6            # base 6, not an indel
    with pytest.raises(B.BvcError, match="base > 3"):
        caller.call_sites(bad)
    bad = tile.slice_sites(0, tile.n_sites)
    assert bad.pack8()
    bad.packed8 = bad.packed8.copy(); bad.packed8[11] = (1 << 6) | (29 << 1)      # base 1, dictionary index 29 of a 21-entry dictionary
    with pytest.raises(B.BvcError, match="qual > 127"):
        caller.call_sites(bad)
    assert caller.call_sites(tile).tobytes() == ok.tobytes()                   # the context stays usable
    caller.close()


def test_split_result_equals_full_records():
    """bvc_enqueue_split: the 72-byte core of every site + the full record of the LRT-positive sites only, against the
    176-byte records of bvc_enqueue on the same tile (integer fields identical, FP64 within the parity tolerance -- the two
    calls are separate runs of the EM), two tickets in flight, and the capacity error."""
    tile = synth_tile(500, 6000, 1.0, "highpoly", seed=31, indel_frac=0.01)
    assert tile.pack8()
    caller = B.BaseTypeCaller(500, max_sites=4000, max_entries=tile.n_entries)
    halves = [tile.slice_sites(0, 3500), tile.slice_sites(3500, 6000)]
    full = np.concatenate([caller.call_sites(t).copy() for t in halves])
    cores = [np.zeros(t.n_sites, B.SITE_CORE_DTYPE) for t in halves]
    vcfs = [np.zeros(t.n_sites, B.SITE_RESULT_DTYPE) for t in halves]
    # four enqueues before the first wait: the third and fourth retire the first two tickets (their records must be home)
    scratch = [(np.zeros(t.n_sites, B.SITE_CORE_DTYPE), np.zeros(t.n_sites, B.SITE_RESULT_DTYPE)) for t in halves]
    tks = [caller.enqueue_split(t, c, v) for t, c, v in zip(halves, cores, vcfs)]
    tks2 = [caller.enqueue_split(t, c, v) for t, (c, v) in zip(halves, scratch)]
    ns = [caller.wait_split(tk) for tk in tks]
    assert [caller.wait_split(tk) for tk in tks2] == ns
    off = 0
    for t, c, v, n in zip(halves, cores, vcfs, ns):
        ref = full[off:off + t.n_sites]
        off += t.n_sites
        for f in ("n_entries", "depth", "cvg_strand", "em_fits", "cvg_alt_base", "lrt_ok", "n_alt"):
            assert np.array_equal(c[f], ref[f]), f
        for f in ("cvg_fs", "cvg_sor"):
            assert c[f].tobytes() == ref[f].tobytes(), f
        ok = ref["lrt_ok"] == 1
        assert n == int(ok.sum()) and n > 50
        assert (c["vcf_index"][~ok] == -1).all() and sorted(c["vcf_index"][ok].tolist()) == list(range(n))
        nbad, lines = compare_records(v[c["vcf_index"][ok]], ref[ok])
        assert nbad == 0, "\n".join(lines)
        for f in ("af_lrt", "vcf_strand", "ranksum", "vcf_fs"):
            assert not ref[f][~ok].any(), f                       # the half that does not travel is all zero anyway
    core = np.zeros(halves[0].n_sites, B.SITE_CORE_DTYPE)
    tk = caller.enqueue_split(halves[0], core, np.zeros(3, B.SITE_RESULT_DTYPE))
    with pytest.raises(B.BvcError, match="vcf_cap"):
        caller.wait_split(tk)
    assert np.array_equal(core["depth"], full["depth"][:3500])   # the core part is valid all the same
    caller.close()


def test_group_path_parity(oracle):
    """--group: per-group depths (exact) and the group-level LRT allele frequencies against the oracle."""
    from test_host_cpp import group_setup
    tile = synth_tile(60, 900, 3.0, "highpoly", seed=41, indel_frac=0.02)
    q = tile.qual.copy(); q[::97] = 0; q[::89] = 1; tile.qual = q
    p, sg = group_setup(tile.n_samples)
    caller = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries, sample_group=sg, n_groups=3)
    got, gd, ga = caller.call_sites_with_groups(tile)
    caller.close()
    want, wd, wa = oracle.call_sites(tile, p, want_groups=True)
    nbad, lines = compare_records(got, want)
    assert nbad == 0, "\n".join(lines)
    assert np.array_equal(gd, wd)
    ok = (want["lrt_ok"] == 1) & (got["flags"] == 0)
    assert ok.sum() > 50 and (wa[ok] != 0).any() and (wa[ok] == -1).any()
    a, b = ga[ok], wa[ok]
    close = (np.abs(a - b) <= 1e-9 * np.maximum(np.abs(a), np.abs(b))) | (np.isnan(a) & np.isnan(b))
    assert close.all(), (a[~close][:5], b[~close][:5])
    assert (ga[want["lrt_ok"] == 0] == 0).all()


def test_double_buffered_tickets(oracle):
    tiles = [synth_tile(500, 700, 0.5, "lowpass", seed=20 + i) for i in range(5)]
    caller = B.BaseTypeCaller(500, max_sites=700, max_entries=700 * 500)
    outs = [np.zeros(t.n_sites, B.SITE_RESULT_DTYPE) for t in tiles]
    tks = [caller.enqueue(t, o) for t, o in zip(tiles, outs)]
    for tk in tks:
        caller.wait(tk)
    caller.close()
    for t, o in zip(tiles, outs):
        want = oracle.call_sites(t, B.default_params(500))
        nbad, lines = compare_records(o, want)
        assert nbad == 0, "\n".join(lines)


def test_capacity_and_argument_errors():
    caller = B.BaseTypeCaller(100, max_sites=10, max_entries=100)
    tile = synth_tile(100, 50, 0.5, seed=1)
    with pytest.raises(B.BvcError):
        caller.call_sites(tile)
    caller.close()
    with pytest.raises(B.BvcError):
        B.BaseTypeCaller(100, maf=0.0)


@pytest.mark.parametrize("name", ["synth_ref_lowpass", "synth_ref_highpoly"])
def test_gpu_vs_reference_binary_golden(name):
    """CUDA path against the text the unmodified reference binary printed (tests/golden, no oracle involved)."""
    import refrun
    tile, cvg, vcf = refrun.load_golden_tile(name)
    caller = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
    rec = caller.call_sites(tile)
    caller.close()
    bad = refrun.compare_records_to_text(tile, rec, cvg, vcf)
    assert not bad, "\n".join(bad[:5])


def _records(tile, monkeypatch, **env):
    for k, v in env.items():
        monkeypatch.setenv(k, str(v))
    caller = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
    caller.resident_upload(tile)                      # the device-resident path: the work counters belong to it
    caller.resident_run(1, False)
    rec = caller.resident_download().copy()
    work = caller.work_counters()
    caller.close()
    for k in env:
        monkeypatch.delenv(k)
    return rec, work


@pytest.mark.parametrize("N,S,depth,model,seed", [(10000, 400, 0.3, "lowpass", 21), (2000, 1500, 5.0, "highpoly", 22)])
def test_em_long_bound_passes_identical_to_exact(monkeypatch, N, S, depth, model, seed):
    """em_long_kernel decides a pass by delta_decision's bound and evaluates the exact delta (Algorithm.cpp:103-113) only when
    the bound is ambiguous; its E/M arithmetic is the same either way, so the records must be BYTE-identical to those of the
    all-exact kernel (BVC_LONG_EXACT=1).  Tiles this small run every fit in em_long_kernel (fewer fits than K3 has lanes)."""
    tile = synth_tile(N, S, depth, model, seed=seed)
    q = tile.qual.copy()
    q[np.random.default_rng(seed).random(q.size) < 0.002] = 1          # qual <= 1 classes: those fits stay on the exact path
    tile.qual = q
    exact, w_exact = _records(tile, monkeypatch, BVC_LONG_EXACT=1)
    bound, w_bound = _records(tile, monkeypatch)
    assert w_exact["em_class_passes"] == 0 and w_bound["em_class_passes"] == 0 and w_bound["em_long_class_passes"] > 0
    assert exact.tobytes() == bound.tobytes()
    # a repeated (ambiguous) pass is counted twice: the bound-based kernel may only do slightly MORE class-passes
    assert w_exact["em_long_class_passes"] <= w_bound["em_long_class_passes"] <= 1.01 * w_exact["em_long_class_passes"]


@pytest.mark.parametrize("refill_min", [1, 6, 32])
def test_em_fit_batched_refills(oracle, monkeypatch, refill_min):
    """em_fit_kernel (forced on a small tile with BVC_EM_MIN_FITS=1) with lanes waiting for `refill_min` free lanes before
    the warp refills: same calls, same pass counts, AF / LRT within tolerance against the per-read oracle."""
    tile = synth_tile(1000, 6000, 0.5, "lowpass", seed=23)
    got, work = _records(tile, monkeypatch, BVC_EM_MIN_FITS=1, BVC_REFILL_MIN=refill_min)
    assert work["em_class_passes"] > 0
    want = oracle.call_sites(tile, B.default_params(tile.n_samples, 0.001))
    nbad, lines = compare_records(got, want)
    assert nbad == 0, "\n".join(lines)
    assert (got["em_fits"] == want["em_fits"]).all()
    # a pass count may differ only where the convergence test delta < epsilon (src/Algorithm.cpp:124) was met within 1e-9:
    # every such site is listed and must carry BVC_FLAG_EM_DELTA
    diff = np.nonzero(got["em_passes"] != want["em_passes"])[0]
    unflagged = [int(i) for i in diff if not (int(got["flags"][i]) & (FLAG_EM_DELTA | FLAG_ARGMIN_TIE))]
    assert not unflagged, "em_passes differ without BVC_FLAG_EM_DELTA at sites %s (got %s want %s)" % (
        unflagged[:10], got["em_passes"][unflagged[:10]].tolist(), want["em_passes"][unflagged[:10]].tolist())
    assert diff.size <= max(1, tile.n_sites // 1000), "em_passes differ at %d sites: %s" % (diff.size, diff[:20].tolist())
