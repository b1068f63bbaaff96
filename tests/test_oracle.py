"""CPU: the oracle (oracle/basetype_oracle.c) against (1) the committed golden vectors, generated from the
unmodified reference by tests/golden/make_golden.py, and (2) the reference objects / binary themselves
whenever oracle/_ref holds them (build container)."""
import json
import os

import numpy as np
import pytest

import basevarc_b200 as B
import refrun

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def unhex(s):
    return float("nan") if s == "nan" else float.fromhex(s)


def same(a, b):
    return a == b or (a != a and b != b)


@pytest.fixture(scope="module")
def kats():
    with open(os.path.join(GOLD, "leaf_kats.json")) as f:
        return json.load(f)


def test_fisher_golden(oracle, kats):
    P = oracle.port()
    for t, want in kats["fisher"]:
        assert same(P.bvc_oracle_fisher(*t), unhex(want)), t


def test_chisf_normsf_golden(oracle, kats):
    P = oracle.port()
    for x, want in kats["chisf"]:
        assert same(P.bvc_oracle_chisf(unhex(x), 1.0), unhex(want)), x
    for x, want in kats["normsf"]:
        assert same(P.bvc_oracle_normsf(unhex(x)), unhex(want)), x


def test_ranksum_golden(oracle, kats):
    for x, y, want in kats["ranksum"]:
        assert same(oracle.port_ranksum(x, y), unhex(want)), (x, y)


def test_lrt_golden(oracle, kats):
    """BaseType::LRT known answers, bit-exact: emitted flag, ALT set, depths, AF doubles, QUAL double."""
    for c in kats["lrt"]:
        bases = [int(ch) for ch in c["bases"]]
        r = oracle.port_lrt(bases, c["quals"], c["ref"], 0.001, c["comb"])
        assert (r.ok, r.n_alt, list(r.alt_bases)[: r.n_alt], list(r.depth)) == (c["ok"], c["n_alt"], c["alt"], c["depth"])
        assert all(same(a, unhex(w)) for a, w in zip(r.af_lrt, c["af"]))
        assert same(r.var_qual, unhex(c["var_qual"]))


def test_survey_known_answers(oracle):
    """The probe values recorded in SURVEY.md section 8(c)."""
    P = oracle.port()
    assert P.bvc_oracle_fisher(7, 2, 0, 5) == 16.782147827453979
    assert P.bvc_oracle_fisher(149, 154, 39, 45) == 1.4763208238511336
    assert P.bvc_oracle_chisf(24.0, 1.0) == 9.633570086430948e-07
    assert P.bvc_oracle_chisf(0.5, 1.0) == 0.47950012218695282
    assert oracle.port_ranksum([30, 30, 31, 35], [20, 30, 36]) == 1.4045726447967399
    r = oracle.port_lrt([0] * 30 + [2] * 10 + [1], [30] * 40 + [20], 0, 0.001)
    assert (r.ok, list(r.alt_bases)[:1]) == (1, [2])
    assert r.af_lrt[2] == 0.24983311379016532 and r.var_qual == 261.36280569545579
    assert oracle.port_lrt([3] * 20, [30] * 20, 3, 0.001).ok == 0
    r = oracle.port_lrt([1] * 12, [30] * 12, 3, 0.001)
    assert (r.ok, r.var_qual, r.af_lrt[1]) == (1, 5000.0, 1.0)
    r = oracle.port_lrt([1], [30], 3, 0.001)
    assert (r.ok, r.var_qual) == (1, 0.0)


@pytest.mark.parametrize("name", ["synth_ref_lowpass", "synth_ref_highpoly"])
def test_tile_oracle_vs_reference_binary_golden(oracle, name):
    """Whole per-site path (bt_f + WriteVcf numerics) against the text the reference BINARY printed."""
    tile, cvg, vcf = refrun.load_golden_tile(name)
    rec = oracle.call_sites(tile, B.default_params(tile.n_samples))
    bad = refrun.compare_records_to_text(tile, rec, cvg, vcf)
    assert not bad, "\n".join(bad[:5])


def test_live_reference_objects(oracle):
    """Bit-identity with the unmodified reference objects on fresh random inputs (build container only)."""
    R = oracle.ref()
    if R is None:
        pytest.skip("oracle/_ref/libref_oracle.so not built (no /root/reference here)")
    P = oracle.port()
    rng = np.random.default_rng(77)
    for _ in range(3000):
        t = [int(v) for v in rng.integers(0, rng.choice([5, 50, 500, 5000]), 4)]
        assert same(P.bvc_oracle_fisher(*t), R.ref_fisher(*t)), t
    for x in rng.uniform(0, 100, 2000):
        assert same(P.bvc_oracle_chisf(x, 1.0), R.ref_chisf(x, 1.0))
    for i in range(600):
        n = int(rng.choice([1, 3, 12, 40, 400, 3000]))
        ref = int(rng.integers(0, 4))
        bases = (rng.choice(4, n, p=rng.dirichlet([5, 0.3, 0.3, 0.1])) + ref) % 4
        quals = rng.integers(0 if i % 7 == 0 else 1, 42, n)
        a, b = oracle.port_lrt(bases, quals, ref, 0.001), oracle.ref_lrt(bases, quals, ref, 0.001)
        assert (a.ok, a.n_alt, list(a.alt_bases)[: a.n_alt], list(a.depth)) == (b.ok, b.n_alt, list(b.alt_bases)[: b.n_alt], list(b.depth))
        assert all(same(x, y) for x, y in zip(a.af_lrt, b.af_lrt)) and same(a.var_qual, b.var_qual)


def test_live_reference_binary(oracle, tmp_path):
    """End to end through the reference's own text boundary (batch files -> `basetype --rerun`)."""
    if not refrun.have_ref_bin():
        pytest.skip("oracle/_ref/BaseVarC_ref not built (no /root/reference here)")
    tile = B.synth_tile(70, 1200, 1.0, "highpoly", seed=31, first_pos=2001, indel_frac=0.02)
    cvg, vcf = refrun.run_reference(tile, str(tmp_path), threads=4, batch=9)
    rec = oracle.call_sites(tile, B.default_params(70))
    bad = refrun.compare_records_to_text(tile, rec, cvg, vcf)
    assert not bad, "\n".join(bad[:5])
