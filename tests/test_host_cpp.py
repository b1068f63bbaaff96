"""CPU: the C++ host side of the --rerun drop-in (libbvc_host.so) -- symbols, tokeniser/packer against the
tile the batch files were written from, CVG/VCF text against what the UNMODIFIED reference binary printed
(tests/golden), and the command failing loudly without a GPU (no CPU calling path)."""
import os
import re
import subprocess

import numpy as np
import pytest

import basevarc_b200 as B
from basevarc_b200 import batchfile, host

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_host_exports_match_header():
    lib = host.load_host_library()
    src = open(os.path.join(ROOT, "include", "bvc_host.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = sorted(set(re.findall(r"\b(bvc_[a-z0-9_]+)\s*\(", src)))
    assert set(names) == set(host.HOST_EXPORTS)
    for n in names:
        assert getattr(lib, n) is not None


def _site_lines(tile, batch):
    """Per site, the batch-file lines (bytes) the reference's load phase would have written."""
    toks = batchfile.tile_tokens(tile)
    off = tile.entry_off.astype(np.int64)
    N = tile.n_samples
    for s in range(tile.n_sites):
        row = np.full(N, ".", dtype=object)
        row[tile.sample[off[s]:off[s + 1]]] = toks[off[s]:off[s + 1]]
        yield [(" ".join(row[j:j + batch]) + " ").encode() for j in range(0, N, batch)]


def _pack(tile, batch=10):
    pk = host.Packer(tile.n_samples)
    for s, lines in enumerate(_site_lines(tile, batch)):
        pk.add_site(int(tile.pos[s]), int(tile.ref_base[s]), lines)
    return pk


@pytest.mark.parametrize("N,S,depth,model,indel", [(23, 300, 1.0, "highpoly", 0.1), (120, 200, 0.6, "lowpass", 0.01)])
def test_packer_roundtrip(N, S, depth, model, indel):
    tile = B.synth_tile(N, S, depth, model, seed=7, indel_frac=indel, first_pos=100)
    tile.indel_text = {int(e): ("+AC" if e % 2 else "-T") for e in np.nonzero((tile.code >> 4) & 1)[0]}
    pk = _pack(tile, batch=7)
    got = pk.to_tile()
    for k in ("ref_base", "entry_off", "code", "qual", "mapq", "rpr", "sample", "pos"):
        assert np.array_equal(getattr(got, k), getattr(tile, k)), k
    assert got.indel_text == tile.indel_text


def test_packer_token_grammar_quirks():
    """src/BaseVarC.cpp:410-438: N bases skipped with the sample index advancing, indel tokens reuse the stale
    numeric fields (even across sites and from a skipped N base), runs of spaces, short tokens."""
    pk = host.Packer(8)
    n = pk.add_site(10, 0, [b"0,30,35,7,1 . 4,20,11,3,0 +AC  2,60,40,99,1 ", b"-T N . "])
    assert n == 5                       # a bare "N" token takes the indel branch of the reference (str[0] != '.')
    t = pk.to_tile()
    assert t.sample.tolist() == [0, 3, 4, 5, 6]
    # entry 1: "+AC" after the skipped N base "4,20,11,3,0": stale base 4, strand 0, qual 11
    assert (t.code[1] & 7, (t.code[1] >> 3) & 1, (t.code[1] >> 4) & 1, t.qual[1], t.mapq[1], t.rpr[1]) == (4, 0, 1, 11, 20, 3)
    assert (t.code[3] & 7, (t.code[3] >> 4) & 1, t.qual[3]) == (2, 1, 40)       # "-T": stale fields of "2,60,40,99,1"
    # next site starts with an indel token: the state survives the site boundary; "1,5" updates only base and mapq
    assert pk.add_site(11, 1, [b"+G 1,5 "]) == 2
    t = pk.to_tile()
    assert t.entry_off.tolist() == [0, 5, 7]
    assert (t.code[5] & 7, t.qual[5], t.mapq[5]) == (2, 40, 60)
    assert (t.code[6] & 7, t.qual[6], t.mapq[6], t.rpr[6], (t.code[6] >> 3) & 1) == (1, 40, 5, 99, 1)
    assert t.indel_text == {1: "+AC", 3: "-T", 4: "N", 5: "+G"}
    with pytest.raises(B.BvcError):
        pk.add_site(12, 0, [b"0,30,200,1,1 "])                                   # qual > 127 is rejected loudly


def _unpack8(tile):
    """(code, qual) from the compact wire form, the way the device expands it (basevarc_b200/csrc/bvc_kernels.cu expand8_kernel)."""
    p = tile.packed8
    qi, strand, base = (p >> 1) & 31, p & 1, p >> 6
    code = np.where(qi < 30, base | (strand << 3), np.where(qi == 31, base | (strand << 3) | 0x10, 4 | (strand << 3) | 0x10)).astype(np.uint8)
    qual = tile.qual_dict[np.where(qi < 30, qi, 0)]
    return code, qual


@pytest.mark.parametrize("indel", [0.0, 0.05])
def test_packed8_wire_form(indel):
    """include/bvc_gpu.h BVC_PACK8: the packer's compact form and pileup.pack8() expand to the same (code, qual) for every
    counted entry and the same code for indel entries (whose stale qual the calling core never reads)."""
    tile = B.synth_tile(64, 700, 1.2, "highpoly", seed=5, indel_frac=indel, first_pos=100)
    tile.indel_text = {int(e): "+A" for e in np.nonzero((tile.code >> 4) & 1)[0]}
    got = _pack(tile, batch=9).to_tile()
    assert got.packed8 is not None and 1 <= got.qual_dict.size <= 30
    counted = ((got.code >> 4) & 1) == 0
    for t in (got, tile):
        if t is tile:
            assert t.pack8()
        code, qual = _unpack8(t)
        assert np.array_equal(code, t.code)
        assert np.array_equal(qual[counted], t.qual[counted])
    # more than 30 distinct quals: neither producer offers the compact form
    wide = B.synth_tile(64, 300, 1.2, "highpoly", seed=6)
    wide.qual = (np.arange(wide.qual.size) % 40 + 2).astype(np.uint8)
    assert not wide.pack8() and wide.packed8 is None
    assert _pack(wide, batch=9).to_tile().packed8 is None


def _ref_tokenise(lines, state):
    """Plain restatement of src/BaseVarC.cpp:410-438 for one site: strtok on " ", then strtok on "," + atoi per field."""
    import re
    out, sample = [], 0
    for line in lines:
        for tok in [t for t in line.decode().split(" ") if t]:
            if tok[0] not in "+-N.":
                fields = [f for f in tok.split(",") if f][:5]
                for i, f in enumerate(fields):
                    m = re.match(r"\s*([+-]?)(\d*)", f)
                    v = int(m.group(2) or 0) * (-1 if m.group(1) == "-" else 1)
                    state[i] = v & (7, 255, 255, 255, 1)[i]
                if state[0] != 4:
                    out.append((sample, state[0], state[4], 0, state[2], state[1], state[3]))
            elif tok[0] != ".":
                out.append((sample, state[0], state[4], 1, state[2], state[1], state[3]))
            sample += 1
    return out


def test_packer_fast_paths_equal_the_general_grammar():
    """The tokeniser's fast paths (". " and canonical five-field tokens) against the general strtok / atoi grammar on
    random lines that mix canonical tokens with odd ones (short tokens, empty fields, leading zeros, 4-digit numbers,
    trailing junk, runs of spaces), long enough for the 16-byte look-ahead to be in play."""
    rng = np.random.default_rng(11)
    odd = [b"1,5", b"2,,30,,20,7,1", b"03,040,035,012,01", b"0,1000,30,5,1", b"3,40,35,12,1x", b"2,60,33,1,0,9", b"+ACG", b"-T",
           b"N", b".x", b"0,300,99,256,3", b"1, 20,30,4,1"[:4], b"4,20,11,3,0"]
    pk = host.Packer(4096)
    state = [0, 0, 0, 0, 0]
    want_all = []
    for site in range(40):
        lines = []
        for _ in range(3):
            toks = []
            for _ in range(int(rng.integers(1, 60))):
                r = rng.random()
                if r < 0.4:
                    toks.append(b".")
                elif r < 0.85:
                    toks.append(b"%d,%d,%d,%d,%d" % (rng.integers(0, 4), rng.integers(0, 61), rng.integers(1, 94), rng.integers(0, 256), rng.integers(0, 2)))
                else:
                    toks.append(odd[int(rng.integers(0, len(odd)))])
            sep = b"  " if rng.random() < 0.2 else b" "
            lines.append(sep.join(toks) + (b" " if rng.random() < 0.8 else b""))
        want = _ref_tokenise(lines, state)
        n = pk.add_site(100 + site, 0, lines)
        assert n == len(want)
        want_all += want
    t = pk.to_tile()
    got = [(int(t.sample[i]), int(t.code[i] & 7), int(t.code[i] >> 3 & 1), int(t.code[i] >> 4 & 1), int(t.qual[i]), int(t.mapq[i]),
            int(t.rpr[i])) for i in range(t.n_entries)]
    assert got == want_all


def _norm_indels(line):
    f = line.split("\t")
    f[8] = ",".join(sorted(f[8].split(",")))         # the reference's order is a hash-map iteration order
    return f


def _same_number_text(a, b):
    """Printed floats: equal, or one unit in the last printed digit (oracle doubles vs reference doubles are
    bit-identical, so in practice always equal)."""
    if a == b:
        return True
    try:
        fa, fb = float(a), float(b)
    except ValueError:
        return False
    d = len(a.split(".")[1]) if "." in a else 0
    return abs(fa - fb) <= 1.01 * 10.0 ** (-d)


@pytest.mark.parametrize("name", ["synth_ref_lowpass", "synth_ref_highpoly"])
def test_formatter_reproduces_reference_text(oracle, name):
    import refrun
    tile, cvg_ref, vcf_ref = refrun.load_golden_tile(name)
    pk = _pack(tile, batch=16)
    assert np.array_equal(pk.to_tile().code, tile.code)
    rec = oracle.call_sites(tile, B.default_params(tile.n_samples))
    ct = pk.ctile()
    cvg_lines = [l for l in cvg_ref.splitlines() if not l.startswith("#")]
    vcf_lines = [l for l in vcf_ref.splitlines() if not l.startswith("#")]
    chrom = cvg_lines[0].split("\t")[0]
    ci = vi = 0
    for s in range(tile.n_sites):
        if rec["n_entries"][s] == 0:
            continue
        got = host.format_cvg(chrom, int(tile.pos[s]), ct, s, rec[s], pk)
        assert got.endswith("\n")
        assert _norm_indels(got[:-1]) == _norm_indels(cvg_lines[ci]), (got, cvg_lines[ci])    # identical text
        ci += 1
        v = host.format_vcf(chrom, int(tile.pos[s]), ct, s, rec[s], tile.n_samples)
        if rec["lrt_ok"][s]:
            assert v[:-1] == vcf_lines[vi], (v[:300], vcf_lines[vi][:300])                        # identical text
            assert len(v.split("\t")) == 9 + tile.n_samples
            vi += 1
        else:
            assert v == ""
    assert ci == len(cvg_lines) and vi == len(vcf_lines)


GROUP_NAMES = ["AFR", "EAS", "EUR"]          # sorted, as the reference's std::map iterates them


def group_setup(n_samples):
    """sample -> group id of the *_groups fixture (tests/golden/make_golden.py: group_of)."""
    names = ["EAS", "AFR", "EUR", None]
    sg = np.array([GROUP_NAMES.index(names[i % 4]) if names[i % 4] else -1 for i in range(n_samples)], np.int32)
    p = B.default_params(n_samples)
    p.n_groups = 3
    import ctypes as C
    p.sample_group = sg.ctypes.data_as(C.POINTER(C.c_int32))
    return p, sg


def test_formatter_group_path_reproduces_reference_text(oracle):
    """--group: per-group depth columns in the CVG, <grp>_AF keys in the VCF INFO (src/BaseVarC.cpp:617-661)."""
    import refrun
    tile, cvg_ref, vcf_ref = refrun.load_golden_tile("synth_ref_groups")
    pk = _pack(tile, batch=16)
    p, sg = group_setup(tile.n_samples)
    rec, gd, ga = oracle.call_sites(tile, p, want_groups=True)
    ct = pk.ctile()
    cvg_lines = [l for l in cvg_ref.splitlines() if not l.startswith("#")]
    vcf_lines = {int(l.split("\t")[1]): l for l in vcf_ref.splitlines() if not l.startswith("#")}
    assert cvg_ref.splitlines()[2].endswith("\tAFR\tEAS\tEUR")
    ci = 0
    for s in range(tile.n_sites):
        if rec["n_entries"][s] == 0:
            continue
        got = host.format_cvg("chrS", int(tile.pos[s]), ct, s, rec[s], pk, gd[s])
        g, w = _norm_indels(got[:-1]), _norm_indels(cvg_lines[ci])
        assert len(g) == len(w) == 15 and all(_same_number_text(x, y) for x, y in zip(g, w)), (got, cvg_lines[ci])
        ci += 1
        if rec["lrt_ok"][s]:
            v = host.format_vcf("chrS", int(tile.pos[s]), ct, s, rec[s], tile.n_samples, GROUP_NAMES, gd[s], ga[s])
            gf, wf = v[:-1].split("\t"), vcf_lines[int(tile.pos[s])].split("\t")
            gi, wi = gf[7].split(";"), wf[7].split(";")
            assert [x.split("=")[0] for x in gi] == [x.split("=")[0] for x in wi]
            for x, y in zip(gi, wi):
                assert all(_same_number_text(a, b) for a, b in zip(x.split("=")[1].split(","), y.split("=")[1].split(","))), (x, y)
            assert gf[8:] == wf[8:]
    assert ci == len(cvg_lines) and int(rec["lrt_ok"].sum()) == len(vcf_lines)


def test_group_columns_format():
    tile = B.synth_tile(6, 1, 5.0, "lowpass", seed=1, indel_frac=0.0)
    pk = _pack(tile, batch=3)
    rec = np.zeros(1, B.SITE_RESULT_DTYPE)
    rec["depth"][0] = [3, 2, 0, 1]
    rec["lrt_ok"] = 1; rec["n_alt"] = 2; rec["alt_bases"][0] = [1, 3, -1]
    rec["af_lrt"][0] = [0, 0.25, 0, 0.125]; rec["var_qual"] = 61.0
    gd = np.array([[1, 1, 0, 0], [0, 0, 0, 0]], np.int32)
    ga = np.array([[0, 0.5, 0, -1.0], [0, 0, 0, 0]], np.float64)
    cvg = host.format_cvg("c", 5, pk.ctile(), 0, rec[0], pk, gd)
    assert cvg.endswith("\t1:1:0:0\t0:0:0:0\n")
    vcf = host.format_vcf("c", 5, pk.ctile(), 0, rec[0], 6, ["EAS", "AFR"], gd, ga)
    info = vcf.split("\t")[7].split(";")
    assert [k.split("=")[0] for k in info] == sorted(k.split("=")[0] for k in info)
    assert "EAS_AF=0.500000,0" in info and "AFR_AF=0" in info
    assert vcf.split("\t")[4] == "C,T" and vcf.split("\t")[6] == "."


def test_command_fails_loudly_without_gpu(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    tile = B.synth_tile(12, 60, 1.0, "lowpass", seed=3, first_pos=501)
    prefix = str(tmp_path / "out")
    batchfile.write_batch_files(tile, prefix, 2, 5)
    fa = str(tmp_path / "ref.fa")
    batchfile.write_fasta(fa, "chrS", tile.ref_base, 501)
    bl = str(tmp_path / "bam.list")
    open(bl, "w").write("".join("s%d.bam\n" % i for i in range(12)))
    cmd = [host.BIN_PATH, "basetype", "--rerun", "--keep_tmp", "-t", "2", "-b", "5", "-i", bl, "-s",
           batchfile.region_for("chrS", 501, 60), "-r", fa, "-o", prefix]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode != 0 and "no CUDA device" in r.stderr and "no CPU path" in r.stderr
    assert not os.path.exists(prefix + ".vcf.gz")
    # incomplete batch files: told to run the reference's load phase, nothing is computed
    os.remove(prefix + ".tmp.thread.1/batch.0")
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode != 0 and "--load" in r.stderr


def test_fixed_format_matches_the_reference_fmt():
    """bvc_format_fixed against fmt::format("{:.Nf}") of the fmt the reference vendors (tests/golden/make_fmt_kats.py):
    printf's %.Nf for ordinary values, one extra digit for tiny non-zero ones (9.6e-5 -> "0.0001")."""
    import json
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fmt_fixed_kats.json")) as f:
        kats = json.load(f)
    bad, quirk = [], 0
    for h, p, want in kats:
        v = float(h) if "n" in h else float.fromhex(h)
        got = host.format_fixed(v, p)
        if "n" not in h and want != "%.*f" % (p, v):
            quirk += 1
        if got != want:
            bad.append((h, p, got, want))
    assert quirk > 200 and len(kats) > 4000
    assert not bad, bad[:10]
