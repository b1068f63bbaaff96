"""C1 = BASELINE.json configs[0]: the reference's own test/ example (100 BAMs, chr17:41197700-41276155).

tests/golden/c1 holds what the UNMODIFIED reference produced here (tests/golden/make_c1.py): the batch files written by its
load phase and its .cvg / .vcf text.  Real data: qual 1..39, mapq filter -q 20, 57 indel tokens, ~8 reads per site.
  * CPU: reference batch files -> host packer (quirk Q1 per window thread) -> oracle -> host formatter == reference text
  * GPU (-m gpu): `BaseVarC_b200 basetype --rerun` with test.sh's arguments on the same files == reference text
"""
import gzip
import os
import shutil
import subprocess

import numpy as np
import pytest

import basevarc_b200 as B
from basevarc_b200 import batchfile, host

C1 = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c1")
T, BATCH, N = 4, 10, 100


def _window():
    with gzip.open(os.path.join(C1, "chr17_window.txt.gz"), "rt") as f:
        length, lo, hi = (int(x) for x in f.readline().split())
        seq = f.readline().strip()
    assert len(seq) == hi - lo
    return length, lo, seq


def _ref_text(ext):
    with gzip.open(os.path.join(C1, "ref.%s.txt.gz" % ext), "rt") as f:
        return f.read()


def _sites():
    """S1: positions S .. E-2 (1-based) whose reference base is A/C/G/T (src/BaseVarUtils.h:71, RefReader.h:22)."""
    length, lo, seq = _window()
    S, E = 41197700, 41276155
    code = {"A": 0, "C": 1, "G": 2, "T": 3}
    out = []
    for p in range(S, E - 1):
        b = seq[p - 1 - lo] if lo <= p - 1 < lo + len(seq) else "N"
        if b in code:
            out.append((p, code[b]))
    return out


def test_c1_packer_oracle_formatter_reproduce_reference_text(oracle):
    from test_host_cpp import _norm_indels
    pv = _sites()
    window = len(pv) % T + len(pv) // T                       # S2: src/BaseVarC.cpp:399-403
    cvg_got, vcf_got = [], []
    n_indel = 0
    for t in range(T):
        mine = pv[t * window:] if t == T - 1 else pv[t * window:(t + 1) * window]
        files = [batchfile.read_bgzf_text(os.path.join(C1, "test.out.tmp.thread.%d" % t, "batch.%d" % j)).encode().split(b"\n")
                 for j in range(N // BATCH)]
        assert all(len(f) >= len(mine) + 1 for f in files)
        pk = host.Packer(N)                                   # one `ai` per window thread (src/BaseVarC.cpp:393)
        for i, (p, rb) in enumerate(mine):
            pk.add_site(p, rb, [f[i + 1] for f in files])     # line 0 = sample names
        tile = pk.to_tile()
        n_indel += len(tile.indel_text)
        rec = oracle.call_sites(tile, B.default_params(N))
        ct = pk.ctile()
        for s in range(tile.n_sites):
            if rec["n_entries"][s] == 0:
                continue
            cvg_got.append(host.format_cvg("chr17", int(tile.pos[s]), ct, s, rec[s], pk)[:-1])
            if rec["lrt_ok"][s]:
                vcf_got.append(host.format_vcf("chr17", int(tile.pos[s]), ct, s, rec[s], N)[:-1])
        pk.close()
    cvg_ref = [l for l in _ref_text("cvg").splitlines() if not l.startswith("#")]
    vcf_ref = [l for l in _ref_text("vcf").splitlines() if not l.startswith("#")]
    assert len(cvg_got) == len(cvg_ref) == 66613 and len(vcf_got) == len(vcf_ref) == 76 and n_indel == 57
    # the oracle's doubles are the reference's doubles and the formatter restates its fmt: the text is IDENTICAL
    # (only the Indels column, a hash-map iteration order in the reference, is compared as a multiset: SURVEY Q10)
    for g, w in zip(cvg_got, cvg_ref):
        assert _norm_indels(g) == _norm_indels(w), (g, w)
    for g, w in zip(vcf_got, vcf_ref):
        assert g == w, (g[:300], w[:300])


@pytest.mark.gpu
def test_c1_rerun_cli_matches_reference_text(tmp_path):
    from test_gpu_rerun import _compare_text
    length, lo, seq = _window()
    full = bytearray(b"N" * length)
    full[lo:lo + len(seq)] = seq.encode()
    with open(tmp_path / "chr17.fa", "wb") as f:
        f.write(b">chr17\n")
        for i in range(0, length, 60):
            f.write(full[i:i + 60]); f.write(b"\n")
    for t in range(T):
        shutil.copytree(os.path.join(C1, "test.out.tmp.thread.%d" % t), tmp_path / ("test.out.tmp.thread.%d" % t))
    shutil.copy(os.path.join(C1, "bam.list"), tmp_path / "bam.list")
    with open(os.path.join(C1, "ARGS")) as f:
        args = f.read().split()
    cmd = [host.BIN_PATH, "basetype", "--rerun"] + args + ["-i", "bam.list", "-r", "chr17.fa", "-o", "test.out"]
    r = subprocess.run(cmd, cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    cvg = batchfile.read_bgzf_text(str(tmp_path / "test.out.cvg.gz"))
    vcf = batchfile.read_bgzf_text(str(tmp_path / "test.out.vcf.gz"))
    _compare_text(cvg, vcf, _ref_text("cvg"), _ref_text("vcf"), N)
    assert len([l for l in vcf.splitlines() if not l.startswith("#")]) == 76
