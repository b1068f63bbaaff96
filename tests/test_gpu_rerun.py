"""-m gpu: the drop-in command end to end.  `BaseVarC_b200 basetype --rerun` on batch files written in the
reference's format must print what the UNMODIFIED reference binary printed for the same files (tests/golden):
same rows, same integers / alleles / FILTER / per-sample columns, floats equal to the printed digit."""
import os
import subprocess

import numpy as np
import pytest

import basevarc_b200 as B
from basevarc_b200 import batchfile, host
from test_host_cpp import _norm_indels, _same_number_text

pytestmark = pytest.mark.gpu


def _run_cli(tile, workdir, threads, batch, extra=(), chrom="chrS"):
    os.makedirs(workdir, exist_ok=True)
    prefix = os.path.join(workdir, "out")
    first = int(tile.pos[0])
    batchfile.write_batch_files(tile, prefix, threads, batch)
    fa = os.path.join(workdir, "ref.fa")
    batchfile.write_fasta(fa, chrom, tile.ref_base, first)
    bl = os.path.join(workdir, "bam.list")
    with open(bl, "w") as f:
        f.write("".join("s%d.bam\n" % i for i in range(tile.n_samples)))
    cmd = [host.BIN_PATH, "basetype", "--rerun", "-t", str(threads), "-b", str(batch), "-i", bl,
           "-s", batchfile.region_for(chrom, first, tile.n_sites), "-r", fa, "-o", prefix] + list(extra)
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "basetype done" in r.stdout and "basetype computing done" in r.stdout
    return batchfile.read_bgzf_text(prefix + ".cvg.gz"), batchfile.read_bgzf_text(prefix + ".vcf.gz"), prefix


def _compare_text(cvg, vcf, cvg_ref, vcf_ref, n_samples):
    head = lambda t: [l for l in t.splitlines() if l.startswith("#") and not l.startswith("##reference=")]
    body = lambda t: [l for l in t.splitlines() if not l.startswith("#")]
    assert head(cvg) == head(cvg_ref)
    assert head(vcf) == head(vcf_ref)
    gc, wc = body(cvg), body(cvg_ref)
    assert len(gc) == len(wc)
    for g, w in zip(gc, wc):
        gf, wf = _norm_indels(g), _norm_indels(w)
        assert len(gf) == len(wf) and all(_same_number_text(x, y) for x, y in zip(gf, wf)), (g, w)
    gv, wv = body(vcf), body(vcf_ref)
    assert len(gv) == len(wv)
    for g, w in zip(gv, wv):
        gf, wf = g.split("\t"), w.split("\t")
        assert len(gf) == len(wf) == 9 + n_samples
        assert gf[:5] == wf[:5] and gf[6] == wf[6] and gf[8:] == wf[8:], (gf[:9], wf[:9])
        assert _same_number_text(gf[5], wf[5]), (gf[5], wf[5])
        gi, wi = gf[7].split(";"), wf[7].split(";")
        assert [x.split("=")[0] for x in gi] == [x.split("=")[0] for x in wi]
        for x, y in zip(gi, wi):
            assert all(_same_number_text(p, q) for p, q in zip(x.split("=")[1].split(","), y.split("=")[1].split(","))), (x, y)


@pytest.mark.parametrize("name,threads,batch", [("synth_ref_lowpass", 3, 16), ("synth_ref_highpoly", 2, 10)])
def test_rerun_matches_reference_binary_text(tmp_path, name, threads, batch):
    import refrun
    tile, cvg_ref, vcf_ref = refrun.load_golden_tile(name)
    cvg, vcf, prefix = _run_cli(tile, str(tmp_path), threads, batch)
    _compare_text(cvg, vcf, cvg_ref, vcf_ref, tile.n_samples)
    # default: batch files and tmp dirs are removed like the reference does (src/BaseVarC.cpp:458-462,299-303)
    assert not os.path.exists(prefix + ".tmp.thread.0")


def test_rerun_is_invariant_to_tiling_and_partition(tmp_path):
    """Small tiles (many enqueues, stale-token state crossing tile boundaries) and another -t/-b partition give
    the same text."""
    tile = B.synth_tile(64, 1500, 1.5, "highpoly", seed=31, indel_frac=0.05, first_pos=2001)
    a = _run_cli(tile, str(tmp_path / "a"), 1, 64, extra=["--keep_tmp"])
    b = _run_cli(tile, str(tmp_path / "b"), 4, 9, extra=["--tile_sites", "37"])
    assert a[0] == b[0] and a[1].split("##reference")[0] == b[1].split("##reference")[0]
    body = lambda t: [l for l in t.splitlines() if not l.startswith("#")]
    assert body(a[1]) == body(b[1]) and len(body(a[1])) > 10
    assert os.path.exists(a[2] + ".tmp.thread.0/batch.0")


def test_rerun_group_path_matches_reference_binary_text(tmp_path):
    import refrun
    tile, cvg_ref, vcf_ref = refrun.load_golden_tile("synth_ref_groups")
    names = ["EAS", "AFR", "EUR", None]
    gf = str(tmp_path / "groups.txt")
    with open(gf, "w") as f:
        f.write("".join("S%d %s\n" % (i, names[i % 4]) for i in range(tile.n_samples) if names[i % 4]))
    cvg, vcf, _ = _run_cli(tile, str(tmp_path / "w"), 2, 16, extra=["-g", gf, "--tile_sites", "250"])
    _compare_text(cvg, vcf, cvg_ref, vcf_ref, tile.n_samples)
    assert "##INFO=<ID=AFR_AF" in vcf and "EUR_AF=" in vcf
