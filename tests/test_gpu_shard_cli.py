"""-m gpu: `BaseVarC_b200 basetype --rerun --shard_samples W` -- the sample-sharded compute phase driven by the C++ dispatcher
(csrc/host/shard_rerun.inc): rank r reads the batch files [r*J/W, (r+1)*J/W) of every window, the ranks exchange integer
histograms per tile, every rank formats its slice of the sites.  The text must be the reference's: on its own test/ example
(C1; 57 indel tokens exercise quirk Q1 ACROSS the ranks' file ranges) and on a 20,000-sample synthetic cohort against the
unmodified reference binary run here.  One GPU: all ranks on it (in-process exchange); >= W GPUs: one GPU per rank, NCCL."""
import os
import subprocess

import numpy as np
import pytest

import basevarc_b200 as B
from basevarc_b200 import batchfile, host
from test_c1 import N as C1_N, _ref_text
from test_gpu_rerun import _compare_text
from test_packed import _c1_workdir

pytestmark = pytest.mark.gpu


def _n_gpus():
    return B.load_library().bvc_device_count()


@pytest.mark.parametrize("W", [2, 3])
def test_shard_samples_cli_on_c1(tmp_path, W):
    args = _c1_workdir(tmp_path)
    r = subprocess.run([host.BIN_PATH, "basetype", "--rerun", "--keep_tmp", "-v", "--shard_samples", str(W), "--tile_sites", "5000"] + args,
                       cwd=tmp_path, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-2000:])
    assert "basetype sample-sharded : %d ranks" % W in r.stdout
    cvg = batchfile.read_bgzf_text(str(tmp_path / "test.out.cvg.gz"))
    vcf = batchfile.read_bgzf_text(str(tmp_path / "test.out.vcf.gz"))
    _compare_text(cvg, vcf, _ref_text("cvg"), _ref_text("vcf"), C1_N)
    assert len([l for l in vcf.splitlines() if not l.startswith("#")]) == 76


def _wide_cohort(tmp_path, n_samples=20000, n_sites=240):
    import refrun
    if not refrun.have_ref_bin():
        pytest.skip("oracle/_ref/BaseVarC_ref was not built")
    tile = B.synth_tile(n_samples, n_sites, 0.1, "highpoly", seed=77, indel_frac=0.02, first_pos=5001)
    work = str(tmp_path / "w")
    cvg_ref, vcf_ref = refrun.run_reference(tile, work, threads=2, batch=2000, keep_tmp=True)
    prefix = os.path.join(work, "out")
    cmd = [host.BIN_PATH, "basetype", "--rerun", "--keep_tmp", "-v", "-t", "2", "-b", "2000", "-i", os.path.join(work, "bam.list"),
           "-s", batchfile.region_for("chrS", 5001, tile.n_sites), "-r", os.path.join(work, "ref.fa"), "-o", prefix]
    return tile, cmd, prefix, cvg_ref, vcf_ref


@pytest.mark.parametrize("W", [2, 5])
def test_shard_samples_cli_wide_cohort_equals_reference_binary(tmp_path, W):
    tile, cmd, prefix, cvg_ref, vcf_ref = _wide_cohort(tmp_path)
    assert len([l for l in vcf_ref.splitlines() if not l.startswith("#")]) > 20
    r = subprocess.run(cmd + ["--shard_samples", str(W), "--tile_sites", "100"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-2000:])
    _compare_text(batchfile.read_bgzf_text(prefix + ".cvg.gz"), batchfile.read_bgzf_text(prefix + ".vcf.gz"), cvg_ref, vcf_ref, tile.n_samples)


def test_shard_samples_cli_two_gpus_nccl(tmp_path):
    if _n_gpus() < 2:
        pytest.skip("needs a 2-GPU lease (one GPU per rank, NCCL); the one-GPU form of the same path is covered above")
    tile, cmd, prefix, cvg_ref, vcf_ref = _wide_cohort(tmp_path)
    r = subprocess.run(cmd + ["--shard_samples", "2", "--tile_sites", "100"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (r.stdout[-500:], r.stderr[-2000:])
    assert "one GPU each, NCCL" in r.stdout
    _compare_text(batchfile.read_bgzf_text(prefix + ".cvg.gz"), batchfile.read_bgzf_text(prefix + ".vcf.gz"), cvg_ref, vcf_ref, tile.n_samples)


def test_window_sharded_cli_two_gpus(tmp_path):
    """--gpus 2: windows [g*T/G, (g+1)*T/G) on GPU g, one dispatcher thread and one context per GPU (csrc/host/rerun.cc)."""
    if _n_gpus() < 2:
        pytest.skip("needs a 2-GPU lease")
    import refrun
    tile, cvg_ref, vcf_ref = refrun.load_golden_tile("synth_ref_lowpass")
    from test_gpu_rerun import _run_cli
    cvg, vcf, _ = _run_cli(tile, str(tmp_path), 4, 16, extra=["--gpus", "2", "-v"])
    _compare_text(cvg, vcf, cvg_ref, vcf_ref, tile.n_samples)
