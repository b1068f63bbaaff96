"""-m gpu: SURVEY 8(b2) -- the reference's own C++ surface bound to the B200 libraries, compiled and run.

oracle/patch_reference.py copies the reference's src/BaseVarC.cpp out of tree, applies INTEGRATION.md's Level-1 patch (one call
in runBaseType) and, separately, its Level-2 patch (bt_s keeps its loop; tokeniser + bt_f replaced), compiles each against the
reference's headers and links it with the reference's other objects plus libbvc_host.so / libbvc_gpu.so.  Both binaries run the
reference's test/ example (C1) here and must print the text the UNMODIFIED reference printed (tests/golden/c1)."""
import os
import shutil
import subprocess

import pytest

from basevarc_b200 import batchfile
from test_c1 import C1, N, T, _ref_text, _window

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _workdir(tmp_path):
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
        return f.read().split() + ["-i", "bam.list", "-r", "chr17.fa", "-o", "test.out"]


@pytest.mark.gpu
@pytest.mark.parametrize("level", [1, 2])
def test_patched_reference_binary_on_c1(tmp_path, level):
    from test_gpu_rerun import _compare_text
    exe = os.path.join(ROOT, "oracle", "_ref", "BaseVarC_l%d" % level)
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/BaseVarC_l%d was not built (no reference tree at build time)" % level)
    args = _workdir(tmp_path)
    r = subprocess.run([exe, "basetype", "--rerun", "--keep_tmp"] + args, cwd=tmp_path, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, (r.stdout[-1000:], r.stderr[-2000:])
    assert "basetype done" in r.stdout
    cvg = batchfile.read_bgzf_text(str(tmp_path / "test.out.cvg.gz"))
    vcf = batchfile.read_bgzf_text(str(tmp_path / "test.out.vcf.gz"))
    _compare_text(cvg, vcf, _ref_text("cvg"), _ref_text("vcf"), N)
    assert len([l for l in vcf.splitlines() if not l.startswith("#")]) == 76
    assert len([l for l in cvg.splitlines() if not l.startswith("#")]) == 66613


def test_patch_script_inserts_only_the_binding():
    """The patch is the include line + the call (level 1) / the bt_s body (level 2): nothing else of the source moves."""
    import importlib.util
    if not os.path.isdir("/root/reference/src"):
        pytest.skip("no reference tree here")
    spec = importlib.util.spec_from_file_location("patch_reference", os.path.join(ROOT, "oracle", "patch_reference.py"))
    pr = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(pr)
    with open("/root/reference/src/BaseVarC.cpp") as f:
        src = f.read().splitlines()
    l1 = pr.patched_source(1).splitlines()
    assert len(l1) == len(src) + 4 and [x for x in l1 if x not in src] == ['#include "bvc_host.h"'] + pr.L1_INSERT.splitlines()
    l2 = pr.patched_source(2).splitlines()
    kept = [x for x in src if x in l2]
    assert len(kept) > 0.85 * len(src) and "bvc_packer_add_line" in "\n".join(l2) and "strtok_r(buf, \",\", &pti)" not in "\n".join(l2)
