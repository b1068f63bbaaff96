#!/usr/bin/env python
"""oracle/patch_reference.py -- TEST INFRASTRUCTURE: the compiled proof of INTEGRATION.md's bindings.

Copies /root/reference/src/BaseVarC.cpp OUT OF TREE (oracle/_ref/patched/, git-ignored), applies

  level 1: one call -- runBaseType() hands the compute phase to bvc_basetype_main() after the load phase, and
  level 2: bt_s keeps its files / window / loop, the tokeniser + bt_f go through the packer and bvc_enqueue / bvc_wait
           (oracle/patches/level2_bt_s.inc),

compiles each against the reference's own headers and links it with the reference objects oracle/Makefile already built plus
basevarc_b200/libbvc_host.so / libbvc_gpu.so:  oracle/_ref/BaseVarC_l1, oracle/_ref/BaseVarC_l2.
tests/test_gpu_binding.py runs both on C1 (tests/golden/c1) and compares with the reference's text.
Nothing of the reference is committed: only the lines this script inserts.  usage: python oracle/patch_reference.py   (reference tree: $BVC_REFERENCE_DIR or /root/reference)"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("BVC_REFERENCE_DIR", "/root/reference")
OUT = os.path.join(HERE, "_ref")
PATCHED = os.path.join(OUT, "patched")

INCLUDE_ANCHOR = '#include "robin_hood.h"\n'
L1_ANCHOR = "    if (opt::load) exit(EXIT_SUCCESS);\n"
L1_INSERT = ("    // INTEGRATION.md Level 1: the compute phase (worker spawn, bt_s, sub-file merge; :262-311) is the B200 dispatcher.\n"
             "    // argv[0] is \"basetype\" here; it re-parses the same options and re-validates the batch files.\n"
             "    exit(bvc_basetype_main(argc, argv));\n")
L2_BEGIN = "    // begin to call basetype and output\n"
L2_END = '    if (bgzf_close(fpv) < 0) std::cerr << "warning: file cannot be closed" << std::endl;\n'


def patched_source(level):
    with open(os.path.join(REF, "src", "BaseVarC.cpp")) as f:
        src = f.read()
    assert src.count(INCLUDE_ANCHOR) == 1 and src.count(L1_ANCHOR) == 1 and src.count(L2_BEGIN) == 1 and src.count(L2_END) == 1, \
        "the reference source does not look like the revision this script was written against"
    src = src.replace(INCLUDE_ANCHOR, INCLUDE_ANCHOR + '#include "bvc_host.h"\n')
    if level == 1:
        return src.replace(L1_ANCHOR, L1_ANCHOR + L1_INSERT)
    a, b = src.index(L2_BEGIN), src.index(L2_END)
    with open(os.path.join(HERE, "patches", "level2_bt_s.inc")) as f:
        return src[:a] + f.read() + src[b:]


def inserted_lines(level):
    """What the binding adds, for INTEGRATION.md (the removed span is named by its anchors)."""
    if level == 1:
        return '#include "bvc_host.h"\n' + L1_INSERT
    with open(os.path.join(HERE, "patches", "level2_bt_s.inc")) as f:
        return '#include "bvc_host.h"\n' + f.read()


def build():
    if not os.path.isdir(os.path.join(REF, "src")):
        print("no reference tree at %s: nothing to patch" % REF)
        return 0
    objs = [o for o in glob.glob(os.path.join(OUT, "build", "*.o")) if os.path.basename(o) not in ("bvc_BaseVarC.o", "ref_shim.o")]
    if not objs:
        subprocess.run(["make", "-s", "-C", HERE, "ref"], check=True)
        objs = [o for o in glob.glob(os.path.join(OUT, "build", "*.o")) if os.path.basename(o) not in ("bvc_BaseVarC.o", "ref_shim.o")]
    os.makedirs(PATCHED, exist_ok=True)
    S, H = os.path.join(REF, "SeqLib"), os.path.join(REF, "SeqLib", "htslib")
    lib = os.path.join(ROOT, "basevarc_b200")
    for level in (1, 2):
        cpp = os.path.join(PATCHED, "BaseVarC_l%d.cpp" % level)
        with open(cpp, "w") as f:
            f.write(patched_source(level))
        obj = cpp[:-4] + ".o"
        subprocess.run(["g++", "-O2", "-w", "-std=c++11", "-include", "limits", "-I" + os.path.join(REF, "src"), "-I" + S, "-I" + H,
                        "-I" + os.path.join(REF, "fmt", "include"), "-I" + os.path.join(OUT, "build", "hts"), "-I" + os.path.join(ROOT, "include"),
                        "-c", cpp, "-o", obj], check=True)
        exe = os.path.join(OUT, "BaseVarC_l%d" % level)
        subprocess.run(["g++", "-o", exe, obj] + objs + ["-L" + lib, "-lbvc_host", "-lbvc_gpu", "-lz", "-lpthread",
                        "-Wl,-rpath,$ORIGIN/../../basevarc_b200"], check=True)
        print("built", exe)
    return 0


if __name__ == "__main__":
    sys.exit(build())
