#!/usr/bin/env python
"""`BaseVarC_b200 basetype --pack` (text batch files -> packed tile files: inflate + tokenise, no GPU) with T window threads and
the inflate pool on the spare cores against the same command with BVC_INFLATE_THREADS=0 and with zlib only: what the pool and the
whole-buffer decoder buy a run with fewer window threads than cores.  python tools/pack_pool_probe.py [SITES [T]]"""
import os
import shutil
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import basevarc_b200 as B  # noqa: E402
from basevarc_b200 import batchfile, host  # noqa: E402


def main():
    S = int(sys.argv[1]) if len(sys.argv) > 1 else 60000
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    tile = B.synth_config("C2", n_sites=S, seed=5)
    d = tempfile.mkdtemp(prefix="bvc_pool_", dir=os.path.join(ROOT, "gpurun_out"))
    try:
        prefix = os.path.join(d, "out")
        host.write_batch_files(tile, prefix, T, 10)
        fa = os.path.join(d, "ref.fa")
        batchfile.write_fasta(fa, "chrS", tile.ref_base, 1001)
        with open(os.path.join(d, "bam.list"), "w") as f:
            f.write("".join("s%d.bam\n" % i for i in range(tile.n_samples)))
        cmd = [host.BIN_PATH, "basetype", "--pack", "-v", "-t", str(T), "-b", "10", "-i", os.path.join(d, "bam.list"), "-s",
               batchfile.region_for("chrS", 1001, tile.n_sites), "-r", fa, "-o", prefix]
        for name, env in (("pool + whole-buffer decoder", {}), ("no pool", {"BVC_INFLATE_THREADS": "0"}),
                          ("no pool, zlib only", {"BVC_INFLATE_THREADS": "0", "BVC_ZLIB_INFLATE": "1"})):
            best = None
            for _ in range(2):
                t0 = time.perf_counter()
                r = subprocess.run(cmd, cwd=d, env=dict(os.environ, **env), capture_output=True, text=True)
                dt = time.perf_counter() - t0
                assert r.returncode == 0, r.stderr[-500:]
                best = dt if best is None else min(best, dt)
            pool = [l for l in r.stdout.splitlines() if "inflate pool" in l]
            print("%-30s %6.2f s  (%d sites x %d samples, -t %d, %d cores)  %s" % (name, best, S, tile.n_samples, T, os.cpu_count(), pool[0] if pool else ""))
    finally:
        shutil.rmtree(d, ignore_errors=True)


if __name__ == "__main__":
    main()
