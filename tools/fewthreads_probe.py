#!/usr/bin/env python
"""SURVEY 8(f) N1 measured: the drop-in `basetype --rerun -t T` with T smaller than the core count (the spare cores inflate
BGZF blocks ahead of the T tokenising threads) against the same command without the pool, with zlib only, and against the
unmodified reference binary at the same -t.  python tools/fewthreads_probe.py [SITES [T]]  -> one JSON line"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                      # noqa: E402
import basevarc_b200 as B         # noqa: E402


def main():
    S = int(sys.argv[1]) if len(sys.argv) > 1 else 250000
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    tile = B.synth_config("C2", n_sites=S, seed=2)
    run = bench.ReferenceRun(tile, T)
    ours = os.path.join(ROOT, "basevarc_b200", "bin", "BaseVarC_b200")
    out = {"sites": S, "samples": tile.n_samples, "threads": T, "cores": os.cpu_count()}
    try:
        _, out["reference_seconds"] = run.step()
        for name, env in (("dropin_pool", {}), ("dropin_no_pool", {"BVC_INFLATE_THREADS": "0"}),
                          ("dropin_no_pool_zlib", {"BVC_INFLATE_THREADS": "0", "BVC_ZLIB_INFLATE": "1"})):
            best, phases = None, None
            for _ in range(2):
                t0 = time.perf_counter()
                r = subprocess.run([ours] + run.cmd[1:] + ["-v", "--no_packed"], cwd=run.dir, env=dict(os.environ, **env), capture_output=True, text=True)
                dt = time.perf_counter() - t0
                assert r.returncode == 0, (r.stderr or r.stdout)[-400:]
                if best is None or dt < best:
                    best = dt
                    phases = [l.split(":", 1)[1].strip() for l in r.stdout.splitlines() if l.startswith("basetype host phases") or l.startswith("basetype inflate pool")]
            out[name] = {"seconds": best, "phases": phases}
    finally:
        import shutil
        shutil.rmtree(run.dir, ignore_errors=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
