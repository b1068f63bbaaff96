#!/usr/bin/env python
"""Times the drop-in CLI (basevarc_b200/bin/BaseVarC_b200 basetype --rerun -v) on synthetic batch files for several
--tile_sites settings and prints its host-phase report.  usage: cli_sweep.py [sites_per_thread] [tile_sites ...]"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench                      # noqa: E402
import basevarc_b200 as B         # noqa: E402


def main():
    spt = int(sys.argv[1]) if len(sys.argv) > 1 else 6000
    tiles = [int(x) for x in sys.argv[2:]] or [0]
    cores = os.cpu_count() or 1
    import torch
    torch.zeros(1, device="cuda")          # the parent holds a context, as bench.py does: the driver stays initialised
    tile = B.synth_config("C2", n_sites=spt * cores, seed=2)
    run = bench.ReferenceRun(tile, cores)
    ours = os.path.join(ROOT, "basevarc_b200", "bin", "BaseVarC_b200")
    try:
        for ts in tiles:
            extra = ["--tile_sites", str(ts)] if ts else []
            for rep in range(2):
                t0 = time.perf_counter()
                r = subprocess.run([ours] + run.cmd[1:] + ["-v"] + extra, cwd=run.dir, capture_output=True, text=True)
                dt = time.perf_counter() - t0
                ph = [l for l in r.stdout.splitlines() if l.startswith("basetype host phases")]
                print("tile_sites=%d rep=%d rc=%d %.3f s | %s" % (ts, rep, r.returncode, dt, ph[0] if ph else r.stderr[-200:]), flush=True)
    finally:
        run.close()


if __name__ == "__main__":
    main()
