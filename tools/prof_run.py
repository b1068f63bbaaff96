#!/usr/bin/env python
"""The smallest program that runs the resident pipeline: python tools/prof_run.py [WORKLOAD [SITES [ITERS]]]
(what ncu wraps: `ncu --set full -k regex:site_hist8 -c 1 python tools/prof_run.py C2 1000000 1`)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import basevarc_b200 as B  # noqa: E402


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "C2"
    S = int(sys.argv[2]) if len(sys.argv) > 2 else 1000000
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    tile = B.synth_config(wl, n_sites=S, seed=1001)
    if os.environ.get("PROF_WIRE", "packed8") == "packed8":
        tile.pack8()
    c = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
    c.resident_upload(tile)
    ms, per = c.resident_run(iters, False, per_kernel=True)
    print("%s %d sites: %.3f ms" % (wl, S, ms), {k: round(v, 3) for k, v in per.items() if v > 0.005}, c.work_counters())
    c.close()


if __name__ == "__main__":
    main()
