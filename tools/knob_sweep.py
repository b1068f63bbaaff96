#!/usr/bin/env python
"""A/B runs of the library's environment knobs on ONE synthetic window (generated once), device-resident:
    python tools/knob_sweep.py C2 1000000 "" "BVC_REFILL_MIN=4" "BVC_REFILL_MIN=8 BVC_DRAIN_LANES=4"
prints the step time, the per-kernel times and the work counters per setting, and whether the records are
byte-identical to those of the first setting (integers must be; FP64 fields may move in their last bits)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import basevarc_b200 as B  # noqa: E402


def main():
    workload, S = sys.argv[1], int(sys.argv[2])
    tile = B.synth_config(workload, n_sites=S, seed=1001)
    first = None
    for envs in sys.argv[3:]:
        kv = dict(e.split("=", 1) for e in envs.split())
        os.environ.update(kv)
        c = B.BaseTypeCaller(tile.n_samples, max_sites=tile.n_sites, max_entries=tile.n_entries)
        c.resident_upload(tile)
        c.resident_run(3, True)
        ms = min(c.resident_run(5, True)[0] for _ in range(3))
        _, per = c.resident_run(3, True, per_kernel=True)
        rec = c.resident_download().copy()
        same = "first"
        if first is None:
            first = rec
        else:
            raw = lambda r: r.view(np.uint8).reshape(len(r), -1)
            same = "identical" if rec.tobytes() == first.tobytes() else "bytes differ in %d of %d records" % (
                int(np.sum(np.any(raw(rec) != raw(first), axis=1))), len(rec))
        print("%-40s step %.3f ms" % (envs or "(defaults)", ms), {k: round(v, 3) for k, v in per.items() if v > 0.005},
              c.work_counters(), same, flush=True)
        c.close()
        for k in kv:
            del os.environ[k]


if __name__ == "__main__":
    main()
