#!/usr/bin/env python
"""Does running two half-windows on two streams (two contexts, two host threads) beat one full window on one stream?
Wall-clock over ITERS passes each; prints ms per 1M-site pass for 1 context x full window and P contexts x 1/P window."""
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import basevarc_b200 as B  # noqa: E402


def main():
    S = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
    iters = 20
    tile = B.synth_config("C2", n_sites=S, seed=1001)
    for parts in (1, 2, 4):
        bounds = [S * i // parts for i in range(parts + 1)]
        subs = [tile.slice_sites(bounds[i], bounds[i + 1]) for i in range(parts)]
        callers = []
        for t in subs:
            c = B.BaseTypeCaller(tile.n_samples, max_sites=t.n_sites, max_entries=t.n_entries)
            c.resident_upload(t)
            c.resident_run(3, flush_l2=False)
            callers.append(c)
        best = None
        for rep in range(3):
            ths = [threading.Thread(target=c.resident_run, args=(iters, False)) for c in callers]
            t0 = time.perf_counter()
            for th in ths:
                th.start()
            for th in ths:
                th.join()
            dt = (time.perf_counter() - t0) / iters * 1e3
            best = dt if best is None else min(best, dt)
        print("parts=%d  %.3f ms per %d-site pass (wall)" % (parts, best, S), flush=True)
        for c in callers:
            c.close()


if __name__ == "__main__":
    main()
