#!/usr/bin/env python
"""Where the start-up time of a context goes (what the drop-in command waits for before its first tile can run):
python tools/create_probe.py [N_SAMPLES [TILE_SITES]]"""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import basevarc_b200 as B  # noqa: E402
from basevarc_b200 import caller as bc  # noqa: E402


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    S = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
    L = bc.load_library()
    t = [time.perf_counter()]
    n = L.bvc_device_count(); t.append(time.perf_counter())
    p0 = L.bvc_host_alloc(1 << 12); t.append(time.perf_counter())           # first runtime call that needs the primary context
    c1 = B.BaseTypeCaller(N, max_sites=S, max_entries=S * N); t.append(time.perf_counter())
    c2 = B.BaseTypeCaller(N, max_sites=S, max_entries=S * N); t.append(time.perf_counter())
    R = S * N + 64
    ptrs = [L.bvc_host_alloc(R) for _ in range(10)]; t.append(time.perf_counter())
    tile = B.synth_tile(N, S, 0.5, "lowpass", seed=1)
    tile.pack8()
    t.append(time.perf_counter())
    c1.call_sites(tile); t.append(time.perf_counter())                    # first launch of every kernel: lazy module loading
    c1.call_sites(tile); t.append(time.perf_counter())
    names = ["cudaGetDeviceCount (cuInit)", "first pinned alloc (primary context)", "bvc_create #1", "bvc_create #2",
             "10 pinned staging arrays of %d MB" % (R >> 20), "(synth tile)", "first call_sites (kernel loading)", "second call_sites"]
    print("devices", n)
    for k, nm in enumerate(names):
        print("%-45s %8.1f ms" % (nm, (t[k + 1] - t[k]) * 1e3))
    c1.close(); c2.close()


if __name__ == "__main__":
    main()
