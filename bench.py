#!/usr/bin/env python
"""bench.py -- site*samples/s of the basetype calling core (EM + LRT + strand/Fisher/rank-sum).

A step = one pass of the hot path over one window of the BASELINE.json workload C2 (synthetic
1,000 samples x 1 Mb at 0.5x, SURVEY.md section 8(d) generator) per GPU.  N GPUs = N windows
(sites are independent: window sharding, no collective, weak scaling).

  value : whole-job site*samples/s with the packed pileup already resident in HBM (device events)
  e2e   : the same through the C ABI with HOST buffers (pinned H2D of the tile + D2H of the records
          inside the timed region), double-buffered over sub-tiles
  roofline / fp64 / cpu_baseline : see DESIGN.md "Measurement"

`--impl reference` times the reference's own CPU implementation on the host cores instead.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# NCCL prints its banner on STDOUT at the VERSION and WARN debug levels (the contract is ONE JSON line there), so its log goes
# to a FILE per process at INFO level: the communicator lines ("... nranks N ...") are then observable (comm_nranks below)
NCCL_LOG = os.path.join(os.environ.get("BVC_NCCL_LOG_DIR", "/tmp"), "bvc_nccl_%d_%%h_%%p.log" % os.getpid())
os.environ["NCCL_DEBUG"] = os.environ.get("BVC_NCCL_DEBUG", "INFO")        # (the image presets NCCL_DEBUG=VERSION)
os.environ["NCCL_DEBUG_FILE"] = os.environ.get("BVC_NCCL_DEBUG_FILE", NCCL_LOG)

METRIC = "site*samples/sec for basetype EM+LRT"
# FP64 operations per unit of work, FMA = 2, counted on the code that ships:
#  em_fit_kernel, per (base, qual) class per singleEM pass (em_class_major, bvc_pipeline.cuh): exact count DADD 1,
#    denominator DFMA 2, one Newton step on the MUFU seed 2 DFMA = 4, weight DMUL 1, V += DADD 1, accT DFMA 2 = 11
#    (7 FP64 instructions + MUFU.RCP64H; `cuobjdump -sass` of the major loop shows exactly these per class)
#  em_long_kernel, per class per pass (em_class_plain): count 1, e3*F 1, ml DFMA 2, two Newton steps 6, weight 1,
#    accT DFMA 2, d*w 1, V += 1 = 15
#  fit_loglik_kernel, per class evaluation: ml 3 + libdevice log + the weighted sums; 36 FP64 instructions per class in
#    the SASS of its 4x unrolled loop (21 DFMA, 6 DMUL, 9 DADD) = 57
FLOP_PER_CLASS_PASS = 11
FLOP_PER_LONG_CLASS_PASS = 15
FLOP_PER_LOGLIK_EVAL = 57
UNIT = "site*samples/s"
B_CONFIG_SAMPLES = {"C2": 1000, "C3": 10000, "C4": 100000, "C5": 2000}


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json, burst copy)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.proc = None
        self.lines = []
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def wait_started(self, timeout=15.0):
        """nvidia-smi's start-up can stall GPU work for seconds: let it settle before the timed region."""
        t0 = time.time()
        while self.proc and len(self.lines) < 3 and time.time() - t0 < timeout:
            time.sleep(0.05)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.th.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def nccl_ranks_seen():
    """(max nranks in this process's NCCL log, whether NVLS shows up) -- None when no communicator line was logged."""
    import glob
    import re
    pat = os.environ.get("NCCL_DEBUG_FILE", NCCL_LOG).replace("%h", "*").replace("%p", "*")
    best, nvls = None, False
    for fn in glob.glob(pat):
        try:
            with open(fn, errors="replace") as f:
                txt = f.read()
        except OSError:
            continue
        for m in re.finditer(r"nranks (\d+)", txt):
            best = max(best or 0, int(m.group(1)))
        nvls = nvls or ("NVLS" in txt)
    return best, nvls


def bind_rank_to_cores(local, world):
    """Give every rank of the node its own slice of the host cores that are local to its GPU (sysfs local_cpulist of the GPU's
    PCI function), before any pinned buffer is allocated (first touch).  Returns a description for the JSON line."""
    try:
        import torch
        pr = torch.cuda.get_device_properties(local)
        bdf = "%04x:%02x:%02x.0" % (getattr(pr, "pci_domain_id", 0), pr.pci_bus_id, pr.pci_device_id)
        with open("/sys/bus/pci/devices/%s/local_cpulist" % bdf) as f:
            txt = f.read().strip()
        cpus = []
        for part in txt.split(","):
            a, _, b = part.partition("-")
            cpus += list(range(int(a), int(b or a) + 1))
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0))) or sorted(os.sched_getaffinity(0))
        share = max(1, len(allowed) // max(world, 1))
        mine = allowed[(local * share) % len(allowed):][:share] or allowed
        os.sched_setaffinity(0, mine)
        numa = None
        try:
            with open("/sys/bus/pci/devices/%s/numa_node" % bdf) as f:
                numa = int(f.read())
        except (OSError, ValueError):
            pass
        return {"gpu_pci": bdf, "gpu_local_cpulist": txt, "gpu_numa_node": numa, "bound_to": "%d-%d" % (mine[0], mine[-1]), "n_cores": len(mine)}
    except Exception as e:                      # no sysfs / not permitted: unbound, and say so
        return {"bound_to": None, "why": repr(e)[:120]}


def copy_probe(local, nbytes_h2d, nbytes_d2h, barrier, reps=5):
    """Copy-only probe of the e2e leg's link: the same number of bytes, page-locked host memory, cudaMemcpyAsync only, all ranks
    at once.  Returns (H2D GB/s, D2H GB/s, both directions at once GB/s each) of THIS rank (events on the copy streams)."""
    import torch
    dev = torch.device("cuda", local)
    h_in = torch.empty(nbytes_h2d, dtype=torch.uint8).pin_memory()
    d_in = torch.empty(nbytes_h2d, dtype=torch.uint8, device=dev)
    d_out = torch.empty(max(nbytes_d2h, 1), dtype=torch.uint8, device=dev)
    h_out = torch.empty(max(nbytes_d2h, 1), dtype=torch.uint8).pin_memory()
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def timed(do_in, do_out):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(s1):
            e0.record()
            for _ in range(reps if do_in else 0):
                d_in.copy_(h_in, non_blocking=True)
            e1.record()
        with torch.cuda.stream(s2):
            f0.record()
            for _ in range(reps if do_out else 0):
                h_out.copy_(d_out, non_blocking=True)
            f1.record()
        torch.cuda.synchronize(dev)
        gi = reps * nbytes_h2d / (e0.elapsed_time(e1) * 1e-3) / 1e9 if do_in else None
        go = reps * nbytes_d2h / (f0.elapsed_time(f1) * 1e-3) / 1e9 if do_out else None
        return gi, go

    timed(True, True)                           # warm-up
    h2d, _ = timed(True, False)
    _, d2h = timed(False, True)
    bi, bo = timed(True, True)
    return {"h2d_gbs": h2d, "d2h_gbs": d2h, "both_h2d_gbs": bi, "both_d2h_gbs": bo}


def _workload(name, n_sites, rank):
    import basevarc_b200 as B
    cfg = B.CONFIGS[name]
    S = n_sites or cfg["n_sites"]
    tile = B.synth_config(name, n_sites=S, seed=int(name[1:]) + 1000 * rank)
    return cfg, tile


def _alg_bytes(tile):
    """4 B per covered entry + 8 B per site header + 192 B result per non-empty site."""
    nonempty = int((np.diff(tile.entry_off.astype(np.int64)) > 0).sum())
    return 4 * tile.n_entries + 8 * tile.n_sites + 192 * nonempty


def cpu_port_baseline(tile, params, budget_sites_per_thread, threads):
    """The oracle's C restatement (per-read, reference operation order) on `threads` host threads."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py
    oracle_py.port()
    S = min(tile.n_sites, budget_sites_per_thread * threads)
    bounds = np.linspace(0, S, threads + 1).astype(int)
    ths = [threading.Thread(target=oracle_py.call_sites, args=(tile, params, int(bounds[i]), int(bounds[i + 1])))
           for i in range(threads)]
    t0 = time.perf_counter()
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    dt = time.perf_counter() - t0
    return S * tile.n_samples / dt, S, dt


REF_BIN = os.path.join(ROOT, "oracle", "_ref", "BaseVarC_ref")


class ReferenceRun:
    """The UNMODIFIED reference binary (oracle/_ref/BaseVarC_ref, built from /root/reference by oracle/Makefile) on
    batch files written from the tile: `basetype --rerun --keep_tmp -t <host cores>` = its compute phase
    (BGZF inflate, tokenising, bt_f per site, formatting, sub-file merge; src/BaseVarC.cpp:262-311)."""

    def __init__(self, tile, cores, batch=250):
        import tempfile
        from basevarc_b200 import batchfile, host
        self.tile, self.cores = tile, cores
        self.dir = tempfile.mkdtemp(prefix="bvc_ref_")
        self.prefix = os.path.join(self.dir, "out")
        first = 1001
        host.write_batch_files(tile, self.prefix, cores, batch)
        fa = os.path.join(self.dir, "ref.fa")
        batchfile.write_fasta(fa, "chrS", tile.ref_base, first)
        bl = os.path.join(self.dir, "bam.list")
        with open(bl, "w") as f:
            f.write("".join("s%d.bam\n" % i for i in range(tile.n_samples)))
        self.cmd = [REF_BIN, "basetype", "--rerun", "--keep_tmp", "-t", str(cores), "-b", str(batch), "-i", bl,
                    "-s", batchfile.region_for("chrS", first, tile.n_sites), "-r", fa, "-o", self.prefix]

    def step(self):
        t0 = time.perf_counter()
        r = subprocess.run(self.cmd, cwd=self.dir, capture_output=True, text=True)
        dt = time.perf_counter() - t0
        if r.returncode != 0 or "basetype done" not in r.stdout:
            raise RuntimeError("reference binary failed: %s" % r.stderr[-500:])
        return self.tile.n_sites * self.tile.n_samples / dt, dt

    def dropin(self):
        """This repo's drop-in (basevarc_b200/bin/BaseVarC_b200, same command line) on the same batch files: wall time of
        the whole command (CUDA context creation included) and whether its .cvg / .vcf text equals the reference's."""
        import gzip
        ours = os.path.join(ROOT, "basevarc_b200", "bin", "BaseVarC_b200")
        if not os.path.exists(ours):
            return None
        ref_txt = {}
        for ext in ("cvg", "vcf"):
            with gzip.open(self.prefix + "." + ext + ".gz", "rb") as f:
                ref_txt[ext] = f.read()
            os.rename(self.prefix + "." + ext + ".gz", self.prefix + ".ref." + ext + ".gz")
        best = None
        for _ in range(2):                                   # second run: page cache and driver warm, as for the reference
            t0 = time.perf_counter()
            r = subprocess.run([ours] + self.cmd[1:] + ["-v"], cwd=self.dir, capture_output=True, text=True)
            dt = time.perf_counter() - t0
            if r.returncode != 0:
                return {"error": (r.stderr or r.stdout)[-300:]}
            if best is None or dt < best:
                best = dt
                phases = [l.split(":", 1)[1].strip() for l in r.stdout.splitlines() if l.startswith("basetype host phases")]
        out = {"seconds": best, "value": self.tile.n_sites * self.tile.n_samples / best, "unit": UNIT,
               "host_phases": phases[0] if phases else None}
        out.update(self._text_diff(ref_txt))
        # the same command from PACKED TILE FILES (include/bvc_host.h; `--pack` converts the text batch files once): no inflate,
        # no tokeniser on the rerun
        t0 = time.perf_counter()
        r = subprocess.run([ours, "basetype", "--pack"] + self.cmd[3:], cwd=self.dir, capture_output=True, text=True)
        pack_s = time.perf_counter() - t0
        if r.returncode == 0:
            best = None
            for _ in range(2):
                t0 = time.perf_counter()
                r = subprocess.run([ours] + self.cmd[1:] + ["-v"], cwd=self.dir, capture_output=True, text=True)
                dt = time.perf_counter() - t0
                if r.returncode != 0 or "reading packed tile files" not in r.stderr:
                    best = None
                    break
                if best is None or dt < best:
                    best = dt
                    phases = [l.split(":", 1)[1].strip() for l in r.stdout.splitlines() if l.startswith("basetype host phases")]
            if best is not None:
                pk = {"pack_seconds": pack_s, "seconds": best, "value": self.tile.n_sites * self.tile.n_samples / best,
                      "host_phases": phases[0] if phases else None,
                      "packed_bytes": sum(os.path.getsize(os.path.join(self.prefix + ".tmp.thread.%d" % t, "packed.bvc")) for t in range(self.cores)),
                      "text_bgzf_bytes": sum(os.path.getsize(os.path.join(dp, f)) for t in range(self.cores)
                                             for dp, _, fs in os.walk(self.prefix + ".tmp.thread.%d" % t) for f in fs if f.startswith("batch."))}
                pk.update(self._text_diff(ref_txt))
                out["from_packed_files"] = pk
            else:
                out["from_packed_files"] = {"error": (r.stderr or r.stdout)[-300:]}
        else:
            out["from_packed_files"] = {"error": (r.stderr or r.stdout)[-300:]}
        return out

    def _text_diff(self, ref_txt):
        import gzip
        out = {}
        for ext in ("cvg", "vcf"):
            with gzip.open(self.prefix + "." + ext + ".gz", "rb") as f:
                mine = f.read()
            a, b = mine.split(b"\n"), ref_txt[ext].split(b"\n")
            body = lambda ls: [l for l in ls if not l.startswith(b"##")]      # header lines carry the command / date
            a, b = body(a), body(b)
            if ext == "cvg":
                # column 9 (Indels) comes out in the reference's hash-map iteration order: compared as a multiset
                def norm(l):
                    f = l.split(b"\t")
                    if len(f) > 8:
                        f[8] = b",".join(sorted(f[8].split(b",")))
                    return b"\t".join(f)
                a, b = [norm(l) for l in a], [norm(l) for l in b]
            out[ext + "_lines"] = len(b)
            diff = [(x, y) for x, y in zip(a, b) if x != y]
            out[ext + "_lines_differing"] = len(diff) + abs(len(a) - len(b))
            if diff:
                out[ext + "_diff_sample"] = [[x.decode()[:160], y.decode()[:160]] for x, y in diff[:3]]
        return out

    def close(self):
        import shutil
        shutil.rmtree(self.dir, ignore_errors=True)


def cpu_reference_baseline(tile, sites_per_thread, cores, steps=1, warmup=0, dropin=None):
    """(value, sites, seconds per step, kind): the reference binary when it was built, else the oracle port."""
    S = min(tile.n_sites, sites_per_thread * cores)
    if os.path.exists(REF_BIN):
        sub = tile.slice_sites(0, S)
        run = ReferenceRun(sub, cores)
        try:
            for _ in range(warmup):
                run.step()
            res = [run.step() for _ in range(max(steps, 1))]
            if dropin is not None:
                dropin["ref_seconds"] = float(np.mean([d for _, d in res]))
                dropin.update(run.dropin() or {"error": "BaseVarC_b200 not built"})
        finally:
            run.close()
        return float(np.mean([v for v, _ in res])), S, float(np.mean([d for _, d in res])), "reference"
    import basevarc_b200 as B
    params = B.default_params(tile.n_samples)
    res = [cpu_port_baseline(tile, params, sites_per_thread, cores) for _ in range(max(steps, 1))]
    return float(np.mean([r[0] for r in res])), S, float(np.mean([r[2] for r in res])), "port"


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on all host cores (rank 0 only)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    # the sample is sized in site*samples: --ref-sites-per-thread is quoted for 1,000 samples (C2)
    per_thread = max(1, args.ref_sites_per_thread * 1000 // B_CONFIG_SAMPLES[args.workload])
    cfg, tile = _workload(args.workload, per_thread * cores, 0)
    value, S, dt, kind = cpu_reference_baseline(tile, per_thread, cores, steps=args.steps, warmup=args.warmup)
    what = ("unmodified reference binary `BaseVarC basetype --rerun -t %d` (compute phase: inflate, tokenise, bt_f, format, merge)" % cores
            if kind == "reference" else "oracle C restatement, per-read, %d threads" % cores)
    sample = "%d sites x %d samples of %s per step; %s" % (S, tile.n_samples, args.workload, what)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "%s: synthetic %d samples x 1 Mb at %.1fx (bounded sample)" % (args.workload, tile.n_samples, cfg["depth"])},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))

def run_sharded(args, return_line=False):
    """--workload C4: the SAMPLE-SHARDED path.  Every rank holds its own sample range of the same
    window; per tile one reduce-scatter of integer histograms + all-gather + all-reduce over NCCL.  Total work is
    fixed (100,000 samples), so scaling is "strong"."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    import torch
    import torch.distributed as dist
    import basevarc_b200 as B
    from basevarc_b200.caller import PinnedArray
    torch.cuda.set_device(local)
    own_pg = not dist.is_initialized()
    if own_pg:
        if world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        else:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29555")
            dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", local))

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    cfg = B.CONFIGS[args.workload]
    N = cfg["n_samples"]
    S = args.sites or 16384
    tile = B.synth_tile(N, S, cfg["depth"], cfg["model"], seed=int(args.workload[1:]), shard=(rank, world))
    R = tile.n_entries
    uid = [B.shard_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    qlo, qhi = args.qual_window
    caller = B.ShardedCaller(N, rank, world, uid[0], qual_lo=qlo, qual_hi=qhi, device=local, max_sites=S, max_entries=max(R, 1))
    passes = 1 if not args.window_sites else max(1, -(-args.window_sites // S))      # the whole window = passes x the resident tile
    units = S * N * passes                          # whole job: every rank works on the same sites x N samples

    # ---- resident leg: local entries already in HBM; device events per phase (max over ranks of the step time) ----
    caller.shard_upload(tile)
    for _ in range(args.warmup):
        caller.shard_run()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.wait_started()
    barrier()
    l0 = caller.kernel_launches
    t0 = time.perf_counter()
    phases = {}
    for _ in range(args.steps * passes):
        first, recs = caller.shard_run()
        st = caller.shard_stats()
        for k, v in st["phase_ms"].items():
            phases[k] = phases.get(k, 0.0) + v / args.steps
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    launches = caller.kernel_launches - l0
    dev_ms = sum(v for k, v in phases.items() if k not in ("h2d", "d2h"))
    ms_step = max_over_ranks(dev_ms)
    value = units / (ms_step * 1e-3)
    phases = {k: max_over_ranks(v) for k, v in phases.items()}

    # ---- e2e leg: pinned host buffers through bvc_shard_call (H2D + collectives + D2H inside the region) --------
    pins = {}
    packed = args.wire in ("packed8", "coded") and tile.pack8()
    for k in ("ref_base", "entry_off", "mapq", "rpr") + (("packed8",) if packed else ("code", "qual")):
        src = getattr(tile, k)
        pins[k] = PinnedArray(src.shape, src.dtype)
        pins[k].array[...] = src
        setattr(tile, k, pins[k].array)
    out = PinnedArray(((S + world - 1) // world,), B.SITE_RESULT_DTYPE)
    for _ in range(args.warmup):
        caller.shard_call(tile, out.array)
    barrier()
    h2d0, d2h0 = caller.transfer_bytes
    x0 = caller.shard_stats()["exchange_bytes"]
    t0 = time.perf_counter()
    for _ in range(args.steps * passes):
        first2, recs2 = caller.shard_call(tile, out.array)
    barrier()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / args.steps)
    h2d1, d2h1 = caller.transfer_bytes
    st = caller.shard_stats()
    clocks = sampler.stop() if sampler else None
    same = recs2.tobytes() == recs.tobytes()
    h2d = sum_over_ranks((h2d1 - h2d0) // args.steps)
    d2h = sum_over_ranks((d2h1 - d2h0) // args.steps)
    xbytes = sum_over_ranks((st["exchange_bytes"] - x0) // args.steps)
    n_pos = st["n_positive"]
    R_all = sum_over_ranks(R)

    peak, peak_src = _peaks()
    B_words = 16 + 8 * (qhi - qlo + 1)
    n_pad = world * ((S + world - 1) // world)
    hist_bytes = (2 * R + 5 * S + 4 * B_words * n_pad) * passes      # per rank and step: code + qual in, offsets + ref, dense records out
    roof = {"bound": "hbm", "kernel": "shard_hist_kernel", "kernel_ms": phases["shard_hist"],
            "achieved": hist_bytes / (phases["shard_hist"] * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "peak_source": peak_src,
            "algorithmic_bytes_per_launch": int(hist_bytes), "traffic": None,
            "kernel_share_of_step": phases["shard_hist"] / max(ms_step, 1e-9)}
    roof["frac"] = roof["achieved"] / roof["peak"]
    if rank != 0:
        caller.close()
        if own_pg:
            dist.destroy_process_group()
        return None
    comm_nranks, nvls = nccl_ranks_seen()
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        v, s_cpu, dt, kind = cpu_reference_baseline(tile, max(1, args.ref_sites_per_thread * 1000 // N), cores)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": "first %d sites x %d samples of the same window, %.1f s (%s)" % (
                   s_cpu, N, dt, "unmodified reference binary, basetype --rerun -t %d" % cores if kind == "reference" else "oracle C restatement, per-read")}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "%s: synthetic %d samples x %d sites at %.1fx depth, sample-sharded over %d GPU(s)%s" % (
                       args.workload, N, S * passes, cfg["depth"], world, "" if passes == 1 else ": %d passes over a resident %d-site tile" % (passes, S)),
                   "sites": S * passes, "tile_sites": S, "passes_per_step": passes, "n_samples": N, "entries_all_ranks": int(R_all), "qual_window": [qlo, qhi],
                   "parallelism": "sample-sharded x%d: reduce-scatter of per-site histograms + all-gather + all-reduce (NCCL)" % world,
                   "l2": "inputs (%.0f MB per rank) exceed the 126 MB L2; no flush needed" % (tile.nbytes() / 1e6) if tile.nbytes() > 126e6
                         else "inputs (%.0f MB per rank) fit the L2: use --sites to grow the tile" % (tile.nbytes() / 1e6)},
        "e2e": {"value": units / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_s * 1e3, "wire": "packed8" if packed else "plain", "matches_resident": bool(same)},
        "gpu_launches": int(launches),
        "roofline": roof,
        "phases_ms": {k: round(v, 4) for k, v in phases.items()},
        "exchange_bytes_per_step": int(xbytes), "wall_ms_per_step": wall_ms,
        "exchange_gbs_per_rank": (xbytes / max(world, 1)) / max((phases.get("reduce_scatter", 0) + phases.get("all_gather", 0) + phases.get("all_reduce", 0)) * 1e-3, 1e-12) / 1e9,
        "comm_nranks": comm_nranks, "nccl_nvls": nvls,
        "cpu_baseline": cpu, "clocks": clocks, "sites_lrt_positive": int(n_pos),
    }
    caller.close()
    if return_line:
        return {k: line[k] for k in ("value", "unit", "ms_per_step", "scaling", "config", "phases_ms", "exchange_bytes_per_step", "comm_nranks", "nccl_nvls")}
    print(json.dumps(line))
    if own_pg:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C2", choices=["C2", "C3", "C4", "C5"])
    ap.add_argument("--sites", type=int, default=0, help="sites per GPU (default: the configuration's full window, C2 = 1e6)")
    ap.add_argument("--subtiles", type=int, default=4, help="sub-tiles per step on the host-buffer (e2e) path (measured: 2-6 within 5 %, 8 +14 %, 16 +38 %)")
    ap.add_argument("--result", default="split", choices=["split", "full"],
                    help="device->host form on the e2e path: 72-byte core per site + full records of the LRT-positive sites, or 176 bytes per site")
    ap.add_argument("--wire", default="coded", choices=["packed8", "plain", "coded"],
                    help="host-buffer form of (code, qual) on the e2e path: one byte per entry + qual dictionary, or the two arrays")
    ap.add_argument("--ref-sites-per-thread", type=int, default=6000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-drain", action="store_true", help="e2e leg: wait for every ticket at the end of each step (default: the tile stream runs on across steps)")
    ap.add_argument("--qual-window", type=int, nargs=2, default=[1, 63], help="exchange window of the sample-sharded mode (C4)")
    ap.add_argument("--window-sites", type=int, default=0,
                    help="the WHOLE job in sites (e.g. 10000000 for C3's 10 Mb): region-sharded over the ranks (strong scaling), every rank "
                         "running ceil(window / ranks / --sites) passes over its resident --sites tile (the generator's sites are i.i.d.)")
    ap.add_argument("--no-bind", action="store_true", help="do not bind the rank to its GPU's local cores")
    ap.add_argument("--no-also", action="store_true", help="N > 1: skip the short sample-sharded (NCCL) run appended as `also`")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "C4":
        return run_sharded(args)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    import torch
    import basevarc_b200 as B
    from basevarc_b200.caller import PinnedArray
    torch.cuda.set_device(local)
    binding = None if args.no_bind else bind_rank_to_cores(local, world)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    cfg, tile = _workload(args.workload, args.sites, rank)
    S, N, R = tile.n_sites, tile.n_samples, tile.n_entries
    caller = B.BaseTypeCaller(N, device=local, max_sites=S, max_entries=R)
    # --window-sites: the named job (e.g. C3's 10 Mb) region-sharded over the ranks -- every rank runs `passes` passes over its
    # resident tile (sites are independent and the generator's sites i.i.d., so a pass over the same tile is the work of the next
    # S sites of the window); total work is fixed: strong scaling.  Default: one pass = one window per GPU (weak scaling).
    passes = 1 if not args.window_sites else max(1, -(-args.window_sites // (world * S)))
    units_per_gpu = S * N * passes

    # ---- kernel-only leg: inputs resident in HBM, device events on the launching stream -----------
    # the tile is resident in the form it travels in: one byte per entry + qual dictionary (--wire packed8, what the host
    # packer emits for <= 30 distinct quals; K1 reads the byte directly), or the two (code, qual) arrays (--wire plain)
    resident_packed = args.wire in ("packed8", "coded") and tile.pack8()
    caller.resident_upload(tile)
    for _ in range(args.warmup):
        caller.resident_run(1, flush_l2=False)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.wait_started()
    barrier()
    l0 = caller.kernel_launches
    ms, _ = caller.resident_run(args.steps * passes, flush_l2=False)
    ms *= passes                                       # a step = all the passes of this rank
    barrier()
    launches = caller.kernel_launches - l0
    ms_step = max_over_ranks(ms)
    recs = caller.resident_download()
    value = world * units_per_gpu / (ms_step * 1e-3)
    # per-kernel breakdown of the same pass (separate, untimed-for-value run: events sit between the launches)
    _, per_kernel = caller.resident_run(max(2, min(args.steps, 5)), flush_l2=False, per_kernel=True)
    work = caller.work_counters()

    # ---- e2e leg: host buffers through bvc_enqueue / bvc_wait (pinned H2D + D2H inside the region) --
    nsub = max(1, min(args.subtiles, S))
    bounds = np.linspace(0, S, nsub + 1).astype(int)
    subs, pins, coded = [], [], []
    for i in range(nsub):
        t = tile.slice_sites(int(bounds[i]), int(bounds[i + 1]))
        # the compact wire form (one byte per entry + qual dictionary; what the host packer emits for a tile with <= 30
        # distinct quals) unless --wire plain: the e2e path is PCIe-bound, this halves the bytes of the two big arrays
        packed = args.wire in ("packed8", "coded") and t.pack8()
        wire = "packed8" if packed else "plain"
        for k in ("ref_base", "entry_off", "mapq", "rpr") + (("packed8",) if packed else ("code", "qual")):
            src = getattr(t, k)
            pa = PinnedArray(src.shape, src.dtype)
            pa.array[...] = src
            pins.append(pa)
            setattr(t, k, pa.array)
        subs.append(t)
        # --wire coded: the packed8 bytes as 4 / 5 / 6-bit codes relative to the site's reference base + escapes (include/bvc_gpu.h,
        # "coded wire form"), decoded back into the resident packed8 bytes by a kernel inside the enqueue; split result only
        cod = None
        if packed and args.wire == "coded" and args.result == "split":
            cod = B.CodedArrays(t, alloc=lambda shape, dtype: PinnedArray(shape, dtype))
            if not cod.bits:
                cod = None
        coded.append(cod)
        if cod is not None:
            wire = "coded%d" % cod.bits
    # result form: "split" = the 72-byte core of every site + the full 176-byte record of the LRT-positive sites only
    # (bvc_enqueue_split; the half that stays behind is zero for every other site), "full" = 176 bytes for every site
    split = args.result == "split"
    if split:
        outs = [PinnedArray((t.n_sites,), B.SITE_CORE_DTYPE) for t in subs]
        # capacity = every site (C5's high-polymorphism windows call > 1/8 of their sites); only the filled part is copied
        vcfs = [PinnedArray((t.n_sites,), B.SITE_RESULT_DTYPE) for t in subs]
        n_vcf = [0] * nsub
    else:
        outs = [PinnedArray((t.n_sites,), B.SITE_RESULT_DTYPE) for t in subs]

    # the dispatcher's pattern (csrc/host/rerun.cc): two tickets in flight, the older one retired before a third is issued.  A
    # window is a STREAM of tiles, so by default the stream runs on across the step boundary (the result of step s's last tile
    # is read while step s+1's first tile uploads) and is drained once, before the closing barrier; --e2e-drain empties it at
    # the end of every step instead (one upload and one compute + read-back per step are then not overlapped).
    fly = []

    def retire():
        i, tk = fly.pop(0)
        if split:
            n_vcf[i] = caller.wait_split(tk)
        else:
            caller.wait(tk)

    def e2e_step(drain):
        for i, (t, o) in enumerate(zip(subs, outs)):
            if len(fly) == 2:
                retire()
            if coded[i] is not None:
                fly.append((i, caller.enqueue_split_coded(t, coded[i], o.array, vcfs[i].array)))
            else:
                fly.append((i, caller.enqueue_split(t, o.array, vcfs[i].array) if split else caller.enqueue(t, o.array)))
        while drain and fly:
            retire()

    for _ in range(args.warmup):
        e2e_step(True)
    barrier()
    h2d0, d2h0 = caller.transfer_bytes
    t0 = time.perf_counter()
    for _ in range(args.steps * passes):
        e2e_step(args.e2e_drain)
    while fly:
        retire()
    barrier()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / args.steps)
    h2d1, d2h1 = caller.transfer_bytes
    # copied bytes (counted by the library from the tensors it copies) + the pinned mapq/rpr words the
    # rank-sum kernel reads in place over PCIe at LRT-positive sites (2 arrays, rounded up to 32-bit words)
    zc = int((recs["n_entries"][recs["lrt_ok"] == 1].astype(np.int64) + 6).sum()) * 2 * passes
    h2d = (h2d1 - h2d0) // args.steps + zc
    d2h = (d2h1 - d2h0) // args.steps
    # ---- copy-only probe of the same link: the bytes of one pass, cudaMemcpyAsync only, every rank at the same time --------
    probe = copy_probe(local, int(h2d // passes), int(d2h // passes), barrier)
    probe["rank"] = rank
    probe["e2e_h2d_gbs_achieved"] = h2d / e2e_s / 1e9
    if binding:
        probe["binding"] = binding
    probes = [probe]
    if dist is not None:
        probes = [None] * world
        dist.all_gather_object(probes, probe)
    clocks = sampler.stop() if sampler else None
    e2e_value = world * units_per_gpu / e2e_s
    got = np.concatenate([o.array for o in outs])
    same = all(np.array_equal(got[f], recs[f]) for f in ("depth", "lrt_ok", "cvg_strand"))
    if split:
        alt = np.concatenate([v.array[o.array["vcf_index"][o.array["lrt_ok"] == 1]]["alt_bases"] for o, v in zip(outs, vcfs)])
        same = same and np.array_equal(alt, recs["alt_bases"][recs["lrt_ok"] == 1]) and sum(n_vcf) == int((recs["lrt_ok"] == 1).sum())
    else:
        same = same and np.array_equal(got["alt_bases"], recs["alt_bases"])

    # ---- rooflines (DESIGN.md "Measurement"): one entry per kernel of the step, each against ITS bound with ITS algorithmic
    # work; `roofline` = the entry of the dominant kernel; `step` = the whole step against its floors ---------------------
    peak, peak_src = _peaks()
    fp64_peak = B.measure_fp64_peak(local)
    alg_bytes = _alg_bytes(tile)                       # whole step, SURVEY 8(d): 4R + 8S + 192 S_out
    n_active = int((recs["em_fits"] > 0).sum())        # sites that ran at least one EM fit (the ladder's sites)
    n_pos = int(recs["lrt_ok"].sum())
    r_pos = int(recs["n_entries"][recs["lrt_ok"] == 1].astype(np.int64).sum())
    entry_bytes = 1 if resident_packed else 2          # K1 reads one byte per entry in the packed8 form, code + qual otherwise
    kernel_work = {
        # K1: the entries once + offsets / ref (5 B per site) in, the 64-byte site summary out (the class pool, ~4 B per class, not counted)
        "site_hist_kernel": ("hbm", entry_bytes * R + 5 * S + 64 * S),
        # K2: summary in; a 176-byte record for the sites finalised here, a 96-byte ladder + 4-byte queue item for the others
        "site_prepare_kernel": ("hbm", 64 * S + 176 * (S - n_active) + 100 * n_active),
        "queue_sort_kernels": ("hbm", 3 * 4 * int(work["em_fits"]) + 2 * 64 * int(work["em_fits"])),   # items read twice + written, their site summaries
        "em_fit_kernel": ("fp64", FLOP_PER_CLASS_PASS * int(work["em_class_passes"])),
        "em_long_kernel": ("fp64", FLOP_PER_LONG_CLASS_PASS * int(work["em_long_class_passes"])),
        "fit_loglik_kernel": ("fp64", FLOP_PER_LOGLIK_EVAL * int(work["loglik_class_evals"])),
        # K4: per active site the ladder state + its fits in, the 176-byte record out
        "ladder_kernel": ("hbm", (96 + 64 + 96) * n_active + 176 * n_active),
        # K5: the four entry arrays of the LRT-positive sites in, 64 bytes of record out
        "vcf_stats_kernel": ("hbm", (entry_bytes + 2) * r_pos + (64 + 64) * n_pos),
    }
    rooflines = {}
    for k, k_ms in per_kernel.items():
        if k not in kernel_work or k_ms <= 0:
            continue
        bound, amount = kernel_work[k]
        if bound == "fp64":
            ach, pk, unit = amount / (k_ms * 1e-3) / 1e12, fp64_peak, "TFLOP/s"
        else:
            ach, pk, unit = amount / (k_ms * 1e-3) / 1e9, peak, "GB/s"
        rooflines[k] = {"bound": bound, "kernel_ms": round(k_ms, 4), "algorithmic_" + ("flop" if bound == "fp64" else "bytes") + "_per_step": int(amount),
                        "achieved": ach, "peak": pk, "unit": unit, "frac": ach / pk}
    dom = max(per_kernel, key=per_kernel.get)
    k_ms = per_kernel[dom]
    roof = dict(rooflines.get(dom, {"bound": "hbm", "achieved": 0.0, "peak": peak, "unit": "GB/s", "frac": 0.0}))
    roof["peak_source"] = ("measured here (bvc_measure_fp64_peak: dependent-free DFMA chains on all SMs; MEASURED_PEAKS.json has no FP64 entry)"
                           if roof["bound"] == "fp64" else peak_src)
    if dom == "em_fit_kernel":
        roof.update({"class_passes_per_step": int(work["em_class_passes"]), "flop_per_class_pass": FLOP_PER_CLASS_PASS})
    hist_bytes = kernel_work["site_hist_kernel"][1]
    hist_ms = per_kernel.get("site_hist_kernel", 0.0)
    em_flop = (FLOP_PER_CLASS_PASS * int(work["em_class_passes"]) + FLOP_PER_LONG_CLASS_PASS * int(work["em_long_class_passes"])
               + FLOP_PER_LOGLIK_EVAL * int(work["loglik_class_evals"]))
    # DRAM bytes per launch of the dominant kernel and its FP64-pipe utilisation: from this round's committed ncu --set full
    # capture of the same workload (profiles/r2_traffic.json names the .csv it was read from and the HEAD it was taken at)
    traffic, ncu_note = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
            tj = json.load(f)
        ent = tj.get("%s:%d" % (args.workload, S), {}).get(dom)
        if ent:
            traffic, ncu_note = ent.get("dram_bytes_per_launch"), {k: v for k, v in ent.items() if k != "dram_bytes_per_launch"}
            ncu_note["source"] = tj.get("source")
    except OSError:
        pass
    roof.update({"traffic": traffic, "ncu": ncu_note, "kernel": dom, "kernel_ms": k_ms,
                 "kernel_share_of_step": k_ms / max(sum(per_kernel.values()), 1e-9)})
    n_reads = recs["depth"].sum(1).astype(np.float64)
    f_alg = float((n_reads * (21.0 * recs["em_passes"] - 3.0 * recs["em_fits"])).sum())   # per-READ formulation, SURVEY 8(d)

    # ---- N > 1: a short run of the SAMPLE-SHARDED path (the one with a collective) so that the scaling record carries NCCL
    # evidence; its own line is what `--workload C4` prints
    also = None
    if world > 1 and not args.no_also:
        import copy
        a2 = copy.copy(args)
        a2.workload, a2.sites, a2.steps, a2.warmup, a2.no_cpu_baseline, a2.window_sites = "C4", 16384, 2, 1, True, 0
        caller.close()
        try:
            also = run_sharded(a2, return_line=True)
        except Exception as e:                      # the C2 line must survive a failing side run
            also = {"error": repr(e)[:300]}
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    comm_nranks, nvls = nccl_ranks_seen()
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        dropin = {}
        v, s_cpu, dt, kind = cpu_reference_baseline(tile, max(1, args.ref_sites_per_thread * 1000 // N), cores, dropin=dropin)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": "first %d sites x %d samples of the same window, %.1f s (%s)" % (
                   s_cpu, N, dt, "unmodified reference binary, basetype --rerun -t %d" % cores if kind == "reference"
                   else "oracle C restatement, per-read")}
        if dropin:
            # the whole command line as a drop-in: same batch files, same -t; host inflate / tokenise / format included
            cpu["rerun_cli"] = dropin
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if args.window_sites else "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": ("%s: synthetic %d samples x %d sites at %.1fx depth per GPU" % (args.workload, N, S, cfg["depth"])) if not args.window_sites else
                               ("%s: synthetic %d samples x %d sites at %.1fx depth, region-sharded over %d GPU(s): %d passes per rank over a resident %d-site tile" % (
                                   args.workload, N, passes * world * S, cfg["depth"], world, passes, S)),
                   "sites_per_gpu": S * passes, "tile_sites": S, "passes_per_step": passes, "n_samples": N, "entries_per_gpu": R * passes,
                   "parallelism": "window-sharded x%d, no collective" % world,
                   "l2": "inputs (%.0f MB per launch) exceed the 126 MB L2; no flush needed" % (tile.nbytes() / 1e6)},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": e2e_s * 1e3, "subtiles": nsub, "stream": "drained every step" if args.e2e_drain else "tiles stream across step boundaries, drained once before the closing barrier", "wire": wire, "result": args.result, "matches_resident": bool(same),
                "h2d_gbs_per_gpu": h2d / e2e_s / 1e9},
        "copy_probe": {"what": "cudaMemcpyAsync only, pinned host memory, the bytes of one pass, all ranks at once; GB/s per rank",
                       "ranks": probes,
                       "aggregate_h2d_gbs": float(sum(p["h2d_gbs"] for p in probes)),
                       "aggregate_e2e_h2d_gbs": float(sum(p["e2e_h2d_gbs_achieved"] for p in probes))},
        "comm_nranks": comm_nranks, "nccl_nvls": nvls, "also": also,
        "gpu_launches": int(launches),
        "roofline": roof,
        "rooflines": rooflines,
        "hbm_stream": {"kernel": "site_hist_kernel (%s form)" % ("one-byte packed8" if resident_packed else "two-array"), "kernel_ms": hist_ms,
                       "algorithmic_bytes_per_launch": int(hist_bytes),
                       "achieved": hist_bytes / max(hist_ms, 1e-9) / 1e6, "peak": peak, "unit": "GB/s",
                       "frac": hist_bytes / max(hist_ms, 1e-9) / 1e6 / peak},
        # (per PASS over the resident tile: kernels_ms, rooflines and the work counters are those of one pass; a step is `passes` of them)
        "step": {"algorithmic_bytes": int(alg_bytes), "hbm_floor_ms": alg_bytes / peak / 1e6,
                 "frac_of_hbm_floor": (alg_bytes / peak / 1e6) / (ms / passes), "fp64_floor_ms": em_flop / fp64_peak / 1e9,
                 "frac_of_fp64_floor": (em_flop / fp64_peak / 1e9) / (ms / passes), "fp64_peak_tflops": fp64_peak,
                 "em_flop_floor_ms": em_flop / fp64_peak / 1e9, "measured_ms": ms / passes, "passes_per_step": passes,
                 "per_read_flop_survey": f_alg, "per_read_tflops_equiv": f_alg / (ms / passes * 1e-3) / 1e12,
                 "note": "per_read_* counts the flops of the reference's per-read EM (SURVEY 8d); the class-collapsed "
                         "kernels do ~n_reads/n_classes fewer, so it is an equivalence figure, not pipe utilisation"},
        "work": {k: int(v) for k, v in work.items()},
        "kernels_ms": {k: round(v, 4) for k, v in per_kernel.items()},
        "cpu_baseline": cpu,
        "clocks": clocks,
        "sites_lrt_positive": int(recs["lrt_ok"].sum()), "em_passes_per_site": float(recs["em_passes"].mean()),
    }
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
