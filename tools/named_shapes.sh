# usage (GPU box): bash tools/named_shapes.sh N [full]  -> the bench lines of the shapes BASELINE.json names, at N GPUs of the box
N=${1:-1}
run() { if [ "$N" = 1 ]; then python bench.py "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@"; fi; }
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo_n$N.txt 2>&1
run --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_c2_n$N.json 2> gpurun_out/r2_bench_c2_n$N.err || tail -5 gpurun_out/r2_bench_c2_n$N.err
run --workload C3 --sites 150000 --window-sites 10000000 --steps 3 --warmup 3 --no-cpu-baseline --no-also > gpurun_out/r2_bench_c3_n$N.json 2> gpurun_out/r2_bench_c3_n$N.err || tail -5 gpurun_out/r2_bench_c3_n$N.err
run --workload C4 --sites $((32768 * N)) --window-sites 1000000 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_c4_n$N.json 2> gpurun_out/r2_bench_c4_n$N.err || tail -5 gpurun_out/r2_bench_c4_n$N.err
if [ "$N" = 1 ]; then
  run --workload C5 --sites 200000 --window-sites 1000000 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_c5_1m.json 2> gpurun_out/r2_bench_c5_1m.err || tail -5 gpurun_out/r2_bench_c5_1m.err
fi
if [ "$2" = full ]; then
  # the reference binary and the drop-in on the FULL 1,000,000-site C2 window (not the 96k-site sample)
  python bench.py --steps 3 --warmup 3 --ref-sites-per-thread $((1000000 / $(nproc) + 1)) > gpurun_out/r2_bench_c2_reffull.json 2> gpurun_out/r2_bench_c2_reffull.err || tail -5 gpurun_out/r2_bench_c2_reffull.err
fi
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/r2_bench_*n$N.json") + glob.glob("gpurun_out/r2_bench_c5_1m.json") + glob.glob("gpurun_out/r2_bench_c2_reffull.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, d["n_gpus"], d["scaling"], "%.3e"%d["value"], "ms %.3f"%d["ms_per_step"], "e2e %.3e"%d["e2e"]["value"], "nranks", d.get("comm_nranks"), d.get("clocks"))
        if d.get("cpu_baseline"): print("   cpu", json.dumps(d["cpu_baseline"])[:1500])
    except Exception as e:
        print(f, "FAILED", repr(e)[:200])
PY
