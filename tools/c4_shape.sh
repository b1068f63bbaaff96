# usage (GPU box): bash tools/c4_shape.sh N TAG -> the C4 job (100,000 samples x 1 Mb, sample-sharded) at N GPUs into gpurun_out/TAG_bench_c4_nN.json
N=${1:-1}; TAG=${2:-r2b}
run() { if [ "$N" = 1 ]; then python bench.py "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@"; fi; }
mkdir -p gpurun_out
f=gpurun_out/${TAG}_bench_c4_n$N
run --workload C4 --sites $((32768 * N)) --window-sites 1000000 --steps 3 --warmup 3 --no-cpu-baseline > $f.json 2> $f.err || tail -5 $f.err
python - <<PY
import json
d=json.loads(open("$f.json").read().strip().splitlines()[-1])
print(d["n_gpus"], "%.3e"%d["value"], "ms %.3f"%d["ms_per_step"], "e2e ms %.3f"%d["e2e"]["ms_per_step"], "nranks", d.get("comm_nranks"), "nvls", d.get("nccl_nvls"), d["phases_ms"], d["exchange_bytes_per_step"], d["config"]["workload"])
PY
