# usage (GPU box): bash tools/bench_smoke.sh N  -> small bench lines through every code path of bench.py at N GPUs
N=${1:-1}
run() { if [ "$N" = 1 ]; then python bench.py "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N "$@"; fi; }
mkdir -p gpurun_out
run --sites 100000 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/smoke_c2_n$N.json 2> gpurun_out/smoke_c2_n$N.err || tail -5 gpurun_out/smoke_c2_n$N.err
run --workload C3 --sites 30000 --window-sites 240000 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/smoke_c3_n$N.json 2> gpurun_out/smoke_c3_n$N.err || tail -5 gpurun_out/smoke_c3_n$N.err
run --workload C4 --sites 8192 --window-sites 32768 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/smoke_c4_n$N.json 2> gpurun_out/smoke_c4_n$N.err || tail -5 gpurun_out/smoke_c4_n$N.err
python - <<PY
import json
for w in ("c2","c3","c4"):
    try:
        d=json.loads(open("gpurun_out/smoke_%s_n$N.json"%w).read().strip().splitlines()[-1])
        print(w, d["n_gpus"], d["scaling"], "%.3e"%d["value"], round(d["ms_per_step"],3), d["config"]["workload"][:90], "e2e %.3e"%d["e2e"]["value"], "nranks", d.get("comm_nranks"))
        if "copy_probe" in d: print("   probe", [(round(p["h2d_gbs"],1), round(p["both_h2d_gbs"],1), round(p["e2e_h2d_gbs_achieved"],1), (p.get("binding") or {}).get("bound_to")) for p in d["copy_probe"]["ranks"]])
        if d.get("also"): print("   also", {k: d["also"].get(k) for k in ("value","ms_per_step","comm_nranks","error")})
    except Exception as e:
        print(w, "FAILED", repr(e)[:200])
PY
