# usage: bash tools/sweep2.sh TAG WORKLOAD SITES "ENV=.." ...
tag=$1; wl=$2; sites=$3; shift 3
i=0
for envs in "$@"; do
  i=$((i+1))
  env $envs python bench.py --workload $wl --sites $sites --no-cpu-baseline --steps 3 --warmup 3 > gpurun_out/${tag}_$i.json 2> gpurun_out/${tag}_$i.err
  python - <<PY
import json
d=json.load(open('gpurun_out/${tag}_$i.json'))
print("$wl $envs", "step %.3f ms" % d['ms_per_step'], {k: round(v,3) for k,v in d['kernels_ms'].items() if k.startswith(('em_','fit_'))})
PY
done
