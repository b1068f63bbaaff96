# usage (on the GPU box): bash tools/sweep.sh TAG "ENV1=a ENV2=b" "ENV1=c" ...   -> one short bench line per setting
tag=$1; shift
i=0
for envs in "$@"; do
  i=$((i+1))
  env $envs python bench.py --no-cpu-baseline --steps 4 --warmup 3 > gpurun_out/${tag}_$i.json 2> gpurun_out/${tag}_$i.err
  python - <<PY
import json
d=json.load(open('gpurun_out/${tag}_$i.json'))
print("$envs", "step %.3f ms  e2e %.2f ms" % (d['ms_per_step'], d['e2e']['ms_per_step']), {k: round(v,3) for k,v in d['kernels_ms'].items()}, d['work'])
PY
done
