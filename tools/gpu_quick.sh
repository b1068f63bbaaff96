# usage (on the GPU box): bash tools/gpu_quick.sh TAG  -> GPU tests, a C2 bench line (no CPU baseline), a 200k-site line
tag=$1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
python -c "
import json; d=json.load(open('gpurun_out/${tag}_bench.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['kernel'], d['roofline']['frac'], d['kernels_ms'], d['work'])"
tail -5 gpurun_out/${tag}_bench.err
