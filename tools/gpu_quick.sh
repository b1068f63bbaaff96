tag=$1
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --no-cpu-baseline > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
python -c "
import json; d=json.load(open('gpurun_out/${tag}_bench.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['kernel'], d['roofline']['frac'], d['kernels_ms'], d['work'])"
tail -5 gpurun_out/${tag}_bench.err
python bench.py --no-cpu-baseline --sites 200000 --steps 3 > gpurun_out/${tag}_bench200k.json 2>&1
python -c "
import json; d=json.load(open('gpurun_out/${tag}_bench200k.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['kernel'], d['roofline']['frac'], d['kernels_ms'], d['work'])"
