timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3; timeout 200 python bench.py > gpurun_out/bench_x.json 2>gpurun_out/bench_x.err; python -c "
import json;d=json.load(open('gpurun_out/bench_x.json'));print(d['value'],d['ms_per_step'],d['e2e']['value']);[print(k,round(v['ms_per_step'],4)) for k,v in d['kernels'].items()]"
