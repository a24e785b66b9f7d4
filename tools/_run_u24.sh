mkdir -p gpurun_out/r02s
python -m pytest tests/test_gpu_route.py -x -q -m gpu -k "three_bytes or pipelined or empty" 2>&1 | tail -5
python bench.py --no-cpu-baseline --no-configs --steps 50 > gpurun_out/r02s/bench.json 2> gpurun_out/r02s/bench.err || tail -5 gpurun_out/r02s/bench.err
python -c "
import json
d=json.loads(open('gpurun_out/r02s/bench.json').read().strip().splitlines()[-1])
print('headline', d['value']/1e9, d['ms_per_step'], 'e2e', d['e2e']['value']/1e6, d['e2e']['ms_per_step'], 'u24', d['e2e_ids_u24']['value']/1e6, d['e2e_ids_u24']['ms_per_step'])
"
