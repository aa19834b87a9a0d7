timeout 900 python -m pytest tests -m gpu -x -q -k "ctc" 2>&1 | tail -3
python bench.py --workload c3peaky --no-extras --no-cpu --no-comparator --steps 10 > gpurun_out/r2_bench_c3peaky.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/r2_bench_c3peaky.json')); print('c3peaky', d['ms_per_step'], d['config']['tiers'])"
