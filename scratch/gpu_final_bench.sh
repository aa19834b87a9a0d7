python bench.py > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err; tail -c 300 gpurun_out/r2_bench_final.err
python bench.py --workload c2 --no-extras > gpurun_out/r2_bench_c2.json 2>/dev/null
python bench.py --workload c3peaky --no-extras --steps 10 > gpurun_out/r2_bench_c3peaky.json 2>/dev/null
python bench.py --workload c4 > gpurun_out/r2_bench_c4.json 2>/dev/null
python bench.py --workload c5 > gpurun_out/r2_bench_c5.json 2>/dev/null
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference.json 2>/dev/null
for f in final c2 c3peaky c4 c5 reference; do python -c "
import json; d=json.load(open('gpurun_out/r2_bench_$f.json')); print('$f', d['metric'], '%.4g'%d['value'], d.get('ms_per_step'), (d.get('roofline') or {}).get('frac'), (d.get('e2e') or {}).get('value'), (d.get('config') or {}).get('tiers'))"; done
