#!/bin/bash
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_pair.json 2> gpurun_out/bench_pair.err; python scratch/show.py "default" < gpurun_out/bench_pair.json
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu --extras > gpurun_out/bench_extras.json 2> gpurun_out/bench_extras.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_extras.json').read().strip().splitlines()[-1]); print(json.dumps(d.get('extras'), indent=1))"
python bench.py --steps 2 --warmup 1 --no-cpu > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_pair.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ctc_pair -c 1 -s 2 -o gpurun_out/prof_pair_final -f python bench.py --steps 2 --warmup 1 --no-cpu --algo 2 > gpurun_out/prof_pair_final.log 2>&1; tail -1 gpurun_out/prof_pair_final.log
python scratch/run_lse.py && ncu --set full --clock-control none --import-source on -k regex:row_lse -c 1 -s 1 -o gpurun_out/prof_row_lse -f python scratch/run_lse.py > gpurun_out/prof_row_lse.log 2>&1; tail -1 gpurun_out/prof_row_lse.log
ncu --set full --clock-control none --import-source on -k regex:ctc_pair -c 1 -s 2 -o gpurun_out/prof_pair_logits -f python scratch/run_lse.py > gpurun_out/prof_pair_logits.log 2>&1; tail -1 gpurun_out/prof_pair_logits.log
