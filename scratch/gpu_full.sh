#!/bin/bash
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -6
timeout 300 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_pair.json 2> gpurun_out/bench_pair.err; python scratch/show.py "algo0 default" < gpurun_out/bench_pair.json
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --workload c2 2>/dev/null | python scratch/show.py "c2 algo0"
timeout 300 python bench.py --steps 20 --warmup 3 --no-cpu --workload c2 --algo 3 2>/dev/null | python scratch/show.py "c2 algo3"
