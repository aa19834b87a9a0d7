#!/bin/bash
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -15
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu --extras > gpurun_out/bench_extras.json 2> gpurun_out/bench_extras.err; tail -3 gpurun_out/bench_extras.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_extras.json').read().strip().splitlines()[-1])
print("ms", d["ms_per_step"], "frac", d["roofline"]["frac"])
print(json.dumps(d.get("extras"), indent=1))
PY
