python -m pytest tests -m gpu -x -q -k "ctc" 2>&1 | tail -15 > gpurun_out/r2c_pytest.txt
tail -6 gpurun_out/r2c_pytest.txt
python bench.py --steps 10 --warmup 3 --no-extras --no-cpu --no-comparator > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; tail -c 600 gpurun_out/r2c_bench.err; python -c "
import json; d=json.load(open('gpurun_out/r2c_bench.json')); print(d['ms_per_step'], d['roofline']['frac'], d['e2e']['ms_per_step'], d['config']['tiers'])"
