mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2j_pytest.txt
tail -6 gpurun_out/r2j_pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py --workload c4 --no-cpu > gpurun_out/r2_bench_c4.json 2> gpurun_out/r2j_c4.err; tail -c 400 gpurun_out/r2j_c4.err
python -c "
import json; d=json.load(open('gpurun_out/r2_bench_c4.json')); print(d['value'], d['ms_per_step'], d['e2e'], d.get('gpu_launches'))"
