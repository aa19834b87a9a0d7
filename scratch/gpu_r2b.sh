set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2b_pytest.txt
tail -4 gpurun_out/r2b_pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2b_smoke.txt 2>&1; tail -3 gpurun_out/r2b_smoke.txt
python bench.py --steps 10 --warmup 3 > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err; tail -c 600 gpurun_out/r2b_bench.err; head -c 3000 gpurun_out/r2b_bench.json
