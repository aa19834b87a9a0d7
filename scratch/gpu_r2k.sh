mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2k_pytest.txt
tail -6 gpurun_out/r2k_pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
python bench.py > gpurun_out/r2k_bench.json 2> gpurun_out/r2k_bench.err; tail -c 600 gpurun_out/r2k_bench.err
cat gpurun_out/r2k_bench.json | head -c 3000
