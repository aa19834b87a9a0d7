set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r2a_pytest.txt
python scratch/err_where.py > gpurun_out/r2a_err_where.txt 2>&1
tail -5 gpurun_out/r2a_pytest.txt
cat gpurun_out/r2a_err_where.txt
