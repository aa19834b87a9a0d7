mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/zz_alone.txt
tail -4 gpurun_out/zz_alone.txt
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/zz_full.txt
tail -4 gpurun_out/zz_full.txt
