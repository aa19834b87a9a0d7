python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python scratch/time_mid.py 2>&1 | tail -8
bash scratch/gpu_r2y.sh
python scratch/time_small.py > gpurun_out/r2_small_batch_times.txt 2>&1
python scratch/time_mid.py >> gpurun_out/r2_small_batch_times.txt 2>&1
