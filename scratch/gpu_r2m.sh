mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r2m_pytest.txt
tail -25 gpurun_out/r2m_pytest.txt
python scratch/time_team.py 2>&1 | tail
MDD_CTC_DUAL_MAX_PAIRS=0 python scratch/time_team.py 2>&1 | head -4
