timeout 600 python -m pytest tests -m gpu -x -q -k "ctc" 2>&1 | tail -15 > gpurun_out/r2d_pytest.txt
tail -6 gpurun_out/r2d_pytest.txt
timeout 120 python scratch/time_team.py > gpurun_out/r2d_team.txt 2>&1
MDD_CTC_TEAM_MAX_PAIRS=0 timeout 120 python scratch/time_team.py >> gpurun_out/r2d_team.txt 2>&1
MDD_CTC_TEAM_MAX_PAIRS=100000 timeout 120 python scratch/time_team.py >> gpurun_out/r2d_team.txt 2>&1
cat gpurun_out/r2d_team.txt
