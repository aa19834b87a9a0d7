mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
MDD_SO_PATH=scratch/var/libmddctc_timing.so python scratch/dual_phases.py 2>&1 | grep -E "T=|setup|loop done"
python scratch/time_team.py 2>&1 | head -4
