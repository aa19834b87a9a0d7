timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for m in 0 100000; do MDD_CTC_TEAM_MAX_PAIRS=$m timeout 120 python scratch/time_team.py 2>&1; done
