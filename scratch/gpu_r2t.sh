python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python scratch/time_team.py 2>&1 | tail -7
echo "--- MDD_PDL=0"
MDD_PDL=0 python scratch/time_team.py 2>&1 | tail -7
