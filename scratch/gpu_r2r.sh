python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --steps 20 --warmup 5 --no-cpu --no-extras --no-comparator 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('EROW on ', d['ms_per_step'], d['roofline']['frac'], d['config']['tiers'])"
MDD_CTC_EROW=0 python bench.py --steps 20 --warmup 5 --no-cpu --no-extras --no-comparator 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('EROW off', d['ms_per_step'], d['roofline']['frac'])"
python scratch/time_team.py 2>&1 | tail -3
