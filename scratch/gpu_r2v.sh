for m in 1 0 1 0; do
MDD_PDL=$m python bench.py --workload c3peaky --no-cpu --no-extras --no-comparator 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('PDL=$m c3peaky', round(d['ms_per_step'],4), d['config']['tiers'])"
done
python bench.py --no-cpu --no-extras --no-comparator 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c3', round(d['ms_per_step'],4), d['roofline']['frac'], d['roofline']['traffic'])"
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
