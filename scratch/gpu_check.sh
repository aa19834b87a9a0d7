#!/bin/bash
python scratch/dbg_flags.py 2>&1 | tail -8
for kf in 8 16; do
  MDD_CTC_FAST_KF=$kf python bench.py --steps 10 --warmup 3 --no-cpu --algo 2 2>/dev/null | python scratch/show.py "algo2 kf=$kf"
done
python bench.py --steps 10 --warmup 3 --no-cpu 2>/dev/null | python scratch/show.py "algo0"
timeout 500 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
