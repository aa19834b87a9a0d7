#!/bin/bash
timeout 300 python scratch/dbg_pair.py 2>&1 | grep -v "^   " | tail -14
timeout 900 python -m pytest tests/test_gpu_parity.py -q -k ctc 2>&1 | tail -5
for w in 0 8; do
  MDD_CTC_PAIR_WPS=$w timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --algo 2 2>/dev/null | python scratch/show.py "pair regbuf=1 wps=$w"
done
