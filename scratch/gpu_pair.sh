#!/bin/bash
timeout 300 python scratch/dbg_pair.py 2>&1 | grep -v "^   " | tail -4
for w in 0 8; do
  MDD_CTC_PAIR_WPS=$w timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --algo 2 2>/dev/null | python scratch/show.py "pair regbuf=1 wps=$w"
done
