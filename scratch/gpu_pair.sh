#!/bin/bash
timeout 300 python scratch/dbg_pair.py 2>&1 | grep -v "^   " | tail -14
MDD_CTC_PAIR_REGBUF=1 timeout 300 python scratch/dbg_pair.py 2>&1 | grep -v "^   " | tail -4
timeout 900 python -m pytest tests/test_gpu_parity.py -q -k ctc 2>&1 | tail -5
MDD_CTC_PAIR_REGBUF=1 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --algo 2 2>/dev/null | python scratch/show.py "pair regbuf=1"
for mr in 255 168 144 128; do for w in 0 8; do
  MDD_CTC_PAIR_MAXR=$mr MDD_CTC_PAIR_WPS=$w timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --algo 2 2>/dev/null | python scratch/show.py "pair maxr=$mr wps=$w"
done; done
