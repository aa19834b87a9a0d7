#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_parity.py -q -k ctc 2>&1 | tail -25
for rb in 1 0; do for w in 1 4; do
  MDD_CTC_PAIR_REGBUF=$rb MDD_CTC_PAIR_WPC=$w timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --algo 2 2>/dev/null | python scratch/show.py "pair regbuf=$rb wpc=$w"
done; done
