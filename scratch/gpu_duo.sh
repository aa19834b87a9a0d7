#!/bin/bash
export MDD_CTC_DUO=1
timeout 300 python scratch/dbg_pair.py 2>&1 | grep -v "^   " | tail -12
timeout 900 python -m pytest tests/test_gpu_parity.py -q -k ctc -x 2>&1 | tail -5
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --algo 2 2>/dev/null | python scratch/show.py "duo"
MDD_CTC_DUO=0 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --algo 2 2>/dev/null | python scratch/show.py "pair"
