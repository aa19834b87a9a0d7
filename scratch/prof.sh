#!/bin/bash
# usage: scratch/prof.sh NAME  -> gpurun_out/NAME.ncu-rep (full set, source on) of the CTC pair kernel at config 3
/usr/local/graft/bin/gpurun --timeout 900 -- "python bench.py --steps 2 --warmup 1 --no-cpu --algo 2 > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:ctc_pair -c 1 -s 2 -o gpurun_out/$1 -f python bench.py --steps 2 --warmup 1 --no-cpu --algo 2 > gpurun_out/$1.log 2>&1; tail -2 gpurun_out/$1.log" 2>&1 | tail -4
ncu -i gpurun_out/$1.ncu-rep --page source --csv --print-source cuda,sass > /tmp/$1.csv 2>/dev/null
python scratch/ncu_lines.py /tmp/$1.csv 30
