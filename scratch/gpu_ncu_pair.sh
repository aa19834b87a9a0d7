CMD="python bench.py --algo 2 --steps 2 --warmup 3 --no-cpu --no-extras --no-comparator"
$CMD > gpurun_out/ncu_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:ctc_pair -s 1 -c 1 -f -o gpurun_out/prof_pair_r2 $CMD > gpurun_out/ncu_run.log 2>&1
tail -3 gpurun_out/ncu_run.log
