CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extras --no-comparator"
$CMD > gpurun_out/ncu_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:ctc_pair_kernel -s 4 -c 2 -f -o gpurun_out/prof_split_r2 $CMD > gpurun_out/ncu_run.log 2>&1
tail -2 gpurun_out/ncu_run.log
