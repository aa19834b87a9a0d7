mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extras --no-comparator"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_c3.csv $CMD > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ctc_pair_kernel -s 5 -c 2 -f -o gpurun_out/prof_pair_r2 $CMD > gpurun_out/ncu_run.log 2>&1; tail -1 gpurun_out/ncu_run.log
ncu --set full --clock-control none --import-source on -k 'regex:ctc_dual|ctc_pair_kernel' -s 3 -c 2 -f -o gpurun_out/prof_dual_c2 python scratch/one_shape.py 200 32 42 30 4 > gpurun_out/ncu_run2.log 2>&1
ncu --set full --clock-control none --import-source on -k 'regex:ctc_dual|ctc_pair_kernel' -s 6 -c 2 -f -o gpurun_out/prof_dual_b512 python scratch/one_shape.py 1000 512 42 120 4 > gpurun_out/ncu_run3.log 2>&1
for r in prof_pair_r2 prof_dual_c2 prof_dual_b512; do
  ncu -i gpurun_out/$r.ncu-rep --page raw --csv > gpurun_out/$r.raw.csv 2>/dev/null
  python scratch/ncu_dyn.py gpurun_out/$r.ncu-rep > gpurun_out/$r.dyn.txt 2>&1
  rm -f gpurun_out/$r.ncu-rep
done
python bench.py > gpurun_out/r2_bench_c3.json 2> gpurun_out/r2x_c3.err; tail -c 200 gpurun_out/r2x_c3.err
python bench.py --workload c2 > gpurun_out/r2_bench_c2.json 2>/dev/null
python scratch/time_small.py > gpurun_out/r2_small_batch_times.txt 2>&1; tail -15 gpurun_out/r2_small_batch_times.txt
