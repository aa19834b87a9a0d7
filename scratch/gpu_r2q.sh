mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extras --no-comparator"
ncu --set full --clock-control none --import-source on -k regex:ctc_pair_kernel -s 5 -c 2 -f -o gpurun_out/prof_pair_r2 $CMD > gpurun_out/ncu_run.log 2>&1; tail -2 gpurun_out/ncu_run.log
ncu --set full --clock-control none --import-source on -k 'regex:ctc_dual|ctc_pair_kernel' -s 3 -c 2 -f -o gpurun_out/prof_dual_c2 python scratch/one_shape.py 200 32 42 30 4 > gpurun_out/ncu_run2.log 2>&1; tail -2 gpurun_out/ncu_run2.log
for r in prof_pair_r2 prof_dual_c2; do
  ncu -i gpurun_out/$r.ncu-rep --page raw --csv > gpurun_out/$r.raw.csv 2>/dev/null
  python scratch/ncu_dyn.py gpurun_out/$r.ncu-rep > gpurun_out/$r.dyn.txt 2>&1
  rm -f gpurun_out/$r.ncu-rep
done
grep -c . gpurun_out/prof_pair_r2.raw.csv
