mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extras --no-comparator"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_c3.csv $CMD > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:ctc_pair_kernel -s 5 -c 2 -f -o gpurun_out/prof_pair_r2 $CMD > gpurun_out/ncu_run.log 2>&1; tail -1 gpurun_out/ncu_run.log
ncu -i gpurun_out/prof_pair_r2.ncu-rep --page raw --csv > gpurun_out/prof_pair_r2.raw.csv 2>/dev/null
python scratch/ncu_dyn.py gpurun_out/prof_pair_r2.ncu-rep > gpurun_out/prof_pair_r2.dyn.txt 2>&1
rm -f gpurun_out/prof_pair_r2.ncu-rep
python bench.py > gpurun_out/r2_bench_c3.json 2> gpurun_out/r2y_c3.err; tail -c 200 gpurun_out/r2y_c3.err
python bench.py --workload c2 > gpurun_out/r2_bench_c2.json 2>/dev/null
python -c "
import json; d=json.load(open('gpurun_out/r2_bench_c3.json')); print(d['ms_per_step'], d['roofline']['frac'], d['roofline']['traffic'], d['extras']['c2']['ms_per_call'])"
