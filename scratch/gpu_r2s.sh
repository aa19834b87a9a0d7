CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-extras --no-comparator"
for m in 1 0; do
MDD_CTC_EROW=$m ncu --metrics gpu__time_duration.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed --clock-control none -k regex:ctc_pair_kernel -s 5 -c 2 --csv --log-file gpurun_out/erow_$m.csv $CMD > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/erow_$m.csv")) if len(r)>5 and r[0].isdigit()]
for r in rows: print("EROW=$m", r[4][40:75], r[-3], r[-2], r[-1])
PY
done
