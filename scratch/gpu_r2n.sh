python scratch/dbg_dual.py 2>&1 | tail -20
for s in "200 32 42 30" "1000 512 42 120"; do
  n=$(echo $s | tr ' ' '_')
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2n_launches_$n.csv python scratch/one_shape.py $s 3 > /dev/null 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/r2n_launches_$n.csv")) if len(r)>5 and r[0].isdigit()]
for r in rows[-6:]:
    print(r[4][:90], r[-1])
PY
done
