mkdir -p gpurun_out
for s in "200 32 42 30" "1000 512 42 120"; do
  n=$(echo $s | tr ' ' '_')
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2l_launches_$n.csv python scratch/one_shape.py $s 3 > /dev/null 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/r2l_launches_$n.csv")) if len(r)>5 and r[0].isdigit()]
hdr=None
for r in rows[-14:]:
    print(r[4][:70], r[-1])
PY
done
