n=$1
mkdir -p gpurun_out
for sc in strong weak; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 20 --warmup 5 --no-cpu --no-extras --no-comparator --scaling $sc 2>gpurun_out/scale_err_$n.txt | tail -1 > gpurun_out/r2_scale_${sc}_n$n.json
python -c "
import json; d=json.load(open('gpurun_out/r2_scale_${sc}_n$n.json')); print(d['n_gpus'], d['scaling'], 'ms', round(d['ms_per_step'],4), 'value %.4g'%d['value'], 'e2e ms', round(d['e2e'].get('ms_per_step',0),3), d['e2e'].get('frac_of_h2d_ceiling'))"
done
