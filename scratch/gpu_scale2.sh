run() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 10 --warmup 3 --no-cpu --no-extras --no-comparator "$@" 2>gpurun_out/scale_err_$n.txt | tail -1; }
run 2 --scaling weak > gpurun_out/r2_scale_weak_n2.json
run 2 --scaling strong > gpurun_out/r2_scale_strong_n2.json
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload c5 --steps 5 --no-cpu 2>>gpurun_out/scale_err_2.txt | tail -1 > gpurun_out/r2_c5_n2.json
tail -3 gpurun_out/scale_err_2.txt
for f in gpurun_out/r2_scale_weak_n2.json gpurun_out/r2_scale_strong_n2.json gpurun_out/r2_c5_n2.json; do python -c "
import json,sys; d=json.load(open('$f')); print('$f', d['n_gpus'], d['scaling'], 'ms', round(d['ms_per_step'],4), 'value %.4g'%d['value'], 'e2e ms', round(d['e2e'].get('ms_per_step',0),3), d['config'].get('numa_binding'), d['e2e'].get('h2d_ceiling_gbs_per_rank'))"; done
