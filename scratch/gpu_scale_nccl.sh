n=8
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 30 --warmup 5 --no-cpu --no-extras --no-comparator --scaling strong 2>gpurun_out/scale_err_$n.txt | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$1', d['n_gpus'], d['scaling'], 'ms', round(d['ms_per_step'],4), 'value %.4g'%d['value'], 'kernel_ms', round(d['roofline']['kernel_ms'],4))"; }
run default
NCCL_MAX_NCHANNELS=1 NCCL_MIN_NCHANNELS=1 run nchannels1
NCCL_MAX_NCHANNELS=1 NCCL_MIN_NCHANNELS=1 NCCL_PROTO=LL run nch1_LL
