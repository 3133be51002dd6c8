# A/B of the codebook-gradient exchange in bench.py at N GPUs: peer kernel on a side stream, peer kernel
# inside the step graph, NCCL async (the default).  Usage: N=2 bash tools/ab_exchange.sh
N=${N:-2}
for v in "RQ_BENCH_PEER=1" "RQ_BENCH_PEER=inline" "RQ_X=0"; do
  env $v timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus $N --steps 30 --warmup 5 2>gpurun_out/ab_ex.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('N=%d us/step %.1f  G items/s %.2f  e2e M/s %.0f  %s' % (d['n_gpus'], d['ms_per_step']*1e3, d['value']/1e9, d['e2e']['value']/1e6, d['extra']['grad_exchange']))" || tail -5 gpurun_out/ab_ex.err
done
