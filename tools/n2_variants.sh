# N = 2 timing variants of the bench step (profiling only): prints us/step, forward-only, backward-only
run() { # name, nproc, env...
  name=$1; np=$2; shift; shift
  if [ "$np" = 1 ]; then
    env "$@" timeout 250 python bench.py --gpus 1 --steps 20 --warmup 5 --no-extras > gpurun_out/n2_$name.json 2> gpurun_out/n2_$name.err
  else
    env "$@" timeout 250 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $np --steps 20 --warmup 5 --no-extras > gpurun_out/n2_$name.json 2> gpurun_out/n2_$name.err
  fi
  python -c "
import json;d=json.loads(open('gpurun_out/n2_$name.json').read().strip().splitlines()[-1]);print('$name', round(d['ms_per_step']*1e3,2),'us/step', round(d['extra']['fwd_ms']*1e3,2), round(d['extra']['bwd_ms']*1e3,2))"
}
run n1 1 A=1
run n1_res8 1 RQ_SM_RESERVE=8
run n2_base 2 A=1
run n2_noexchange 2 RQ_BENCH_WHATIF_NO_EXCHANGE=1
run n2_noexchange_res0 2 RQ_BENCH_WHATIF_NO_EXCHANGE=1 RQ_SM_RESERVE=0
