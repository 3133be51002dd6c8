run() { label=$1; shift; env "$@" timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 4 --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('$label', 'value %.3f G'%(d['value']/1e9), 'ms/step %.4f'%d['ms_per_step'], 'e2e %.1f M'%(d['e2e']['value']/1e6))"; }


run "eager-comm reserve8" X=1
run "eager-comm reserve0" RQ_SM_RESERVE=0
