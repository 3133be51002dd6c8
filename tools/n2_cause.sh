# where does the N = 2 step lose its 5 us?  (profiling only)
p() { python -c "
import json,sys;d=json.loads(open('$1').read().strip().splitlines()[-1]);print('$2', round(d['ms_per_step']*1e3,2),'us/step fwd', round(d['extra']['fwd_ms']*1e3,2),'bwd', round(d['extra']['bwd_ms']*1e3,2))"; }
python bench.py --steps 20 --warmup 5 --no-extras > gpurun_out/c_a.json 2>/dev/null; p gpurun_out/c_a.json "N=1 block graph:"
RQ_BENCH_BLOCK_GRAPH=0 python bench.py --steps 20 --warmup 5 --no-extras > gpurun_out/c_b.json 2>/dev/null; p gpurun_out/c_b.json "N=1 one graph per step:"
CUDA_VISIBLE_DEVICES=0 python bench.py --steps 20 --warmup 5 --no-extras > gpurun_out/c_c0.json 2>/dev/null &
CUDA_VISIBLE_DEVICES=1 python bench.py --steps 20 --warmup 5 --no-extras > gpurun_out/c_c1.json 2>/dev/null &
wait
p gpurun_out/c_c0.json "two independent N=1 runs at once, GPU 0:"; p gpurun_out/c_c1.json "two independent N=1 runs at once, GPU 1:"
CUDA_VISIBLE_DEVICES=0 RQ_BENCH_BLOCK_GRAPH=0 python bench.py --steps 20 --warmup 5 --no-extras > gpurun_out/c_d0.json 2>/dev/null &
CUDA_VISIBLE_DEVICES=1 RQ_BENCH_BLOCK_GRAPH=0 python bench.py --steps 20 --warmup 5 --no-extras > gpurun_out/c_d1.json 2>/dev/null &
wait
p gpurun_out/c_d0.json "two independent per-step-graph runs at once, GPU 0:"; p gpurun_out/c_d1.json "... GPU 1:"
RQ_BENCH_WHATIF_NO_EXCHANGE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-extras > gpurun_out/c_e.json 2>/dev/null; p gpurun_out/c_e.json "N=2 torchrun, no exchange:"
