"""Multi-GPU check of parallel.PeerAllReduce (run under torchrun, N >= 2): results against NCCL, in place and out of
place, eager and replayed from a CUDA graph together with the fused backward; then a latency comparison."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import rq_vae_recommender_b200 as R
from rq_vae_recommender_b200 import _lib

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ok = True
for shape, slots in (((3, 256, 32), 4), ((4, 1024, 64), 2), ((8,), 1)):
    peer = R.parallel.PeerAllReduce(shape, slots=slots)
    for it in range(6):
        s = it % slots
        g = torch.Generator(device=dev).manual_seed(1000 * it + rank)
        x = torch.randn(shape, device=dev, generator=g)
        peer.buffer[s].copy_(x)
        want = x.clone()
        dist.all_reduce(want)
        if it % 2 == 0:
            got = peer.all_reduce_(s, average=False)
        else:
            got = peer.all_reduce_(s, average=False, out=torch.empty(shape, device=dev))
            assert torch.equal(peer.buffer[s], x)
        err = float((got - want).abs().max() / want.abs().max().clamp_min(1e-30))
        # bit-identical on every rank
        ref = got.clone()
        dist.broadcast(ref, 0)
        same = bool(torch.equal(ref, got))
        if err > 1e-6 or not same:
            ok = False
            print(f"rank {rank} shape {shape} it {it}: rel err {err:.2e} identical_across_ranks={same}", flush=True)
# graph replay: producer kernel + all-reduce captured together
shape = (3, 256, 32)
peer = R.parallel.PeerAllReduce(shape, slots=2)
src = [torch.randn(shape, device=dev, generator=torch.Generator(device=dev).manual_seed(77 + rank + 10 * s)) for s in range(2)]
graphs = []
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    for s in range(2):
        peer.buffer[s].copy_(src[s]); peer.all_reduce_(s)
torch.cuda.current_stream().wait_stream(side)
torch.cuda.synchronize(); dist.barrier()
for s in range(2):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        peer.buffer[s].copy_(src[s])
        peer.all_reduce_(s)
    graphs.append(g)
for it in range(10):
    graphs[it % 2].replay()
torch.cuda.synchronize()
for s in range(2):
    want = src[s].clone(); dist.all_reduce(want); want /= world
    err = float((peer.buffer[s] - want).abs().max() / want.abs().max())
    if err > 1e-6:
        ok = False
        print(f"rank {rank} graph slot {s}: rel err {err:.2e}", flush=True)
# latency: 100 graph replays vs 100 NCCL all-reduces of the same buffer
def timed(fn, n=200):
    torch.cuda.synchronize(); dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(n):
        fn(i)
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
t_peer = timed(lambda i: graphs[i % 2].replay())
buf = src[0].clone()
t_nccl = timed(lambda i: dist.all_reduce(buf))
flag = _lib.lib().rq_debug_peer_timeout_flag()
print(f"rank {rank}: ok={ok} timeout_flag={flag} peer (copy + all-reduce, graph replay) {t_peer:.1f} us  nccl eager {t_nccl:.1f} us", flush=True)
dist.barrier(); torch.cuda.synchronize()
os._exit(0 if ok and not flag else 1)
