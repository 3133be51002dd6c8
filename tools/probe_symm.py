"""Probe of torch.distributed._symmetric_memory on this box (run under torchrun, N >= 2)."""
import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
t = symm.empty(8 * 24576, dtype=torch.float32, device=torch.device("cuda", local))
h = symm.rendezvous(t, dist.group.WORLD)
t.fill_(rank + 1)
h.barrier()
peer = h.get_buffer((rank + 1) % world, (4,), torch.float32)
print(rank, "ptrs", [hex(p) for p in h.buffer_ptrs], "sig", [hex(p) for p in h.signal_pad_ptrs], "pad bytes", h.signal_pad_size,
      "mc", h.has_multicast_support, hex(h.multicast_ptr) if h.has_multicast_support else None, "peer val", peer.tolist(),
      "dev arrays", hex(h.buffer_ptrs_dev), hex(h.signal_pad_ptrs_dev), flush=True)
pad = h.get_signal_pad(rank, (8,), torch.uint32)
print(rank, "pad", pad.tolist(), flush=True)
h.barrier()
dist.barrier()
torch.cuda.synchronize()
os._exit(0)
