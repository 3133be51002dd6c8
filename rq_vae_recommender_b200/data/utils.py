"""Host -> device feeding of the quantisation path.

`cycle`, `batch_to`, `next_batch` keep the reference's names and meaning (data/utils.py:4-17).
`DevicePrefetcher` is the B200 form of `next_batch`: the reference copies each batch synchronously
on the compute stream right before the step (train_rqvae.py:143 -> data/utils.py:15-17); here the
copy of batch i+1 runs on its own stream from pinned memory while batch i is being quantised, so a
step costs max(copy, compute) instead of their sum."""
from typing import Iterable, Iterator

import torch


def cycle(dataloader):
    while True:
        for data in dataloader:
            yield data


def _to(v, device, non_blocking=False):
    return v.to(device, non_blocking=non_blocking) if torch.is_tensor(v) else v


def batch_to(batch, device, non_blocking: bool = False):
    """Tensor, NamedTuple (SeqBatch / TokenizedSeqBatch), tuple/list or dict of tensors -> `device`."""
    if torch.is_tensor(batch):
        return _to(batch, device, non_blocking)
    if hasattr(batch, "_asdict"):
        return type(batch)(*[_to(v, device, non_blocking) for v in batch._asdict().values()])
    if isinstance(batch, dict):
        return {k: _to(v, device, non_blocking) for k, v in batch.items()}
    if isinstance(batch, (tuple, list)):
        return type(batch)(_to(v, device, non_blocking) for v in batch)
    raise TypeError(f"batch_to: unsupported batch type {type(batch).__name__}")


def next_batch(dataloader, device):
    return batch_to(next(dataloader), device)


def _record_stream(batch, stream) -> None:
    vals = [batch] if torch.is_tensor(batch) else (
        batch._asdict().values() if hasattr(batch, "_asdict") else
        batch.values() if isinstance(batch, dict) else batch)
    for v in vals:
        if torch.is_tensor(v) and v.is_cuda:
            v.record_stream(stream)


class DevicePrefetcher:
    """Iterator over device-resident batches, one copy ahead of the consumer.

    `source` yields host batches (pinned memory gives asynchronous copies; pageable memory still
    works but the copy then blocks the host).  Each `next()` (1) makes the consumer's current
    stream wait for the copy of the batch it returns, (2) starts the copy of the following batch on
    the prefetcher's stream, and returns.  Buffers are handed to the caching allocator with
    `record_stream`, so a batch's memory is not reused while the consumer stream still reads it.

    `into` (tensor batches only): a ring of at least depth + 1 preallocated device tensors the batches are
    copied into, in order — the static inputs of CUDA-graph-captured steps, which must not move.  Before a
    buffer is overwritten the copy stream waits for everything the consumer stream had been given when the
    copy was issued, which includes the step that last read that buffer."""

    def __init__(self, source: Iterable, device, depth: int = 1, into=None):
        if depth < 1:
            raise ValueError("depth must be >= 1")
        if into is not None and len(into) < depth + 1:
            raise ValueError("into needs at least depth + 1 buffers")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("DevicePrefetcher feeds the CUDA engine; got device " + str(device))
        self._it: Iterator = iter(source)
        self._stream = torch.cuda.Stream(device=self.device)
        self._queue = []
        self._free_events = []
        self._depth = depth
        self._into = list(into) if into is not None else None
        self._n_issued = 0
        for _ in range(depth):
            self._issue()

    def _issue(self) -> None:
        try:
            host = next(self._it)
        except StopIteration:
            return
        # events are recycled (creating one costs more than recording it)
        ev = self._free_events.pop() if self._free_events else torch.cuda.Event()
        if self._into is not None:
            dev = self._into[self._n_issued % len(self._into)]
            self._n_issued += 1
            # the consumer's earlier reads of this buffer come first
            self._stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self._stream), torch.no_grad():
                dev.copy_(host, non_blocking=True)
                ev.record(self._stream)
        else:
            with torch.cuda.stream(self._stream):
                dev = batch_to(host, self.device, non_blocking=True)
                ev.record(self._stream)
        self._queue.append((dev, ev))

    def __iter__(self):
        return self

    def __next__(self):
        if not self._queue:
            raise StopIteration
        dev, ev = self._queue.pop(0)
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ev)
        self._free_events.append(ev)
        if self._into is None:
            _record_stream(dev, cur)
        self._issue()
        return dev
