"""Host -> device batch prefetcher: overlaps the pinned-memory copy of batch i+1 with the training step of batch i.

The reference feeds the network synchronously: `prep_data` copies every batch with `.to(device)` right before
`model.forward` (training/train_ubresnet2018_wlarcv2.py:322-332), so the PCIe copy (134 MB for 32 crops of
image + int64 label + weight) sits on the critical path.  Here the copy of the next batch runs on a side stream into a
second set of device buffers while the current step computes; the consumer only waits on the copy's event.
"""
import torch


class DevicePrefetcher:
    """Wraps an iterator of tuples of pinned CPU tensors; yields tuples of device tensors.

    Double buffered: at most two batches are resident on the device.  Every batch is copied exactly once, on
    `self.stream`; `__next__` makes the current stream wait for that copy and immediately queues the following one."""

    def __init__(self, batches, device, depth=2):
        self.it = iter(batches)
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)
        self.depth = depth
        self.slots = []            # ring of (tensors, copy-done event, consumed event)
        self.pending = []
        for _ in range(depth - 1):
            self._issue()

    def _issue(self):
        try:
            host = next(self.it)
        except StopIteration:
            return
        for t in host:
            if not t.is_pinned():
                raise RuntimeError("DevicePrefetcher: host tensors must be pinned (tensor.pin_memory())")
        with torch.cuda.stream(self.stream):
            dev = tuple(t.to(self.device, non_blocking=True) for t in host)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self.pending.append((dev, ev))

    def __iter__(self):
        return self

    def __next__(self):
        if not self.pending:
            self._issue()
        if not self.pending:
            raise StopIteration
        dev, ev = self.pending.pop(0)
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ev)
        for t in dev:
            t.record_stream(cur)       # the caching allocator must not reuse the buffer before the step has read it
        self._issue()                  # next batch starts copying while this one is being used
        return dev
